"""CPU: the C-ABI library loads and exports every symbol include/wost.h declares (no compute calls)."""
import os
import re

from elaina_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "wost.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(wost_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_exported():
    lib = _lib.load()
    names = declared_symbols()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), "symbol %s declared in include/wost.h is not exported by libwost.so" % n
    bound = {s[0] for s in _lib.SYMBOLS}
    assert set(names) == bound, "ctypes table and header disagree: %s" % (set(names) ^ bound)


def test_version_and_error_string():
    lib = _lib.load()
    assert b"sm_100a" in lib.wost_version()
    assert lib.wost_last_error() is not None


def test_invalid_arguments_are_reported_not_thrown():
    import ctypes
    lib = _lib.load()
    h = ctypes.c_void_p()
    rc = lib.wost_scene_create(4, None, 0, None, 0, ctypes.byref(h))
    assert rc == 1 and b"invalid" in lib.wost_last_error()
