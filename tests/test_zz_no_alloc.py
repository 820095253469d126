"""GPU: wost_lbvh_nearest_static, the closest-point query for embedders that may not allocate device memory.

Some hosts of libwost.so cannot let it call cudaMalloc once they are running: an application that has already handed the
whole card to its own memory pool, a frame loop that cannot tolerate the device-wide stall an allocation causes, a sandbox
that forbids allocation after initialisation. For them the library stages the query points and the results through one
fixed array inside its own device image (include/wost.h states the limit and the sharing rules).

What is checked here, through the C ABI with host buffers: the static-scratch call returns exactly what the allocating
wost_lbvh_nearest returns (closest primitive equal, distance bit-identical -- the tolerance of every LBVH parity test in
this repo), on two calls in a row, the second with fewer points and on another scene, so that left-overs of the first call
in the shared array (points, results, control words) would show up as a mismatch. Over-long batches must be refused.

The file name sorts last on purpose: `pytest -x` stops at the first failure, and a failure here must not hide the parity
tests. It passes alone in a fresh process (first call = per-device set-up) and after the rest of the suite.
"""
import ctypes
import re
import os

import numpy as np
import pytest

from helpers import bits, random_queries

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _max_points():
    src = open(os.path.join(ROOT, "include", "wost.h")).read()
    return int(re.search(r"#define\s+WOST_NEAREST_STATIC_MAX_POINTS\s+(\d+)", src).group(1))


def _nearest_both(lib, scene, q):
    """(idx, dist) from the allocating entry point and from the static-scratch one, both straight through ctypes."""
    q = np.ascontiguousarray(q, np.float32)
    n = len(q)
    out = []
    for fn in (lib.wost_lbvh_nearest, lib.wost_lbvh_nearest_static):
        idx = np.full(n, 0xDEADBEEF, np.uint32)
        dist = np.full(n, np.nan, np.float32)
        rc = fn(scene.h, q.ctypes.data_as(ctypes.c_void_p), n, idx.ctypes.data_as(ctypes.c_void_p), dist.ctypes.data_as(ctypes.c_void_p))
        assert rc == 0, lib.wost_last_error()
        out.append((idx, dist))
    return out


def test_static_scratch_nearest_matches_allocating_entry_point_twice():
    from elaina_b200 import _lib, api, scenes
    lib = _lib.load()
    cap = _max_points()
    # call 1: a 3D mesh, the largest batch the scratch takes
    v3, p3 = scenes.icosphere(4, 1.0, 0.2)
    s3 = api.Scene(3, v3, p3)
    q3 = random_queries(s3.verts, cap, 11)[0]
    (i_ref, d_ref), (i_st, d_st) = _nearest_both(lib, s3, q3)
    assert np.array_equal(i_ref, i_st)
    assert np.array_equal(bits(d_ref), bits(d_st))
    assert np.isfinite(d_st).all() and (i_st < len(p3)).all()
    # call 2, straight after: a 2D polyline, a ragged batch; the array still holds call 1's points and results
    v2, p2 = scenes.star_polygon(5000)
    s2 = api.Scene(2, v2, p2)
    q2 = random_queries(s2.verts, 12345, 12)[0]
    (i_ref, d_ref), (i_st, d_st) = _nearest_both(lib, s2, q2)
    assert np.array_equal(i_ref, i_st)
    assert np.array_equal(bits(d_ref), bits(d_st))
    # idx is optional, as in wost_lbvh_nearest
    dist = np.full(100, np.nan, np.float32)
    q = np.ascontiguousarray(q2[:100])
    assert lib.wost_lbvh_nearest_static(s2.h, q.ctypes.data_as(ctypes.c_void_p), 100, None, dist.ctypes.data_as(ctypes.c_void_p)) == 0
    assert np.array_equal(bits(dist), bits(d_ref[:100]))
    # the Python wrapper goes through the same symbol
    i_w, d_w = s2.nearest_static(q2[:1000])
    assert np.array_equal(i_w, i_ref[:1000]) and np.array_equal(bits(d_w), bits(d_ref[:1000]))


def test_static_scratch_nearest_refuses_batches_beyond_its_limit():
    from elaina_b200 import _lib, api, scenes
    lib = _lib.load()
    cap = _max_points()
    v2, p2 = scenes.star_polygon(64)
    s2 = api.Scene(2, v2, p2)
    n = cap + 1
    q = np.zeros((n, 2), np.float32)
    idx = np.zeros(n, np.uint32)
    dist = np.full(n, -7.0, np.float32)
    rc = lib.wost_lbvh_nearest_static(s2.h, q.ctypes.data_as(ctypes.c_void_p), n, idx.ctypes.data_as(ctypes.c_void_p), dist.ctypes.data_as(ctypes.c_void_p))
    assert rc == 1                                    # WOST_ERR_INVALID
    assert str(cap).encode() in lib.wost_last_error()  # the message names the limit
    assert (dist == -7.0).all()                        # nothing was written
    # an empty batch is fine and touches nothing
    assert lib.wost_lbvh_nearest_static(s2.h, q.ctypes.data_as(ctypes.c_void_p), 0, None, dist.ctypes.data_as(ctypes.c_void_p)) == 0
