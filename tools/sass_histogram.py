#!/usr/bin/env python
"""Opcode evidence for the tcgen05 / TMA / vector-atomic claims: per-kernel counts of the SASS mnemonics that matter, read from
the in-tree library with cuobjdump.  Usage: python tools/sass_histogram.py > profiles/sass_histogram_r3.txt"""
import collections
import os
import re
import subprocess
import sys

LIB = sys.argv[1] if len(sys.argv) > 1 else os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "elaina_b200", "libwost.so")
WATCH = ("UTCHMMA", "UTCQMMA", "UTCBAR", "LDTM", "STTM", "UBLKCP", "UTMALDG", "UTMASTG", "SYNCS", "HMMA", "REDG", "RED", "ATOMG", "ATOMS", "LDGSTS")
sass = subprocess.run(["cuobjdump", "-sass", LIB], stdout=subprocess.PIPE, text=True, check=True).stdout
demangle = {}
per = collections.defaultdict(collections.Counter)
size = collections.Counter()
fn = None
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        fn = m.group(1)
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Za-z0-9_.]+)", line)
    if not m or fn is None:
        continue
    op = m.group(1)
    size[fn] += 1
    base = op.split(".")[0]
    if base in WATCH:
        key = base + (".F32x4" if "F32x4" in op else "")
        per[fn][key] += 1
names = subprocess.run(["c++filt"], input="\n".join(size), stdout=subprocess.PIPE, text=True).stdout.splitlines()
short = {k: re.sub(r"\(.*", "", n)[:110] for k, n in zip(size, names)}
tot = collections.Counter()
for c in per.values():
    tot.update(c)
print("# sm_100a SASS opcode counts, cuobjdump -sass elaina_b200/libwost.so (%d kernels, %d instructions)" % (len(size), sum(size.values())))
print("## totals over the library (0 = absent)")
for k in sorted(set(tot) | {"UTCHMMA", "LDTM", "STTM", "UTCBAR", "UBLKCP", "UTMALDG", "UTMASTG", "REDG.F32x4"}):
    print("%8d  %s" % (tot.get(k, 0), k))
print("## per kernel: instructions | watched opcodes")
for f in sorted(size, key=lambda f: short[f]):
    if per[f]:
        print("%-110s %7d | %s" % (short[f], size[f], ", ".join("%s %d" % kv for kv in sorted(per[f].items()))))
