"""Bin an ncu source page by active lanes per instruction: where do the warp instructions of a divergent kernel go?
usage: ncu -i rep.ncu-rep --page source --csv --print-source sass,cuda | python tools/ncu_bins.py"""
import collections
import csv
import sys

rows = list(csv.reader(sys.stdin))
fpath, hdr, out = "", None, []
for r in rows:
    if len(r) == 2 and r[0] in ("File Path", "File Name"):
        fpath = r[1].split("/")[-1]
    elif len(r) > 5 and r[0] == "Line No":
        hdr = r
    elif hdr and len(r) == len(hdr) and r[0].isdigit():
        d = dict(zip(hdr, r))
        try:
            out.append((fpath, int(r[0]), int(d["Instructions Executed"]), int(d["Thread Instructions Executed"])))
        except Exception:
            pass
tot = sum(o[2] for o in out) or 1
names = ["32 lanes", "16-31 lanes", "8-16 lanes", "4-8 lanes", "<4 lanes"]
bins = {k: [0, 0] for k in names}
perfile = collections.Counter()
lowfile = collections.Counter()
for f, l, wi, ti in out:
    if wi == 0:
        continue
    lanes = ti / wi
    k = names[0] if lanes > 31.5 else names[1] if lanes >= 16 else names[2] if lanes >= 8 else names[3] if lanes >= 4 else names[4]
    bins[k][0] += wi
    bins[k][1] += ti
    perfile[f] += wi
    if lanes < 8:
        lowfile[f] += wi
for k in names:
    wi, ti = bins[k]
    print("%-12s %5.1f%% of warp instructions, avg lanes %.1f" % (k, 100 * wi / tot, ti / max(wi, 1)))
print("by file:", {f: round(100 * v / tot, 1) for f, v in perfile.most_common(10)})
print("below 8 lanes, by file:", {f: round(100 * v / tot, 1) for f, v in lowfile.most_common(10)})
