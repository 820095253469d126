"""Aggregate an ncu source page (--page source --csv --print-source sass,cuda) into per-source-line totals.
usage: ncu -i rep.ncu-rep --page source --csv --print-source sass,cuda | python tools/ncu_lines.py [topN]"""
import csv, sys
top = int(sys.argv[1]) if len(sys.argv) > 1 else 40
rows = list(csv.reader(sys.stdin))
out = []
fpath = ""
hdr = None
for r in rows:
    if len(r) == 2 and r[0] in ("File Path", "File Name"):
        fpath = r[1].split("/")[-1]
    elif len(r) > 5 and r[0] == "Line No":
        hdr = r
    elif hdr and len(r) == len(hdr) and r[0].isdigit():
        d = dict(zip(hdr, r))
        try:
            out.append((int(d["Instructions Executed"]), int(d["Thread Instructions Executed"]), int(d.get("# Samples", 0) or 0), fpath, r[0], r[1].strip()[:90]))
        except Exception:
            pass
tot = sum(o[0] for o in out)
tott = sum(o[1] for o in out)
print("total warp-inst %d thread-inst %d" % (tot, tott))
for o in sorted(out, reverse=True)[:top]:
    print("%6.2f%% winst  thr/inst %5.1f  samples %6d  %s:%s  %s" % (100.0 * o[0] / max(tot, 1), o[1] / max(o[0], 1), o[2], o[3], o[4], o[5]))
