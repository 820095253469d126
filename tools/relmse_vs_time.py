#!/usr/bin/env python
"""N4 (second half): relMSE-vs-time curves from the progressive dumps of a run directory.

Both binaries -- the reference's elaina-exec and elaina_b200/elaina-exec -- write `frames_time/<elapsed ms>.exr` every
`saveTimeMetricsDuration` samples (integrator/uniform/integrator.cu:577-609). This tool turns such a directory into the
curve the BASELINE metric asks for ("relMSE at equal time"): relMSE of every dump against a reference field, with the
error measure of SURVEY 8(d): mean((a - b)^2 / (b^2 + 1e-2)) over the evaluated pixels.

  python tools/relmse_vs_time.py <run_dir> --workload c3                 # reference field = the workload's analytic solution
  python tools/relmse_vs_time.py <run_dir> --reference other_run/solution.f32 --frame 1024 1024 [--mask mask.pgm]
  python tools/relmse_vs_time.py <run_a> <run_b> --workload c3           # several runs -> one table, plus the equal-time comparison

Output: one JSON line per run {"run": ..., "curve": [{"ms": t, "relmse": e}, ...]} and, for two runs, the relMSE of each at
the time the slower one finished (linear interpolation in log-log space)."""
import argparse
import json
import math
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))


def relmse(a, b, mask=None):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    e = (a - b) ** 2 / (b ** 2 + 1e-2)
    if mask is not None:
        e = e[np.asarray(mask, bool)]
    return float(np.mean(e))


def read_pgm(path):
    with open(path, "rb") as f:
        assert f.readline().strip() == b"P5"
        line = f.readline()
        while line.startswith(b"#"):
            line = f.readline()
        w, h = [int(x) for x in line.split()]
        f.readline()
        return np.frombuffer(f.read(w * h), np.uint8).reshape(h, w)


def curve(run_dir, ref, mask):
    from exr_read import read_exr
    d = os.path.join(run_dir, "frames_time")
    pts = []
    for fn in sorted(os.listdir(d), key=lambda s: int(s.split(".")[0]) if s.split(".")[0].isdigit() else -1):
        if not fn.endswith(".exr"):
            continue
        img = read_exr(os.path.join(d, fn))[::-1]  # files are written vertically flipped (core/texture.cu:92-93)
        pts.append({"ms": int(fn.split(".")[0]), "relmse": relmse(img[..., 0], ref, mask)})
    return pts


def at_time(pts, ms):
    """relMSE at `ms` by log-log interpolation (Monte-Carlo error falls like 1/t); None outside the curve"""
    pts = [p for p in pts if p["ms"] > 0 and p["relmse"] > 0]
    if not pts or ms < pts[0]["ms"]:
        return None
    for a, b in zip(pts, pts[1:]):
        if a["ms"] <= ms <= b["ms"]:
            t = (math.log(ms) - math.log(a["ms"])) / max(1e-12, math.log(b["ms"]) - math.log(a["ms"]))
            return math.exp(math.log(a["relmse"]) * (1 - t) + math.log(b["relmse"]) * t)
    last = pts[-1]
    return last["relmse"] * last["ms"] / ms  # beyond the last dump: extrapolate with the 1/t law


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("runs", nargs="+", help="run directories (<base_path>/<exp_name>) holding frames_time/")
    ap.add_argument("--workload", help="take the reference field and the mask from elaina_b200.workloads (analytic solution: c3 family)")
    ap.add_argument("--reference", help="raw fp32 film {r,g,b,a,w} x W*H (the `raw` export) used as the reference field")
    ap.add_argument("--frame", nargs=2, type=int)
    ap.add_argument("--mask")
    a = ap.parse_args()
    mask = None
    if a.workload:
        from elaina_b200 import workloads
        import b1_run
        wl = workloads.workload(a.workload)
        ref = b1_run.analytic(wl)
        if ref is None:
            raise SystemExit("workload %s has no analytic solution: pass --reference" % a.workload)
        mask = wl.get("mask")
    elif a.reference and a.frame:
        W, H = a.frame
        f = np.fromfile(a.reference, np.float32).reshape(H, W, 5)
        ref = f[..., 0] / np.maximum(f[..., 4], 1e-30)
        mask = read_pgm(a.mask) if a.mask else None
    else:
        raise SystemExit("need --workload or --reference with --frame")
    curves = {}
    for r in a.runs:
        curves[r] = curve(r, ref, mask)
        print(json.dumps({"run": r, "curve": curves[r]}), flush=True)
    if len(a.runs) == 2:
        ca, cb = curves[a.runs[0]], curves[a.runs[1]]
        if ca and cb:
            t = min(ca[-1]["ms"], cb[-1]["ms"])
            print(json.dumps({"equal_time_ms": t, a.runs[0]: at_time(ca, t), a.runs[1]: at_time(cb, t)}), flush=True)


if __name__ == "__main__":
    main()
