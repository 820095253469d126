#!/usr/bin/env python
"""relMSE at equal time, uniform vs guided, through elaina_b200/elaina-exec and its frames_time/ dumps (N4 harness).

  python tools/equal_time.py --workload c1 --spp 1024 --ref-spp 16384 --every 64 --out gpurun_out/equal_time

1. reference field: the analytic solution where the workload has one (c3 family), else a uniform run at --ref-spp (its own Monte-Carlo
   error, ~ spp / ref_spp of the curves' last point, is stated in the output);
2. uniform and guided runs at --spp with saveTimeMetricsDuration = --every -> frames_time/<elapsed ms>.exr;
3. tools/relmse_vs_time.py turns both directories into relMSE(t) and compares them at the time the faster one finished.
Prints one JSON object."""
import argparse
import json
import os
import shutil
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
from elaina_b200 import workloads  # noqa: E402
import relmse_vs_time as rvt  # noqa: E402
import b1_run  # noqa: E402

BIN = os.path.join(ROOT, "elaina_b200", "elaina-exec")


def run(conf, cwd):
    p = subprocess.run([BIN, conf], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=3000, cwd=cwd)
    if p.returncode != 0:
        raise SystemExit("elaina-exec failed: " + p.stdout[-1500:])


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="c1")
    ap.add_argument("--spp", type=int, default=0)
    ap.add_argument("--ref-spp", type=int, default=16384)
    ap.add_argument("--every", type=int, default=64)
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "equal_time"))
    a = ap.parse_args()
    wl = workloads.workload(a.workload)
    spp = a.spp or wl["spp"]
    W, H = wl["frame"]
    case = os.path.abspath(os.path.join(a.out, a.workload))
    shutil.rmtree(case, ignore_errors=True)
    os.makedirs(case)
    mask = wl.get("mask")
    out = {"workload": wl["desc"], "spp": spp, "dump_every_samples": a.every}
    ref = b1_run.analytic(wl)
    if ref is None:
        conf = workloads.write_case(wl, case, integrator="uniform", exp_name="reference", spp=a.ref_spp, exports=("raw",))
        run(conf, case)
        f = np.fromfile(os.path.join(case, "reference", "solution.f32"), np.float32).reshape(H, W, 5)
        ref = (f[..., 0] / np.maximum(f[..., 4], 1e-30)).astype(np.float64)
        res = json.load(open(os.path.join(case, "reference", "result.json")))
        out["reference"] = {"kind": "uniform run at %d spp" % a.ref_spp, "duration_ms": res.get("duration"),
                            "note": "its own Monte-Carlo error is ~ %d/%d of the last point of the uniform curve" % (spp, a.ref_spp)}
    else:
        out["reference"] = {"kind": "analytic solution"}
    curves = {}
    for integ in ("uniform", "guided"):
        conf = workloads.write_case(wl, case, integrator=integ, exp_name=integ, spp=spp, extra_setting={"saveTimeMetricsDuration": a.every}, exports=("raw",))
        run(conf, case)
        res = json.load(open(os.path.join(case, integ, "result.json")))
        f = np.fromfile(os.path.join(case, integ, "solution.f32"), np.float32).reshape(H, W, 5)
        sol = f[..., 0] / np.maximum(f[..., 4], 1e-30)
        curves[integ] = rvt.curve(os.path.join(case, integ), ref, mask)
        out[integ] = {"duration_ms": res.get("duration"), "walk_steps_per_sec": res.get("walk_steps_per_sec"), "final_relmse": rvt.relmse(sol, ref, mask),
                      "curve": curves[integ]}
    t = min(out["uniform"]["duration_ms"], out["guided"]["duration_ms"])
    out["equal_time"] = {"ms": t, "uniform": rvt.at_time(curves["uniform"], t), "guided": rvt.at_time(curves["guided"], t),
                         "note": "durations include the progressive dumps (D2H + EXR/PNG write per dump) of both runs alike"}
    print(json.dumps(out), flush=True)
    shutil.rmtree(case, ignore_errors=True)


if __name__ == "__main__":
    main()
