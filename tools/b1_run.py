#!/usr/bin/env python
"""BASELINE B1: run the reference's own CUDA build (baseline/_ref/elaina-exec-ref, built by baseline/build_ref_cuda.sh from
the sources where they lie) and our elaina-exec on the SAME elaina-exec case directory, on the GPU of this box.

  python tools/b1_run.py --workload c3 [--integrator uniform|guided] [--spp N] [--out gpurun_out/b1] [--time-frames N]
  python tools/b1_run.py --pin            # the small oracle-pinning cases -> gpurun_out/b1/b1_<case>.npz (copied to tests/golden/)

Prints one JSON line per run:
  reference_cuda  walk-steps/s = B1_WALK_STEPS (queue-size counter patch) / result.json duration (exec.cu:157-158)
  ours            walk-steps/s through elaina_b200/elaina-exec on the same conf.json (duration = the same wall-clock quantity)
  fields          relMSE between the two films, relMSE of each against the analytic solution where one exists
TEST / BASELINE INFRASTRUCTURE: nothing under elaina_b200/ uses the reference binary.
"""
import argparse
import json
import os
import re
import shutil
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from elaina_b200 import workloads  # noqa: E402

REF_BIN = os.path.join(ROOT, "baseline", "_ref", "elaina-exec-ref")
OUR_BIN = os.path.join(ROOT, "elaina_b200", "elaina-exec")


def relmse(a, b, mask=None):
    """mean((a - b)^2 / (b^2 + 1e-2)) over the evaluated pixels"""
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    e = (a - b) ** 2 / (b ** 2 + 1e-2)
    if mask is not None:
        e = e[np.asarray(mask, bool)]
    return float(np.mean(e))


def run_reference(conf, case_dir, exp, timeout):
    raw = os.path.join(case_dir, exp, "solution_b1.f32")
    env = dict(os.environ, B1_DUMP_RAW=raw)
    t0 = time.perf_counter()
    p = subprocess.run([REF_BIN, conf], env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=timeout, cwd=case_dir)
    wall = time.perf_counter() - t0
    out = p.stdout
    m = re.findall(r"B1_WALK_STEPS (\d+)", out)
    res = {}
    try:
        res = json.load(open(os.path.join(case_dir, exp, "result.json")))
    except Exception:
        pass
    return dict(rc=p.returncode, wall_s=wall, walk_steps=int(m[-1]) if m else None, duration_ms=res.get("duration"), raw=raw if os.path.exists(raw) else None,
                log_tail=out[-1500:])


def run_ours(conf, case_dir, exp, timeout):
    t0 = time.perf_counter()
    p = subprocess.run([OUR_BIN, conf], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=timeout, cwd=case_dir)
    wall = time.perf_counter() - t0
    res = {}
    try:
        res = json.load(open(os.path.join(case_dir, exp, "result.json")))
    except Exception:
        pass
    raw = os.path.join(case_dir, exp, "solution.f32")
    return dict(rc=p.returncode, wall_s=wall, walk_steps=res.get("walk_steps"), duration_ms=res.get("duration"), device_steps_per_sec=res.get("walk_steps_per_sec"),
                raw=raw if os.path.exists(raw) else None, log_tail=p.stdout[-800:])


def load_film(path, W, H):
    f = np.fromfile(path, np.float32).reshape(H, W, 5)
    with np.errstate(divide="ignore", invalid="ignore"):
        return f[..., :3] / f[..., 4:5]


def analytic(wl):
    if wl["name"] in ("c3", "c3-small", "c5", "pin-dirichlet2d"):
        from elaina_b200 import scenes
        W, H = wl["frame"]
        g = 1.5 * (2.0 * np.arange(W, dtype=np.float64) / W - 1.0)
        X, Y = np.meshgrid(g, g)
        return scenes.harmonic_2d(np.stack([X, Y], -1), 3)
    return None


def frames_time_curve(case_dir, exp, W, H, ref_field, mask):
    """N4: relMSE(t) from the frames_time/<ms>.exr dumps of either binary -- reads the raw sibling when present, else skips."""
    d = os.path.join(case_dir, exp, "frames_time")
    pts = []
    if not os.path.isdir(d):
        return pts
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    from exr_read import read_exr
    for fn in sorted(os.listdir(d), key=lambda s: int(s.split(".")[0]) if s.split(".")[0].isdigit() else 0):
        if not fn.endswith(".exr"):
            continue
        try:
            img = read_exr(os.path.join(d, fn))[::-1]  # files are written vertically flipped
        except Exception as e:  # noqa: BLE001
            pts.append({"ms": int(fn.split(".")[0]), "error": str(e)})
            continue
        pts.append({"ms": int(fn.split(".")[0]), "relmse": relmse(img[..., 0], ref_field, mask)})
    return pts


def one(wl_name, integrator, spp, out_root, time_frames=0, timeout=1500, skip_ref=False, skip_ours=False, depth=None):
    wl = workloads.workload(wl_name)
    if depth:
        wl["depth"] = depth
    spp = spp or wl["spp"]
    W, H = wl["frame"]
    case_dir = os.path.abspath(os.path.join(out_root, "%s_%s" % (wl_name, integrator)))
    shutil.rmtree(case_dir, ignore_errors=True)
    os.makedirs(case_dir)
    extra = {"saveTimeMetricsDuration": int(time_frames)} if time_frames else None
    line = {"workload": wl_name, "integrator": integrator, "spp": spp, "frame": [W, H], "desc": wl["desc"]}
    mask = wl.get("mask")
    ref_field = analytic(wl)
    fields = {}
    for who in ("reference_cuda", "ours"):
        if (who == "reference_cuda" and skip_ref) or (who == "ours" and skip_ours):
            continue
        exp = "ref" if who == "reference_cuda" else "ours"
        conf = workloads.write_case(wl, case_dir, integrator=integrator, exp_name=exp, spp=spp, extra_setting=extra)
        r = (run_reference if who == "reference_cuda" else run_ours)(conf, case_dir, exp, timeout)
        if r["walk_steps"] and r["duration_ms"]:
            r["walk_steps_per_sec"] = r["walk_steps"] / (r["duration_ms"] * 1e-3)
        if r["raw"]:
            fields[who] = load_film(r["raw"], W, H)
            if ref_field is not None:
                r["relmse_vs_analytic"] = relmse(fields[who][..., 0], ref_field, mask)
                if time_frames:
                    r["relmse_vs_time"] = frames_time_curve(case_dir, exp, W, H, ref_field, mask)
        if r["rc"] == 0:
            r.pop("log_tail")
        r.pop("raw")
        line[who] = r
    if len(fields) == 2:
        a, b = fields["ours"], fields["reference_cuda"]
        m = np.ones((H, W), bool) if mask is None else np.asarray(mask, bool)
        ok = m & np.isfinite(a[..., 0]) & np.isfinite(b[..., 0])
        line["fields"] = {"relmse_ours_vs_reference": relmse(a[ok], b[ok]), "mean_abs_diff": float(np.mean(np.abs(a[ok] - b[ok]))),
                          "mean_ours": float(np.mean(a[ok])), "mean_reference": float(np.mean(b[ok])), "pixels": int(ok.sum())}
        ro, rr = line["ours"].get("walk_steps_per_sec"), line["reference_cuda"].get("walk_steps_per_sec")
        if ro and rr:
            line["speedup_vs_reference_cuda"] = ro / rr
    print(json.dumps(line), flush=True)
    return line, fields, case_dir


def pin(out_root):
    """small cases -> b1_<case>.npz: the reference's film (fp32), its walk-step count, and the case parameters"""
    os.makedirs(out_root, exist_ok=True)
    for name in workloads.PIN_CASES:
        line, fields, case_dir = one(name, "uniform", None, out_root, skip_ours=False)
        if "reference_cuda" not in fields:
            print("PIN FAILED", name, flush=True)
            continue
        r = line["reference_cuda"]
        np.savez_compressed(os.path.join(out_root, "b1_%s.npz" % name), field=fields["reference_cuda"].astype(np.float32),
                            ours=fields.get("ours", np.zeros(0)).astype(np.float32), walk_steps=np.int64(r["walk_steps"] or 0),
                            spp=np.int32(line["spp"]), frame=np.asarray(line["frame"], np.int32))
        shutil.rmtree(case_dir, ignore_errors=True)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="c3")
    ap.add_argument("--integrator", default="uniform")
    ap.add_argument("--spp", type=int, default=0)
    ap.add_argument("--depth", type=int, default=0)
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "b1"))
    ap.add_argument("--time-frames", type=int, default=0, help="saveTimeMetricsDuration: dump frames_time/<ms> every N samples and report relMSE(t)")
    ap.add_argument("--pin", action="store_true")
    ap.add_argument("--skip-ref", action="store_true")
    ap.add_argument("--skip-ours", action="store_true")
    ap.add_argument("--keep", action="store_true")
    a = ap.parse_args()
    if a.pin:
        pin(a.out)
    else:
        _, _, cd = one(a.workload, a.integrator, a.spp, a.out, a.time_frames, skip_ref=a.skip_ref, skip_ours=a.skip_ours, depth=a.depth)
        if not a.keep:
            shutil.rmtree(cd, ignore_errors=True)
