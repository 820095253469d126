#!/usr/bin/env python
"""BASELINE B2: the reference's own SNCH-LBVH (oracle/_ref/libref_lbvh.so = its headers host-compiled where they lie, Thrust CPP
backend, TEST INFRASTRUCTURE) timed on the host cores as the query-level CPU baseline, beside the device query kernels of
libwost.so (`wost_lbvh_*_dev`, the renderDirichletSDF / renderSilhouetteSDF channel kernels' traversal) on the same queries.

  python tools/b2_queries.py [--workloads c3 c1 c4-small] [--grid 1024] [--cpu-queries 40000]

One JSON line per (workload, query kind):
  cpu_baseline {value (queries/s), cores: 1, kind: "reference", sample}   -- the reference's code, single thread (its host
               build is sequential), on a bounded sample of the evaluation grid
  value        queries/s on the GPU over the whole grid, inputs resident in HBM, CUDA events on the launching stream
  e2e          the same through the host-buffer entry points (H2D of the query points + D2H of the results inside the timed region)
  parity       the GPU results on the CPU sample are bit-identical to the reference's (indices and distances)
  roofline     algorithmic bytes per query (ceil(log2 N) visits x 56 / 72 B + leaf 48 / 76 B; silhouette 88 / 112 B
               per visit + 144 / 280 B) x queries/s against the measured HBM peak
"""
import argparse
import ctypes
import json
import math
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from elaina_b200 import workloads  # noqa: E402


def grid_points(wl, n):
    s = wl["grid_scale"]
    g = (2.0 * (np.arange(n, dtype=np.float64) + 0.5) / n - 1.0) * s
    X, Y = np.meshgrid(g + wl["grid_pos"][0], g + wl["grid_pos"][1])
    if wl["dim"] == 2:
        return np.stack([X.ravel(), Y.ravel()], 1).astype(np.float32)
    return np.stack([X.ravel(), Y.ravel(), np.zeros(n * n)], 1).astype(np.float32)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workloads", nargs="+", default=["c3", "c1", "c4-small"])
    ap.add_argument("--grid", type=int, default=1024)
    ap.add_argument("--cpu-queries", type=int, default=40000)
    a = ap.parse_args()
    import torch
    from elaina_b200 import _lib, api
    from oracle_lib import RefLBVH, have_ref
    L = _lib.load()
    dev = torch.device("cuda", 0)
    peak = 6650.0
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        pass
    ref = RefLBVH() if have_ref() else None
    for name in a.workloads:
        wl = workloads.workload(name)
        dim = wl["dim"]
        for which in ("dirichlet", "neumann"):
            v, p = (wl["dverts"], wl["dprims"]) if which == "dirichlet" else (wl.get("nverts"), wl.get("nprims"))
            if v is None:
                continue
            sg = api.Scene(dim, v, p)
            sr = ref.scene(dim, v, p) if ref is not None else None
            q = grid_points(wl, a.grid)
            rng = np.random.RandomState(1)
            sel = rng.choice(len(q), min(a.cpu_queries, len(q)), replace=False)
            qs = np.ascontiguousarray(q[sel])
            d = rng.randn(len(q), dim).astype(np.float32)
            d /= np.linalg.norm(d, axis=1, keepdims=True)
            md = np.full(len(q), 4.0 * wl["grid_scale"], np.float32)
            N = len(p)
            depth = max(1, math.ceil(math.log2(max(2, N))))
            kinds = ["nearest", "ray"] + (["silhouette"] if which == "neumann" or name != "c3" else [])
            tq, td, tm = torch.from_numpy(q).to(dev), torch.from_numpy(d).to(dev), torch.from_numpy(md).to(dev)
            idx = torch.empty(len(q), dtype=torch.int32, device=dev)
            dist = torch.empty(len(q), dtype=torch.float32, device=dev)
            hit = torch.empty(len(q), dtype=torch.int32, device=dev)
            uv = torch.empty(len(q) * 2, dtype=torch.float32, device=dev)
            stream = torch.cuda.current_stream().cuda_stream
            for kind in kinds:
                bpq = {"nearest": depth * (56 if dim == 2 else 72) + (48 if dim == 2 else 76),
                       "ray": depth * (56 if dim == 2 else 72) + (48 if dim == 2 else 76),
                       "silhouette": depth * (88 if dim == 2 else 112) + (144 if dim == 2 else 280)}[kind]

                def gpu_once():
                    if kind == "nearest":
                        _lib.check(L.wost_lbvh_nearest_dev(sg.h, tq.data_ptr(), len(q), idx.data_ptr(), dist.data_ptr(), ctypes.c_void_p(stream)))
                    elif kind == "silhouette":
                        _lib.check(L.wost_lbvh_silhouette_dev(sg.h, tq.data_ptr(), len(q), 0, dist.data_ptr(), ctypes.c_void_p(stream)))
                    else:
                        _lib.check(L.wost_lbvh_ray_dev(sg.h, tq.data_ptr(), td.data_ptr(), tm.data_ptr(), len(q), hit.data_ptr(), dist.data_ptr(), uv.data_ptr(),
                                                       idx.data_ptr(), ctypes.c_void_p(stream)))
                for _ in range(3):
                    gpu_once()
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                reps = 5
                e0.record()
                for _ in range(reps):
                    gpu_once()
                e1.record()
                torch.cuda.synchronize()
                gpu_qps = reps * len(q) / (e0.elapsed_time(e1) * 1e-3)
                # end to end from host buffers
                t0 = time.perf_counter()
                if kind == "nearest":
                    gi, gd = sg.nearest(q)
                elif kind == "silhouette":
                    gd = sg.silhouette(q, False)
                else:
                    gh, gd, guv, gi = sg.ray(q, d, md)
                e2e_qps = len(q) / (time.perf_counter() - t0)
                line = {"baseline": "B2", "workload": name, "mesh": which, "prims": int(N), "query": kind, "metric": "queries_per_sec", "value": gpu_qps,
                        "unit": "queries/s", "queries": len(q), "e2e": {"value": e2e_qps, "unit": "queries/s", "h2d_bytes": int(q.nbytes * (3 if kind == "ray" else 1)),
                                                                        "d2h_bytes": int(len(q) * (8 if kind != "ray" else 20))},
                        "roofline": {"bound": "hbm", "bytes_per_query": bpq, "achieved": gpu_qps * bpq / 1e9, "peak": peak, "unit": "GB/s", "frac": gpu_qps * bpq / 1e9 / peak}}
                if sr is not None:
                    t0 = time.perf_counter()
                    if kind == "nearest":
                        ri, rd = sr.nearest(qs)
                        same = bool(np.array_equal(ri, gi[sel]) and np.array_equal(rd.view(np.uint32), gd[sel].view(np.uint32)))
                    elif kind == "silhouette":
                        rd = sr.silhouette(qs, False)
                        same = bool(np.array_equal(rd.view(np.uint32), gd[sel].view(np.uint32)))
                    else:
                        rh, rt, ruv, ro = sr.ray(qs, np.ascontiguousarray(d[sel]), np.ascontiguousarray(md[sel]))
                        same = bool(np.array_equal(rh, gh[sel]) and np.array_equal(ro, gi[sel]) and np.array_equal(rt.view(np.uint32), gd[sel].view(np.uint32)))
                    dt = time.perf_counter() - t0
                    line["cpu_baseline"] = {"value": len(qs) / dt, "unit": "queries/s", "cores": 1, "kind": "reference",
                                            "sample": "%d random points of the %dx%d evaluation grid, the reference's snch_lbvh headers host-compiled (oracle/_ref)" % (len(qs), a.grid, a.grid)}
                    line["parity"] = {"bit_identical_to_reference_on_sample": same}
                    line["speedup_vs_reference_cpu"] = gpu_qps / (len(qs) / dt)
                print(json.dumps(line), flush=True)
            sg.close()


if __name__ == "__main__":
    main()
