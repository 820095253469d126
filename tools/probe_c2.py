"""Why is data/fille (c2) 10x slower per walk-step than data/ladybug (c1)? Timing probes on the GPU."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from elaina_b200 import api, workloads

def solve(wl, spp=16, **over):
    wl = dict(wl, **over)
    dim = wl["dim"]
    sc = api.Scene(dim, wl["dverts"], wl["dprims"])
    sn = api.Scene(dim, wl["nverts"], wl["nprims"]) if wl.get("nverts") is not None else None
    pr = api.Problem(dim, dirichlet=sc, neumann=sn, **workloads.problem_kwargs(wl))
    it = api.UniformIntegrator(pr, api.make_settings(wl["frame"], wl["depth"], spp, wl["eps"], seed=1))
    it.solve(); it.solve()
    return it.last_walk_steps / (it.last_device_ms * 1e-3) / 1e6, it.last_device_ms, it.last_walk_steps / (wl["frame"][0] * wl["frame"][1] * spp)

def nearest_rate(wl, pts):
    sc = api.Scene(2, wl["dverts"], wl["dprims"])
    sc.nearest(pts[:1024])
    t = time.perf_counter(); sc.nearest(pts); dt = time.perf_counter() - t
    return len(pts) / dt / 1e6

for name in (() if "--decades" in sys.argv else ("c1", "c2")):
    wl = workloads.workload(name)
    print(name, "as configured: %.0f M steps/s, %.0f ms, %.1f steps/walk" % solve(wl), flush=True)
    print(name, "depth 64:      %.0f M steps/s, %.0f ms, %.1f steps/walk" % solve(wl, depth=64), flush=True)
    print(name, "scale 250:     %.0f M steps/s, %.0f ms, %.1f steps/walk" % solve(wl, grid_scale=250.0), flush=True)
    print(name, "scale 300:     %.0f M steps/s, %.0f ms, %.1f steps/walk" % solve(wl, grid_scale=300.0), flush=True)
    print(name, "scale 320:     %.0f M steps/s, %.0f ms, %.1f steps/walk" % solve(wl, grid_scale=320.0), flush=True)
    rng = np.random.RandomState(0)
    inner = (rng.rand(1 << 20, 2) * 400 + 50).astype(np.float32)
    band = np.stack([rng.rand(1 << 20) * 80 - 85, rng.rand(1 << 20) * 680 - 90], 1).astype(np.float32)
    wall = np.stack([np.full(1 << 20, -89.0), rng.rand(1 << 20) * 680 - 90], 1).astype(np.float32)
    print(name, "nearest probe M queries/s (incl. H2D/D2H): inner %.1f band %.1f wall %.1f" % (nearest_rate(wl, inner), nearest_rate(wl, band), nearest_rate(wl, wall)), flush=True)


def decade_profile(name="c2", spp=16, depth=None):
    """needs a library built with -DWOST_TRAV_DEBUG (make PTXAS=-DWOST_TRAV_DEBUG): trips / queries per decade of max |coordinate|"""
    import ctypes
    from elaina_b200 import _lib
    wl = workloads.workload(name)
    if depth:
        wl["depth"] = depth
    sc = api.Scene(2, wl["dverts"], wl["dprims"]); sn = api.Scene(2, wl["nverts"], wl["nprims"])
    pr = api.Problem(2, dirichlet=sc, neumann=sn, **workloads.problem_kwargs(wl))
    it = api.UniformIntegrator(pr, api.make_settings(wl["frame"], wl["depth"], spp, wl["eps"], seed=1))
    out = (ctypes.c_uint * 48)()
    _lib.load().wost_debug_counters(it.h, out)
    it.solve()
    _lib.load().wost_debug_counters(it.h, out)
    names = ["<600", "<2e3", "<1e4", "<1e5", "<1e6", "<1e7", "<1e8", ">=1e8"]
    print(name, "device ms %.0f, full/restricted %s" % (it.last_device_ms, it.traversal_stats()))
    for i, n in enumerate(names):
        q, t = out[8 + i], out[i]
        cyc = np.frombuffer(out, np.uint32)[16:].view(np.uint64)
        print("  %6s queries %12d trips %14d trips/query %.1f   warp-Mcycles: trip-iterations %.0f tail-iterations %.0f" % (n, q, t, t / max(1, q), cyc[i] / 1e6, cyc[8 + i] / 1e6))


if "--decades" in sys.argv:
    for d in (64, 96, 128):
        print("depth", d)
        decade_profile(depth=d)
