"""Where does the end-to-end leg of bench.py spend its time? Wall clock per stage of one C-ABI round trip on the C3 workload."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from elaina_b200 import api
wl = bench.workload("c3")
st = api.make_settings(wl["frame"], wl["depth"], wl["spp"], wl["eps"], seed=wl["seed"])
for rep in range(4):
    t = [time.perf_counter()]
    sc = api.Scene(2, wl["dverts"], wl["dprims"]); t.append(time.perf_counter())
    pr = api.Problem(2, dirichlet=sc, **bench.problem_kwargs(wl)); t.append(time.perf_counter())
    it = api.UniformIntegrator(pr, st); t.append(time.perf_counter())
    it.solve(); t.append(time.perf_counter())
    f = it.film(); t.append(time.perf_counter())
    it.close(); pr.close(); t.append(time.perf_counter())
    names = ["scene_create", "problem_create", "integrator_create", "solve", "film_download", "close"]
    print(rep, " ".join("%s=%.1fms" % (n, 1e3 * (b - a)) for n, a, b in zip(names, t[:-1], t[1:])), "device_ms=%.1f" % it.last_device_ms, flush=True)
