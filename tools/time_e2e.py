"""Where does the end-to-end leg of bench.py spend its time? Wall clock per stage of one C-ABI round trip.
  python tools/time_e2e.py [workload] [--torch]     (--torch: with torch's CUDA context and bench.py's 256 MiB flush buffer alive)"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from elaina_b200 import api, workloads
name = next((a for a in sys.argv[1:] if not a.startswith("--")), "c3")
if "--torch" in sys.argv:
    import torch
    torch.cuda.set_device(0)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
wl = workloads.workload(name)
dim = wl["dim"]
st = api.make_settings(wl["frame"], wl["depth"], wl["spp"], wl["eps"], seed=wl["seed"])
keep = None
for rep in range(5):
    t = [time.perf_counter()]
    sc = api.Scene(dim, wl["dverts"], wl["dprims"]); t.append(time.perf_counter())
    sn = api.Scene(dim, wl["nverts"], wl["nprims"]) if wl.get("nverts") is not None else None
    pr = api.Problem(dim, dirichlet=sc, neumann=sn, **workloads.problem_kwargs(wl)); t.append(time.perf_counter())
    it = api.UniformIntegrator(pr, st); t.append(time.perf_counter())
    it.solve(); t.append(time.perf_counter())
    f = it.film(); t.append(time.perf_counter())
    if rep == 0 and "--keep-first" in sys.argv:
        keep = (it, pr)  # like bench.py: the device-timed integrator stays alive while the e2e leg creates its own
    else:
        it.close(); pr.close()
    t.append(time.perf_counter())
    names = ["scene_create", "problem_create", "integrator_create", "solve", "film_download", "close"]
    print(rep, " ".join("%s=%.1fms" % (n, 1e3 * (b - a)) for n, a, b in zip(names, t[:-1], t[1:])), "device_ms=%.1f" % it.last_device_ms, flush=True)
