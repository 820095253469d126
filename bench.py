#!/usr/bin/env python
"""bench.py -- walk-steps/sec of the WoSt walker hot path (BASELINE.json metric) on N B200s of one node.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c3|c3-small|c4] [--impl ours|reference]

One "step" = one solve() of the workload (all pixels x all samples-per-pixel x all bounces) -- the call
UniformIntegrator::solve() of the reference (integrator/uniform/integrator.cu:666-672).
  value     walk-steps/s, device-timed (CUDA events on the integrator's stream inside wost_solve, max over ranks),
            scene/BVH/colours already resident in HBM.
  e2e       the same metric through the C ABI from HOST buffers, per step: wost_scene_create (H2D of vertices and
            primitives + LBVH build) -> wost_problem_create (H2D colours) -> wost_integrator_create -> wost_solve ->
            wost_film_download (D2H), timed by wall clock around the whole sequence (max over ranks).
  roofline  algorithmic bytes per walk-step x walk-steps / summed device time of the
            fused walk-step kernel launches, against MEASURED_PEAKS.json hbm_gbs.
  cpu_baseline  the scalar oracle (oracle/libwost_oracle.so, TEST INFRASTRUCTURE) timed on the host cores on a bounded
            sample of the same workload (a band of pixel rows at reduced spp).
Multi-GPU: evaluation points shard by interleaved row tiles; no data-path collective; the final film is
summed over ranks with one NCCL all-reduce (outside the device-timed region, inside e2e).
--impl reference: the reference has no CPU implementation of the walker (CUDA-only) and its CUDA build is
unbuildable offline; per the tier rules this arm times the CPU oracle port with all host threads.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np

METRIC = "walk_steps_per_sec"
UNIT = "walk-steps/s"

from elaina_b200.workloads import workload, problem_kwargs, guided_settings  # noqa: E402  (BASELINE.json configurations)


class ClockSampler(threading.Thread):
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md 'clocks' line)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        super().__init__(daemon=True)
        self.gpu = gpu_index
        self.rows = []
        self.proc = None

    def run(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            for line in self.proc.stdout:
                self.rows.append([x.strip() for x in line.split(",")])
        except Exception:
            pass

    def stop(self):
        if self.proc is not None:
            try:
                self.proc.terminate()
                self.proc.wait(timeout=5)
            except Exception:
                pass
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[1]))
                mx.append(float(r[2]))
            except Exception:
                continue
            for name, col in (("hw_slowdown", 5), ("hw_thermal_slowdown", 6), ("sw_thermal_slowdown", 7), ("sw_power_cap", 8)):
                if len(r) > col and r[col].lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons), "samples": len(sm)}


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def cpu_oracle_sample(wl, budget_s=12.0, nthreads=0):
    """Time the scalar oracle on a bounded sample of the workload: rows around the frame centre, spp reduced."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from oracle_lib import Oracle
    orc = Oracle()
    orc.set_math_mode(0)
    dim = wl["dim"]
    prob = dict(dim=dim, dirichlet=orc.scene(dim, wl["dverts"], wl["dprims"]),
                neumann=orc.scene(dim, wl["nverts"], wl["nprims"]) if wl.get("nverts") is not None else None, **problem_kwargs(wl))
    W, H = wl["frame"]
    cores = nthreads if nthreads > 0 else (os.cpu_count() or 1)
    st = dict(frame_w=W, frame_h=H, max_walking_depth=wl["depth"], samples_per_pixel=1, epsilon_shell=wl["eps"], seed=wl["seed"])
    rows = 2
    total_steps, total_t = 0, 0.0
    r0 = H // 2
    # grow the sample until ~budget_s of CPU work has been timed
    while total_t < budget_s and r0 + rows <= H:
        t0 = time.perf_counter()
        _, steps = orc.solve_uniform(prob, st, pixel_begin=r0 * W, pixel_end=(r0 + rows) * W, nthreads=cores)
        dt = time.perf_counter() - t0
        total_steps += steps
        total_t += dt
        r0 += rows
        if dt < budget_s / 8:
            rows = min(rows * 2, H - r0) if H - r0 > 0 else rows
        if rows <= 0:
            break
    sample = "%d pixel rows of the %dx%d frame from row %d, 1 spp, depth %d (%d walk-steps)" % (r0 - H // 2, W, H, H // 2, wl["depth"], total_steps)
    return total_steps / max(total_t, 1e-9), cores, sample, total_steps, total_t


def run_reference_arm(args, wl, rank, world):
    if rank != 0:
        return
    vals = []
    sample = ""
    cores = os.cpu_count() or 1
    for i in range(args.warmup + args.steps):
        v, cores, sample, steps, t = cpu_oracle_sample(wl, budget_s=max(2.0, 60.0 / max(1, args.warmup + args.steps)))
        if i >= args.warmup:
            vals.append((steps, t))
    steps = sum(s for s, _ in vals)
    t = sum(t for _, t in vals)
    value = steps / max(t, 1e-9)
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * t / max(1, args.steps), "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "config": {"workload": wl["desc"], "note": "reference walker is CUDA-only and unbuildable offline; CPU oracle port timed (DESIGN.md)"},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": "per step: " + sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c3")
    ap.add_argument("--spp", type=int, default=0, help="override samples per pixel (reported in config)")
    ap.add_argument("--spp-batch", type=int, default=0)
    ap.add_argument("--walk-mode", default="persistent", choices=["persistent", "wavefront", "reforder", "persistent-reforder"],
                    help="reforder: closest-point queries in the reference's visit order (stand-in for the reference's traversal cost)")
    ap.add_argument("--integrator", default="uniform", choices=["uniform", "guided"],
                    help="guided: GuidedIntegrator with the reference's n.json settings (train spp = spp/4, guided depth 10, learned selection)")
    ap.add_argument("--mlp-impl", type=int, default=0, help="guided: 0 = fused tcgen05 bounce kernel, 1 = validation mode (SIMT MLP through HBM)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true", help="profiling runs only: skip the end-to-end leg (line carries e2e: null)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    wl = workload(args.workload)
    if args.spp > 0:
        wl["desc"] = wl["desc"].replace("%d spp" % wl["spp"], "%d spp" % args.spp)
        wl["spp"] = args.spp
    if args.impl == "reference":
        run_reference_arm(args, wl, rank, world)
        return

    import torch
    import torch.distributed as dist
    from elaina_b200 import api
    from elaina_b200 import dist as wdist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA GPU (there is no CPU fallback for the walker)")
    torch.cuda.set_device(local_rank)
    api.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    dev = torch.device("cuda", local_rank)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    dim = wl["dim"]
    W, H = wl["frame"]
    # strong scaling: the frame is fixed, rows are dealt to ranks in interleaved tiles of 4 rows (pixel-keyed RNG ->
    # the summed film does not depend on N)
    shard = wdist.shard_kwargs(rank, world, stride_rows=4)
    guided = args.integrator == "guided"
    gkw = {}
    if guided:
        # data/{ladybug,fille}/n.json: trainSppCount = spp/4, maxGuidedDepth 10 in both phases, uniformFraction 0.5 (= learned selection)
        gkw = dict(guided_settings(wl), mlp_impl=args.mlp_impl)
    settings = api.make_settings((W, H), wl["depth"], wl["spp"], wl["eps"], seed=wl["seed"], spp_batch=args.spp_batch, walk_mode={"persistent": 1, "reforder": 2, "wavefront": 3, "persistent-reforder": 4}[args.walk_mode], **shard, **gkw)
    Integrator = (lambda p, st: api.GuidedIntegrator(p, st, wl["network"])) if guided else api.UniformIntegrator

    def make_problem():
        g_d = api.Scene(dim, wl["dverts"], wl["dprims"])
        g_n = api.Scene(dim, wl["nverts"], wl["nprims"]) if wl.get("nverts") is not None else None
        return api.Problem(dim, dirichlet=g_d, neumann=g_n, **problem_kwargs(wl))

    prob = make_problem()
    it = Integrator(prob, settings)
    if guided and world > 1:
        wdist.install_gradient_allreduce(it, dev)  # NCCL sum of the gradient slab on the integrator's stream, per batch
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    film_t = None
    if world > 1:
        film_t = torch.empty(H * W * 5, dtype=torch.float32, device=dev)

    # ---- device-timed: K solves, scene resident
    for _ in range(args.warmup):
        it.solve()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    barrier()
    dev_ms, steps_total, launches, kern_ms, kern_launches = 0.0, 0, 0, 0.0, 0
    for _ in range(args.steps):
        flush.zero_()
        torch.cuda.synchronize()
        it.solve()
        dev_ms += it.last_device_ms
        steps_total += it.last_walk_steps
        launches += it.launch_count
        n, ms = it.step_kernel_time
        kern_launches += n
        kern_ms += ms
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    dev_ms = max_over_ranks(dev_ms)
    kern_ms_max = max_over_ranks(kern_ms)
    steps_all = sum_over_ranks(float(steps_total))
    value = steps_all / (dev_ms * 1e-3)

    # ---- end to end through the C ABI from host buffers
    h2d = wl["dverts"].nbytes + wl["dprims"].nbytes + (wl["color_dirichlet"].nbytes if wl.get("color_dirichlet") is not None else 0)
    if wl.get("nverts") is not None:
        h2d += wl["nverts"].nbytes + wl["nprims"].nbytes + (wl["color_neumann"].nbytes if wl.get("color_neumann") is not None else 0)
    if wl.get("source") is not None:
        h2d += wl["source"].nbytes
    d2h = W * H * 5 * 4
    e2e_steps = 0

    def e2e_once():
        t_ = [time.perf_counter()]
        p = make_problem()
        t_.append(time.perf_counter())
        i2 = Integrator(p, settings)
        t_.append(time.perf_counter())
        if guided and world > 1:
            wdist.install_gradient_allreduce(i2, dev)
        i2.solve()
        t_.append(time.perf_counter())
        if os.environ.get("WOST_BENCH_DEBUG"):
            print("e2e stages ms: problem %.1f integrator %.1f solve %.1f (device %.1f)" % (1e3 * (t_[1] - t_[0]), 1e3 * (t_[2] - t_[1]), 1e3 * (t_[3] - t_[2]), i2.last_device_ms), file=sys.stderr, flush=True)
        if world > 1:
            # final film gather: shards own disjoint rows and write zeros elsewhere -> one NCCL sum
            ptr = i2.film_device_ptr()
            film = wdist.wrap_device_ptr(ptr, H * W * 5, dev)
            film_t.copy_(film)
            dist.all_reduce(film_t)
            out = film_t.cpu() if rank == 0 else None
            torch.cuda.synchronize()
        else:
            out = i2.film()
        s = i2.last_walk_steps
        i2.close()
        p.close()
        return s, out

    n_e2e = max(1, min(args.steps, 3))
    e2e_value = None
    if not args.no_e2e:
        e2e_once()
        barrier()
        t0 = time.perf_counter()
        for _ in range(n_e2e):
            s, _ = e2e_once()
            e2e_steps += s
        barrier()
        e2e_t = max_over_ranks(time.perf_counter() - t0)
        e2e_value = sum_over_ranks(float(e2e_steps)) / e2e_t

    peak, peak_src = measured_peak()
    bps = wl["bytes_per_step"] + (176 if guided else 0)  # + 12 B network input and 164 B output per guided step
    relmse = None
    if rank == 0 and world == 1 and args.workload in ("c3", "c3-small"):
        # relMSE against the analytic field Re(z^3) over the evaluated pixels (BASELINE metric, second half)
        from elaina_b200 import scenes as _sc
        g = 1.5 * (2.0 * np.arange(W, dtype=np.float64) / W - 1.0)
        X, Y = np.meshgrid(g, g)
        ref = _sc.harmonic_2d(np.stack([X, Y], -1), 3)
        sol = it.film()[..., 0].astype(np.float64)
        m = wl["mask"].astype(bool)
        relmse = float(np.mean((sol[m] - ref[m]) ** 2 / (ref[m] ** 2 + 1e-2)))
    achieved = steps_total * bps / (kern_ms * 1e-3) / 1e9 if kern_ms > 0 else 0.0

    traffic = None
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "traffic_r1d.json")))
        if not guided and tj.get("kernel", "").split("<")[0] in ("k_walk_persistent" if args.walk_mode.startswith("persistent") else "k_step_uniform") and dim == 2:
            traffic = tj["dram_bytes_per_walk_step"] * steps_total / max(1, kern_launches)  # DRAM bytes per launch, from the committed ncu capture
    except Exception:
        pass
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        v, cores, sample, _, _ = cpu_oracle_sample(wl)
        cpu = {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": dev_ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": wl["desc"], "integrator": args.integrator, "walk_mode": args.walk_mode if not guided else ("fused tcgen05 bounce kernel" if args.mlp_impl == 0 else "validation: SIMT MLP through HBM"), "relmse_vs_analytic": relmse, "spp_batch": args.spp_batch or "auto (~16M walks in flight)",
                       "walk_steps_per_step": steps_all / args.steps, "sharding": "interleaved 4-row tiles over %d ranks" % world,
                       "l2": "256 MiB flush between timed steps; live working set (queues + accumulators, >1 GB) exceeds the 126 MB L2"},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h), "steps": n_e2e,
                    "path": "wost_scene_create+LBVH build -> wost_problem_create -> wost_integrator_create -> wost_solve -> wost_film_download"},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                         "kernel": ("k_g_first+k_g_bounce+k_g_finish<%d>" if guided else ("k_walk_persistent<%d>" if args.walk_mode.startswith("persistent") else "k_step_uniform<%d>")) % dim, "bytes_per_walk_step": bps, "launches": int(kern_launches),
                         "avg_launch_ms": kern_ms / max(1, kern_launches), "kernel_share_of_step": kern_ms_max / dev_ms if dev_ms > 0 else None,
                         "peak_source": peak_src},
            "clocks": clocks,
        }
        if guided:
            if os.environ.get("WOST_PROFILE"):
                line["profile_ms_last_solve"] = it.profile
            line["guided"] = dict(it.train_stats, train_spp=gkw["train_spp_count"], network="DenseGrid 8x4 + MLP 32-64-64-64-48 (data/ladybug/n.json)")
        if cpu is not None:
            line["cpu_baseline"] = cpu
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
