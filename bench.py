#!/usr/bin/env python
"""bench.py -- walk-steps/sec of the WoSt walker hot path (BASELINE.json metric) on N B200s of one node.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c3|c1|c2|c4|c4-small|c5] [--integrator uniform|guided]
                  [--impl ours|reference] [--no-sub]

One "step" = one solve() of the workload (all pixels x all samples-per-pixel x all bounces) -- the call
UniformIntegrator::solve() / GuidedIntegrator::solve() of the reference (integrator/uniform/integrator.cu:666-672).
  value     walk-steps/s, device-timed (CUDA events on the integrator's stream inside wost_solve, max over ranks),
            scene/BVH/colours already resident in HBM.
  e2e       the same metric through the C ABI from HOST buffers, per step: wost_scene_create (H2D of vertices and
            primitives + LBVH build) -> wost_problem_create (H2D colours) -> wost_integrator_create -> wost_solve ->
            [N > 1: wost_film_gather to rank 0] -> wost_film_download (D2H), wall clock around the whole sequence, max over ranks.
  roofline  algorithmic bytes per walk-step x walk-steps / summed device time of the fused
            walk-step kernel launches, against MEASURED_PEAKS.json hbm_gbs. The tree of the headline workload is L2-resident
            and the kernel is issue-bound: `bound_actual` and the `issue` block say so with the figures of the committed ncu
            capture of the same kernel (profiles/), labelled as such.
  sub       (default run only) the other single-GPU configurations of BASELINE.json, each with its own value / e2e / roofline:
            C3 guided (network re-initialised before every timed solve, like elaina-exec), C1 on the real data/ladybug
            geometry, C4-small guided. Reduced spp where stated in their `workload` string (walk-steps/s does not depend on spp).
  parity_check  rows of the frame solved by the CPU oracle and by the GPU (same Philox streams, 1 spp): films must be
            bit-identical -- the cpu_baseline leg doubles as a parity test at BASELINE size.
  cpu_baseline  the scalar oracle (oracle/libwost_oracle.so, TEST INFRASTRUCTURE) timed on the host cores on a bounded
            sample of the same workload.
Multi-GPU: evaluation points shard by interleaved row tiles; no data-path collective in the uniform
integrator; guided training all-reduces the gradient slab per optimiser step and the final film is gathered to rank 0 --
both NCCL calls are issued by libwost.so itself (csrc/comm.cu); torch.distributed carries the NCCL id, barriers and timing.
--impl reference: the reference has no CPU implementation of the walker (CUDA-only); per the tier rules this arm times the
CPU oracle port (pinned against the reference's own CUDA build, tests/test_oracle_b1_pin.py) with all host threads, on the
whole frame at 4 spp per step. When the reference's own CUDA binary (baseline/_ref/elaina-exec-ref) and a
GPU are present, the line also carries `reference_cuda`: its walk-steps/s on the same workload on this box.
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


def relmse_analytic(wl, film):
    """relMSE against the analytic field Re(z^3) over the evaluated pixels (BASELINE metric, second half); c3 family only"""
    if wl["name"] not in ("c3", "c3-small", "c5"):
        return None
    from elaina_b200 import scenes as _sc
    W, H = wl["frame"]
    g = 1.5 * (2.0 * np.arange(W, dtype=np.float64) / W - 1.0)
    X, Y = np.meshgrid(g, g)
    ref = _sc.harmonic_2d(np.stack([X, Y], -1), 3)
    sol = film[..., 0].astype(np.float64)
    m = wl["mask"].astype(bool)
    return float(np.mean((sol[m] - ref[m]) ** 2 / (ref[m] ** 2 + 1e-2)))


# ----------------------------------------------------------------------------------------------------------------------
# CPU oracle (TEST INFRASTRUCTURE): the cpu_baseline leg, the parity check and the --impl reference arm
def _oracle_problem(wl):
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from oracle_lib import Oracle
    orc = Oracle()
    orc.set_math_mode(0)
    dim = wl["dim"]
    prob = dict(dim=dim, dirichlet=orc.scene(dim, wl["dverts"], wl["dprims"]),
                neumann=orc.scene(dim, wl["nverts"], wl["nprims"]) if wl.get("nverts") is not None else None, **problem_kwargs(wl))
    return orc, prob


def cpu_oracle_sample(wl, budget_s=12.0, nthreads=0, keep_field=False):
    """Time the scalar oracle on a bounded sample of the workload: rows around the frame centre at 1 spp.
    -> (steps/s, cores, sample description, steps, seconds, (row0, row1, field[rows*W, 3]) | None)"""
    orc, prob = _oracle_problem(wl)
    W, H = wl["frame"]
    cores = nthreads if nthreads > 0 else (os.cpu_count() or 1)
    st = dict(frame_w=W, frame_h=H, max_walking_depth=wl["depth"], samples_per_pixel=1, epsilon_shell=wl["eps"], seed=wl["seed"])
    rows = 2
    total_steps, total_t = 0, 0.0
    r0 = H // 2
    first = None
    while total_t < budget_s and r0 + rows <= H:  # grow the sample until ~budget_s of CPU work has been timed
        t0 = time.perf_counter()
        sol, steps = orc.solve_uniform(prob, st, pixel_begin=r0 * W, pixel_end=(r0 + rows) * W, nthreads=cores)
        dt = time.perf_counter() - t0
        if keep_field:  # the chunk with the most evaluated pixels is kept for the parity check
            npx = int(wl["mask"][r0:r0 + rows].sum()) if wl.get("mask") is not None else rows * W
            if first is None or npx > first[3]:
                first = (r0, r0 + rows, np.asarray(sol, np.float32).reshape(-1, 3)[r0 * W:(r0 + rows) * W].copy(), npx)
        total_steps += steps
        total_t += dt
        r0 += rows
        if dt < budget_s / 8:
            rows = min(rows * 2, H - r0) if H - r0 > 0 else rows
        if rows <= 0:
            break
    sample = "%d pixel rows of the %dx%d frame from row %d, 1 spp, depth %d (%d walk-steps)" % (r0 - H // 2, W, H, H // 2, wl["depth"], total_steps)
    return total_steps / max(total_t, 1e-9), cores, sample, total_steps, total_t, first


REF_CUDA_RECORD = os.path.join(ROOT, "baseline", "_ref", "last_reference_cuda.json")  # written by --impl reference, read by the repo arm


def run_reference_cuda(wl, spp, timeout=600, integrator="uniform"):
    """the reference's own CUDA elaina-exec on this box's GPU, same workload through the same conf.json keys."""
    ref_bin = os.path.join(ROOT, "baseline", "_ref", "elaina-exec-ref")
    if not os.path.exists(ref_bin):
        return {"unavailable": "baseline/_ref/elaina-exec-ref not built (baseline/build_ref_cuda.sh needs the reference tree)"}
    try:
        if subprocess.run(["nvidia-smi", "-L"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, timeout=30).returncode != 0:
            return {"unavailable": "no GPU"}
    except Exception:
        return {"unavailable": "no GPU"}
    import shutil
    import tempfile
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import b1_run
    from elaina_b200 import workloads as wls
    case_dir = tempfile.mkdtemp(prefix="b1_bench_")
    try:
        conf = wls.write_case(wl, case_dir, integrator=integrator, exp_name="ref", spp=spp)
        r = b1_run.run_reference(conf, case_dir, "ref", timeout)
        out = {"rc": r["rc"], "walk_steps": r["walk_steps"], "duration_ms": r["duration_ms"], "spp": spp, "integrator": integrator,
               "sample": "whole frame at %d spp through baseline/_ref/elaina-exec-ref (reference sources built by baseline/build_ref_cuda.sh, "
                         "walk-step counter patch only), duration = result.json (exec.cu:157-158)" % spp}
        if r["walk_steps"] and r["duration_ms"]:
            out["value"] = r["walk_steps"] / (r["duration_ms"] * 1e-3)
            out["unit"] = UNIT
        if r["raw"]:
            W, H = wl["frame"]
            film = np.fromfile(r["raw"], np.float32).reshape(H, W, 5)
            out["relmse_vs_analytic"] = relmse_analytic(wl, film[..., :3] / np.maximum(film[..., 4:5], 1e-30))
        if r["rc"] != 0:
            out["log_tail"] = r["log_tail"][-400:]
        return out
    except Exception as e:  # noqa: BLE001
        return {"unavailable": "run failed: %r" % (e,)}
    finally:
        shutil.rmtree(case_dir, ignore_errors=True)


def run_reference_arm(args, wl, rank, world):
    if rank != 0:
        return
    orc, prob = _oracle_problem(wl)
    W, H = wl["frame"]
    cores = os.cpu_count() or 1
    ref_spp = 4
    st = dict(frame_w=W, frame_h=H, max_walking_depth=wl["depth"], samples_per_pixel=ref_spp, epsilon_shell=wl["eps"], seed=wl["seed"])
    # keep the whole --steps/--warmup run within a few minutes: whole frame at 4 spp when that fits, else a band of centre rows
    t0 = time.perf_counter()
    orc.solve_uniform(prob, dict(st, samples_per_pixel=1), pixel_begin=(H // 2) * W, pixel_end=(H // 2 + 8) * W, nthreads=cores)
    est_frame_s = (time.perf_counter() - t0) / 8.0 * H * ref_spp  # centre rows are the longest ones of a masked frame: an overestimate
    budget = 240.0 / max(1, args.steps + args.warmup)
    rows = H if est_frame_s <= budget else max(8, int(H * budget / est_frame_s) // 8 * 8)
    pb, pe = ((H - rows) // 2) * W, ((H - rows) // 2 + rows) * W
    steps_total, t_total = 0, 0.0
    for i in range(args.warmup + args.steps):
        t0 = time.perf_counter()
        _, steps = orc.solve_uniform(prob, st, pixel_begin=pb, pixel_end=pe, nthreads=cores)
        dt = time.perf_counter() - t0
        if i >= args.warmup:
            steps_total += steps
            t_total += dt
    value = steps_total / max(t_total, 1e-9)
    sample = "per step: %s of the %dx%d frame at %d spp, depth %d (%d walk-steps per step)" % (
        "the whole frame" if rows == H else "%d centre rows" % rows, W, H, ref_spp, wl["depth"], steps_total // max(1, args.steps))
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * t_total / max(1, args.steps), "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": wl["desc"], "integrator": "uniform",
                       "sample": "CPU arm: " + sample + "; reference_cuda: the whole frame at reduced spp (a rate, walk-steps/s)",
                       "note": "the reference walker is CUDA-only (no CPU path); this arm times the CPU oracle port -- pinned against the "
                       "reference's own CUDA build, tests/test_oracle_b1_pin.py -- on all host threads; reference_cuda = the reference's own binary on this box's GPU"},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    if not args.no_reference_cuda:
        line["reference_cuda"] = run_reference_cuda(wl, spp=min(wl["spp"], 32))
        record = {"workload": wl["name"], "box_time": time.time(), "head": line["reference_cuda"], "sub": []}
        if args.workload == "c3" and args.spp == 0 and "value" in line["reference_cuda"]:
            # the repo arm's sub-workloads (same names), at reduced spp: the reference's rate does not depend on spp beyond the
            # guided integrator's training share, which train spp = spp / 4 keeps proportional
            for name, integ, spp in (("c3", "guided", 32), ("c1", "uniform", 16)):
                r = run_reference_cuda(workload(name), spp=spp, integrator=integ)
                r["name"] = "%s-%s" % (name, integ)
                record["sub"].append(r)
            line["reference_cuda_sub"] = record["sub"]
        try:
            with open(REF_CUDA_RECORD, "w") as f:
                json.dump(record, f)
        except OSError:
            pass
    print(json.dumps(line), flush=True)


def vs_reference_cuda(head_value, sub):
    """ratio of this arm's device-timed rates to the reference's own CUDA binary, when `--impl reference` ran on THIS box before
    (the driver runs the reference arm first); None otherwise. Both are rates in walk-steps/s on the same workload."""
    try:
        rec = json.load(open(REF_CUDA_RECORD))
    except Exception:
        return None
    if rec.get("workload") != "c3" or abs(time.time() - rec.get("box_time", 0)) > 6 * 3600 or "value" not in rec.get("head", {}):
        return None
    out = {"c3-uniform": {"ratio": head_value / rec["head"]["value"], "reference_cuda_value": rec["head"]["value"], "reference_spp": rec["head"].get("spp")}}
    ours = {r.get("name"): r for r in sub}
    for r in rec.get("sub", []):
        if "value" in r and r.get("name") in ours:
            out[r["name"]] = {"ratio": ours[r["name"]]["value"] / r["value"], "reference_cuda_value": r["value"], "reference_spp": r.get("spp")}
    out["note"] = "device-timed walk-steps/s of this arm / walk-steps/s of baseline/_ref/elaina-exec-ref (result.json duration) measured earlier on this box"
    return out


# ----------------------------------------------------------------------------------------------------------------------
def issue_profile(wl, guided, walk_mode):
    """issue-slot figures of the dominant kernel from the committed ncu capture of the same kernel (NOT measured in this run)"""
    if guided or wl["name"] not in ("c3", "c3-small", "c5", "c4", "c4-slices") or not walk_mode.startswith("persistent"):
        return None  # committed captures: k_walk_persistent<2, lean variant> on the C3 geometry, k_walk_persistent<3> on the 1.31 M-triangle C4 tree
    for name in (("traffic_r3_c4.json",) if wl["name"].startswith("c4") else ("traffic_r2.json", "traffic_r1d.json")):
        p = os.path.join(ROOT, "profiles", name)
        if not os.path.exists(p):
            continue
        try:
            tj = json.load(open(p))
        except Exception:
            continue
        return tj
    return None


class Bench:
    def __init__(self, args, rank, local_rank, world):
        import torch
        import torch.distributed as dist
        from elaina_b200 import api
        from elaina_b200 import dist as wdist
        self.torch, self.dist, self.api, self.wdist = torch, dist, api, wdist
        self.args, self.rank, self.world = args, rank, world
        self.dev = torch.device("cuda", local_rank)
        self.comm = wdist.Comm.from_torch(self.dev) if world > 1 else None
        self.flush = torch.empty(256 << 20, dtype=torch.uint8, device=self.dev)  # > 126 MB L2
        self.pinned = {}  # (H, W) -> pinned host buffer the e2e leg downloads the film into

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def reduce(self, x, op):
        if self.world == 1:
            return x
        t = self.torch.tensor([x], dtype=self.torch.float64, device=self.dev)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX if op == "max" else self.dist.ReduceOp.SUM)
        return float(t.item())

    def run(self, wl, integrator, steps, warmup, do_e2e=True, walk_mode="persistent", spp_batch=0, mlp_impl=0, sampler=None):
        """device-timed solves + end-to-end solves of one workload -> dict (rank 0 returns the full record)"""
        api, wdist, torch = self.api, self.wdist, self.torch
        rank, world = self.rank, self.world
        dim = wl["dim"]
        W, H = wl["frame"]
        guided = integrator == "guided"
        # strong scaling: the frame is fixed, rows are dealt to ranks in interleaved tiles of 4 rows (pixel-keyed RNG ->
        # the gathered film does not depend on N)
        shard = wdist.shard_kwargs(rank, world, stride_rows=4)
        gkw = dict(guided_settings(wl), mlp_impl=mlp_impl) if guided else {}
        settings = api.make_settings((W, H), wl["depth"], wl["spp"], wl["eps"], seed=wl["seed"], spp_batch=spp_batch,
                                     walk_mode={"persistent": 1, "reforder": 2, "wavefront": 3, "persistent-reforder": 4}[walk_mode], **shard, **gkw)

        slices = wl.get("slices") or [{}]  # a slice set (c4-slices): one problem + integrator per evaluation plane, geometry shared

        def make_problems():
            g_d = api.Scene(dim, wl["dverts"], wl["dprims"])
            g_n = api.Scene(dim, wl["nverts"], wl["nprims"]) if wl.get("nverts") is not None else None
            return [api.Problem(dim, dirichlet=g_d, neumann=g_n, **problem_kwargs(dict(wl, **ov))) for ov in slices]

        def make_integrator(p):
            it_ = api.GuidedIntegrator(p, settings, wl["network"]) if guided else api.UniformIntegrator(p, settings)
            if guided and self.comm is not None:
                self.comm.attach(it_)  # ncclAllReduce of the gradient slab on the integrator's stream, per optimiser step (csrc/comm.cu)
            return it_

        class Group:
            """the integrators of one step (one per slice), driven like a single one"""

            def __init__(g):
                g.probs = make_problems()
                g.its = [make_integrator(p) for p in g.probs]

            def solve(g):
                g.device_ms, g.walk_steps, g.launches, g.kern_n, g.kern_ms = 0.0, 0, 0, 0, 0.0
                for i in g.its:
                    if guided:
                        i.resetTraining()  # like elaina-exec: every solve starts from a freshly initialised network (exec.cu:104-123)
                    i.solve()
                    g.device_ms += i.last_device_ms
                    g.walk_steps += i.last_walk_steps
                    g.launches += i.launch_count
                    n_, ms_ = i.step_kernel_time
                    g.kern_n += n_
                    g.kern_ms += ms_

            def close(g):
                for i in g.its:
                    i.close()
                for p_ in g.probs:
                    p_.close()

        grp = Group()
        it = grp.its[0]
        for _ in range(warmup):
            grp.solve()
        if sampler is not None and rank == 0:
            sampler.start()
        self.barrier()
        dev_ms, steps_total, launches, kern_ms, kern_launches = 0.0, 0, 0, 0.0, 0
        for _ in range(steps):
            self.flush.zero_()
            torch.cuda.synchronize()
            grp.solve()
            dev_ms += grp.device_ms
            steps_total += grp.walk_steps
            launches += grp.launches
            kern_launches += grp.kern_n
            kern_ms += grp.kern_ms
        self.barrier()
        dev_ms = self.reduce(dev_ms, "max")
        kern_ms_max = self.reduce(kern_ms, "max")
        steps_all = self.reduce(float(steps_total), "sum")
        value = steps_all / (dev_ms * 1e-3)
        relmse = None
        if world == 1 and rank == 0:
            relmse = relmse_analytic(wl, it.film())
        l2_window = it.l2_window_bytes
        stats_ = [i.traversal_stats(reset=True) for i in grp.its]  # since creation: warm-up + timed solves
        trav_full, trav_restricted = sum(x[0] for x in stats_), sum(x[1] for x in stats_)
        train_stats = dict(it.train_stats) if guided else None
        profile = it.profile if guided and os.environ.get("WOST_PROFILE") else None

        # ---- end to end through the C ABI from host buffers
        h2d = wl["dverts"].nbytes + wl["dprims"].nbytes + (wl["color_dirichlet"].nbytes if wl.get("color_dirichlet") is not None else 0)
        if wl.get("nverts") is not None:
            h2d += wl["nverts"].nbytes + wl["nprims"].nbytes + (wl["color_neumann"].nbytes if wl.get("color_neumann") is not None else 0)
        if wl.get("source") is not None:
            h2d += wl["source"].nbytes
        d2h = W * H * 5 * 4 * len(slices)

        def e2e_once():
            t_ = [time.perf_counter()]
            g2 = Group()  # scenes (H2D + LBVH build), problems (H2D colours), integrators
            t_.append(time.perf_counter())
            g2.solve()
            t_.append(time.perf_counter())
            out = None
            for i2 in g2.its:
                if self.comm is not None:
                    self.comm.gather_film(i2, 0, 0)  # final grid gather: owned row tiles -> rank 0 (packed rows over NVLink)
                if rank == 0:
                    if (H, W) not in self.pinned:
                        self.pinned[(H, W)] = torch.empty((H, W, 5), dtype=torch.float32, pin_memory=True).numpy()
                    out = i2.film(out=self.pinned[(H, W)])  # D2H of the film into pinned host memory
            t_.append(time.perf_counter())
            if os.environ.get("WOST_BENCH_DEBUG"):
                print("e2e stages ms: scenes+problems+integrators %.1f solve %.1f (device %.1f) gather+download %.1f" % (
                    1e3 * (t_[1] - t_[0]), 1e3 * (t_[2] - t_[1]), g2.device_ms, 1e3 * (t_[3] - t_[2])), file=sys.stderr, flush=True)
            s = g2.walk_steps
            g2.close()
            return s, out

        n_e2e = max(1, min(steps, 3))
        e2e_value, e2e_relmse = None, None
        if do_e2e:
            e2e_once()
            self.barrier()
            e2e_steps = 0
            t0 = time.perf_counter()
            for _ in range(n_e2e):
                s, film = e2e_once()
                e2e_steps += s
            self.barrier()
            e2e_t = self.reduce(time.perf_counter() - t0, "max")
            e2e_value = self.reduce(float(e2e_steps), "sum") / e2e_t
            if rank == 0 and film is not None:
                e2e_relmse = relmse_analytic(wl, film)  # N > 1: checks the gathered film
        grp.close()
        if rank != 0:
            return None
        peak, peak_src = measured_peak()
        bps = wl["bytes_per_step"] + (176 if guided else 0)  # + 12 B network input and 164 B output per guided step
        achieved = steps_total * bps / (kern_ms * 1e-3) / 1e9 if kern_ms > 0 else 0.0
        kernel = ("k_g_arrive+k_g_bounce+k_g_finish<%d>" if guided else ("k_walk_persistent<%d>" if walk_mode.startswith("persistent") else "k_step_uniform<%d>")) % dim
        roof = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
                "kernel": kernel, "bytes_per_walk_step": bps, "launches": int(kern_launches), "avg_launch_ms": kern_ms / max(1, kern_launches),
                "kernel_share_of_step": kern_ms_max / dev_ms if dev_ms > 0 else None, "peak_source": peak_src}
        ip = issue_profile(wl, guided, walk_mode)
        if ip is not None:
            roof["traffic"] = ip["dram_bytes_per_walk_step"] * steps_total / max(1, kern_launches)
            roof["traffic_source"] = "NOT measured in this run: DRAM bytes per walk-step of the committed ncu capture (%s) x this run's walk-steps per launch" % ip.get("source", "profiles/")
            roof["bound_actual"] = ip.get("bound_actual", "issue")
            roof["issue"] = {k: ip.get(k) for k in ("issue_active_pct", "inst_per_cycle_per_sm", "threads_per_instruction", "warp_instructions_per_walk_step",
                                                    "l2_hit_rate_pct", "l2_gbs_achieved", "registers_per_thread", "ctas_per_sm") if k in ip}
            roof["issue"]["source"] = ip.get("source")
            # L2-resident tree: the HBM figure above is nominal. L2 bytes per walk-step come from the committed capture (lts sectors x
            # 32 B), the rate from THIS run's kernel time; the denominators are measured by tools/l2_peak.cu (profiles/l2_peak.json):
            # per-lane random 16-byte loads (the shape of the node fetches: every lane in another record) and fully coalesced loads
            try:
                lp = json.load(open(os.path.join(ROOT, "profiles", "l2_peak.json")))
                l2_bytes_per_step = ip["l2_gbs_achieved"] * 1e9 * ip["kernel_ms_under_ncu"] * 1e-3 / ip["walk_steps_in_launch"]
                l2_gbs = l2_bytes_per_step * steps_total / (kern_ms * 1e-3) / 1e9
                roof["l2"] = {"achieved_gbs": l2_gbs, "l2_bytes_per_walk_step": l2_bytes_per_step,
                              "peak_gbs": lp["random16"]["sector_gbs_best"], "frac": l2_gbs / lp["random16"]["sector_gbs_best"],
                              "peak_coalesced_gbs": lp["warp512"]["sector_gbs_best"], "frac_of_coalesced": l2_gbs / lp["warp512"]["sector_gbs_best"],
                              "source": "bytes per walk-step: committed ncu capture (not this run); kernel time: this run; peaks: profiles/l2_peak.json "
                                        "(tools/l2_peak.cu on a B200 of this pool, %d MiB buffer, L1 bypassed)" % lp.get("buffer_mib", 32)}
            except Exception:
                pass
        rec = {"value": value, "unit": UNIT, "ms_per_step": dev_ms / steps, "steps": steps, "warmup": warmup,
               "config": {"workload": wl["desc"], "integrator": integrator,
                          "walk_mode": walk_mode if not guided else ("fused tcgen05 bounce kernel" if mlp_impl == 0 else "validation: SIMT MLP through HBM"),
                          "relmse_vs_analytic": relmse, "spp_batch": spp_batch or "auto (~64M walks in flight uniform, ~16M guided)",
                          "walk_steps_per_step": steps_all / steps, "sharding": "interleaved 4-row tiles over %d ranks" % world,
                          "l2": "256 MiB flush between timed steps", "l2_window_bytes": int(l2_window),
                          "closest_point_second_passes_per_walk_step": trav_restricted / max(1.0, float(steps_total) * (steps + warmup) / steps),
                          "closest_point_full_reference_order_traversals": int(trav_full)},
               "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h), "steps": n_e2e,
                       "relmse_vs_analytic": e2e_relmse,
                       "path": "wost_scene_create+LBVH build -> wost_problem_create -> wost_integrator_create -> wost_solve -> %swost_film_download"
                               % ("wost_film_gather (NCCL, packed rows to rank 0) -> " if world > 1 else "")},
               "gpu_launches": int(launches), "roofline": roof}
        if guided:
            rec["guided"] = dict(train_stats, train_spp=gkw["train_spp_count"], network="DenseGrid + MLP 32-64-64-64-48 (data/*/n.json)",
                                 gradient_allreduce=("ncclAllReduce inside libwost.so, once per optimiser step" if world > 1 else None))
            if profile:
                rec["profile_ms_last_solve"] = profile
        return rec

    def parity_check(self, wl, band):
        """the oracle's rows (cpu_baseline leg) against the GPU film of the same rows: same Philox streams, 1 spp -> bit-identical"""
        api = self.api
        r0, r1, field = band[:3]
        W, H = wl["frame"]
        dim = wl["dim"]
        g_d = api.Scene(dim, wl["dverts"], wl["dprims"])
        g_n = api.Scene(dim, wl["nverts"], wl["nprims"]) if wl.get("nverts") is not None else None
        prob = api.Problem(dim, dirichlet=g_d, neumann=g_n, **problem_kwargs(wl))
        st = api.make_settings((W, H), wl["depth"], 1, wl["eps"], seed=wl["seed"], spp_batch=1, pixel_begin=r0 * W, pixel_end=r1 * W)
        it = api.UniformIntegrator(prob, st)
        it.solve()
        got = it.film()[r0:r1, :, :3].reshape(-1, 3)
        m = wl["mask"][r0:r1].reshape(-1).astype(bool) if wl.get("mask") is not None else np.ones(len(got), bool)
        same = np.array_equal(got[m].view(np.uint32), field[m].view(np.uint32))
        out = {"rows": [int(r0), int(r1)], "pixels": int(m.sum()), "bit_identical": bool(same),
               "max_abs_diff": float(np.max(np.abs(got[m] - field[m]))) if m.any() else 0.0,
               "what": "GPU film (wost_solve, 1 spp, spp_batch 1) == CPU oracle field on the same rows of the BASELINE-size frame"}
        it.close()
        prob.close()
        return out


    def parity_check_multi(self, wl, spp=8, spp_batch=4, band_rows=64):
        """N > 1: the film the N ranks solve in interleaved row tiles and gather on rank 0, against a band of
        rows rank 0 then solves ALONE (one shard, pixel range) at the same fixed spp_batch: pixel-keyed Philox streams + slot-ordered
        accumulators make the two bit-identical. Collective: every rank calls it."""
        api, wdist = self.api, self.wdist
        W, H = wl["frame"]
        dim = wl["dim"]

        def problem():
            g_d = api.Scene(dim, wl["dverts"], wl["dprims"])
            g_n = api.Scene(dim, wl["nverts"], wl["nprims"]) if wl.get("nverts") is not None else None
            return api.Problem(dim, dirichlet=g_d, neumann=g_n, **problem_kwargs(wl))
        prob = problem()
        st = api.make_settings((W, H), wl["depth"], spp, wl["eps"], seed=wl["seed"], spp_batch=spp_batch, **wdist.shard_kwargs(self.rank, self.world, stride_rows=4))
        it = api.UniformIntegrator(prob, st)
        it.solve()
        self.comm.gather_film(it, 0, 0)
        gathered = it.film()[..., :3].copy() if self.rank == 0 else None
        it.close()
        self.barrier()
        out = None
        if self.rank == 0:
            r0 = (H - band_rows) // 2
            r1 = r0 + band_rows
            st1 = api.make_settings((W, H), wl["depth"], spp, wl["eps"], seed=wl["seed"], spp_batch=spp_batch, pixel_begin=r0 * W, pixel_end=r1 * W)
            it1 = api.UniformIntegrator(prob, st1)
            it1.solve()
            alone = it1.film()[r0:r1, :, :3].reshape(-1, 3)
            it1.close()
            got = gathered[r0:r1].reshape(-1, 3)
            m = wl["mask"][r0:r1].reshape(-1).astype(bool) if wl.get("mask") is not None else np.ones(len(got), bool)
            out = {"ranks": self.world, "rows": [int(r0), int(r1)], "pixels": int(m.sum()), "spp": spp, "spp_batch": spp_batch,
                   "bit_identical": bool(np.array_equal(got[m].view(np.uint32), alone[m].view(np.uint32))),
                   "max_abs_diff": float(np.max(np.abs(got[m] - alone[m]))) if m.any() else 0.0,
                   "nonzero_pixels": int((np.abs(got[m]).sum(1) > 0).sum()),
                   "what": "film solved by %d ranks (interleaved 4-row tiles, wost_film_gather to rank 0) == the same rows solved by rank 0 alone, "
                           "same seed and spp_batch" % self.world}
        prob.close()
        self.barrier()
        return out


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
    ap.add_argument("--no-sub", action="store_true", help="skip the sub-workloads of the default run")
    ap.add_argument("--no-reference-cuda", action="store_true", help="--impl reference: do not run the reference's own CUDA binary")
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
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA GPU (there is no CPU fallback for the walker)")
    torch.cuda.set_device(local_rank)
    api.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    B = Bench(args, rank, local_rank, world)

    sampler = ClockSampler(local_rank)
    head = B.run(wl, args.integrator, args.steps, args.warmup, do_e2e=not args.no_e2e, walk_mode=args.walk_mode, spp_batch=args.spp_batch,
                 mlp_impl=args.mlp_impl, sampler=sampler)
    clocks = sampler.stop() if rank == 0 else None

    # ---- the other single-GPU configurations (default run only)
    sub = []
    default_run = args.workload == "c3" and args.integrator == "uniform" and args.spp == 0 and args.walk_mode == "persistent" and not args.no_sub
    if default_run:
        for name, integ, spp in (("c3", "guided", 0), ("c1", "uniform", 256), ("c4-small", "guided", 0)):
            w2 = workload(name)
            if spp:
                w2["desc"] = w2["desc"].replace("%d spp" % w2["spp"], "%d spp (reduced from %d for the bench)" % (spp, w2["spp"]))
                w2["spp"] = spp
            r = B.run(w2, integ, steps=2, warmup=1, do_e2e=not args.no_e2e)
            if rank == 0:
                r["name"] = "%s-%s" % (name, integ)
                sub.append(r)

    cpu, parity = None, None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:  # reported baseline: rank 0 at N = 1 only
        v, cores, sample, _, _, band = cpu_oracle_sample(wl, keep_field=True)
        cpu = {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample}
        if band is not None:
            parity = B.parity_check(wl, band)

    if world > 1 and default_run:
        parity = B.parity_check_multi(wl)

    if rank == 0:
        line = {"metric": METRIC, "value": head["value"], "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": head["ms_per_step"], "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
                "data": "synthetic", "config": head["config"], "e2e": head["e2e"], "gpu_launches": head["gpu_launches"], "roofline": head["roofline"],
                "clocks": clocks}
        for k in ("guided", "profile_ms_last_solve"):
            if k in head:
                line[k] = head[k]
        if sub:
            line["sub"] = sub
        if cpu is not None:
            line["cpu_baseline"] = cpu
        if parity is not None:
            line["parity_check"] = parity
        if world == 1 and default_run:
            ratio = vs_reference_cuda(head["value"], sub)
            if ratio is not None:
                line["vs_reference_cuda"] = ratio
        print(json.dumps(line), flush=True)
    if world > 1:
        B.barrier()
        if B.comm is not None:
            B.comm.close()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
