#!/usr/bin/env python
"""elaina-exec on N GPUs without Python in the loop: the same conf.json run by one process and by `torchrun --no-python
elaina_b200/elaina-exec conf.json` (one process per GPU, NCCL inside libwost.so). The uniform film must be bit-identical;
the guided film is compared statistically (training batches are composed differently). Prints one JSON line.
  python tools/exec_multi_check.py [--gpus 2] [--workload c3-small] [--spp 16]"""
import argparse
import json
import os
import shutil
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from elaina_b200 import workloads  # noqa: E402

BIN = os.path.join(ROOT, "elaina_b200", "elaina-exec")


def run(cmd, cwd):
    t0 = time.perf_counter()
    p = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=900, cwd=cwd)
    return p.returncode, time.perf_counter() - t0, p.stdout[-1500:]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=2)
    ap.add_argument("--workload", default="c3-small")
    ap.add_argument("--spp", type=int, default=16)
    a = ap.parse_args()
    wl = workloads.workload(a.workload)
    W, H = wl["frame"]
    out = {"workload": a.workload, "gpus": a.gpus, "spp": a.spp}
    for integ in ("uniform", "guided"):
        case = tempfile.mkdtemp(prefix="wost_exec_multi_")
        films, res = {}, {}
        for tag, n in (("single", 1), ("multi", a.gpus)):
            conf = workloads.write_case(wl, case, integrator=integ, exp_name=tag, spp=a.spp)
            cmd = [BIN, conf] if n == 1 else [sys.executable, "-m", "torch.distributed.run", "--no-python", "--nnodes=1", "--nproc-per-node", str(n),
                                             "--master-addr", "127.0.0.1", "--master-port", str(29700 + os.getpid() % 200), BIN, conf]
            rc, wall, tail = run(cmd, case)
            r = {"rc": rc, "wall_s": wall}
            if rc != 0:
                r["log_tail"] = tail
            try:
                r.update(json.load(open(os.path.join(case, tag, "result.json"))))
                films[tag] = np.fromfile(os.path.join(case, tag, "solution.f32"), np.float32).reshape(H, W, 5)
            except Exception as e:  # noqa: BLE001
                r["error"] = repr(e)
            res[tag] = r
        if len(films) == 2:
            a_, b_ = films["single"], films["multi"]
            res["bit_identical"] = bool(np.array_equal(a_.view(np.uint32), b_.view(np.uint32)))
            m = wl["mask"].astype(bool) if wl.get("mask") is not None else np.ones((H, W), bool)
            res["mean_abs_diff"] = float(np.mean(np.abs(a_[m][:, :3] - b_[m][:, :3])))
            res["walk_steps_equal"] = res["single"].get("walk_steps") == res["multi"].get("walk_steps")
        out[integ] = res
        shutil.rmtree(case, ignore_errors=True)
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
