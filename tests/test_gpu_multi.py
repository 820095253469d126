"""GPU, 2 ranks (one process per GPU, torchrun): the NCCL path inside libwost.so -- film gather == the single-GPU film bit for
bit; a data-parallel optimiser step == the single-GPU step on the whole batch within fp32 reduction noise; replicas stay
identical through a guided solve. Skipped on boxes with fewer than two GPUs (run with `gpurun --gpus 2`)."""
import os
import subprocess
import sys
import tempfile

import numpy as np
import pytest

from helpers import bits

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ngpus():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


@pytest.fixture(scope="module")
def two_rank_results():
    if _ngpus() < 2:
        pytest.skip("needs two GPUs")
    out = tempfile.mkdtemp(prefix="wost_multi_")
    port = 29600 + os.getpid() % 300
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1", "--master-port", str(port),
           os.path.join(ROOT, "tests", "multi_gpu_worker.py"), out]
    p = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600, cwd=ROOT)
    assert p.returncode == 0, p.stdout[-3000:]
    return [np.load(os.path.join(out, "rank%d.npz" % r)) for r in range(2)]


def test_film_gather_equals_single_gpu(two_rank_results):
    from elaina_b200 import api
    import multi_gpu_worker as w
    r0, r1 = two_rank_results
    uni, _ = w.cases()
    prob = w.problem_of(api, uni)
    it = api.UniformIntegrator(prob, api.make_settings(**w.UNI_SET))
    it.solve()
    full = it.film()
    assert int(r0["uni_steps"][0]) == it.last_walk_steps == int(r1["uni_steps"][0])
    assert np.array_equal(bits(r0["uni_root_gather"]), bits(full))          # gather to rank 0 (packed rows, send/recv)
    for r in (r0, r1):
        assert np.array_equal(bits(r["uni_all_gather"]), bits(full))        # all-gather variant: every rank holds the film
    # before the gather every rank only holds its own tiles (the others keep the initial {0,0,0,1,1})
    own0, own1 = r0["uni_own"], r1["uni_own"]
    rows0 = (np.arange(full.shape[0]) // 4) % 2 == 0
    assert np.array_equal(bits(own0[rows0]), bits(full[rows0])) and np.array_equal(bits(own1[~rows0]), bits(full[~rows0]))
    assert np.all(own0[~rows0][..., :3] == 0) and np.all(own1[rows0][..., :3] == 0)
    assert int(r0["nccl_version"]) >= 21800


def test_data_parallel_step_equals_single_gpu_step(two_rank_results):
    from elaina_b200 import api
    import multi_gpu_worker as w
    r0, r1 = two_rank_results
    _, mixed = w.cases()
    gp = w.problem_of(api, mixed)
    gi = api.GuidedIntegrator(gp, api.make_settings(**w.GUIDED_SET), w.net_cfg())
    p0 = w.init_params(gi.n_params)
    gi.set_params(p0)
    loss, grad = gi.train_batch(*w.train_samples())
    p1 = gi.get_params()
    # train_batch returns the rank's LOCAL gradient (mean over its half of the batch); the library all-reduces the slab and the
    # optimiser divides by the number of contributing ranks: both ranks take the same step
    assert not np.array_equal(bits(r0["step_grad"]), bits(r1["step_grad"]))
    assert np.array_equal(bits(r0["step_params"]), bits(r1["step_params"]))
    # mean of the two half-batch gradients == the whole-batch gradient up to fp32 summation order (fp16 operands are identical)
    nm = gi.n_mlp_params
    both = 0.5 * (r0["step_grad"].astype(np.float64) + r1["step_grad"].astype(np.float64))
    for lo, hi in ((0, nm), (nm, gi.n_params)):
        scale = np.abs(grad[lo:hi]).max() + 1e-12
        assert np.max(np.abs(both[lo:hi] - grad[lo:hi])) / scale < 1e-3, (lo, hi)
    moved = np.abs(p1 - p0) > 1e-4
    assert moved.sum() > 1000
    strong = np.abs(grad[moved]) > 1e-3 * np.abs(grad).max()
    np.testing.assert_allclose(r0["step_params"][moved][strong], p1[moved][strong], atol=5e-4)
    assert 0.5 * (float(r0["step_loss"]) + float(r1["step_loss"])) == pytest.approx(loss, rel=1e-3)


def test_guided_solve_keeps_replicas_identical(two_rank_results):
    r0, r1 = two_rank_results
    assert np.array_equal(bits(r0["solve_params"]), bits(r1["solve_params"]))
    assert int(r0["solve_stats"][0]) == int(r1["solve_stats"][0]) >= 1      # same number of optimiser steps (= all-reduces) on both ranks
    f = r0["solve_film"]
    assert np.isfinite(f).all() and np.abs(f[..., :3]).sum() > 0
    rows1 = (np.arange(f.shape[0]) // 4) % 2 == 1
    assert np.abs(f[rows1][..., :3]).sum() > 0                              # rank 1's tiles arrived on rank 0
