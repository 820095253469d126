"""CPU, world size 2, gloo: the host-side multi-GPU logic (elaina_b200/dist.py) -- shard rule, film reduction of
disjoint tiles == the single-process film bit for bit, gradient-slab reduction with valid-rank count.
The walks themselves come from the oracle (test infrastructure); no CUDA is touched."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from elaina_b200 import dist as wd
        from elaina_b200 import scenes
        from oracle_lib import Oracle
        orc = Oracle()
        dv, dp = scenes.star_polygon(300)
        prob = dict(dim=2, dirichlet=orc.scene(2, dv, dp), neumann=None, color_dirichlet=scenes.gray_pairs(scenes.harmonic_2d(dv, 3)),
                    grid_scale=1.2, grid_pos=(0.0, 0.0), grid_up=(0.0, 1.0))
        W, H, stride = 24, 20, 4
        st = dict(frame_w=W, frame_h=H, max_walking_depth=16, samples_per_pixel=3, epsilon_shell=3e-3, seed=5)
        film = np.zeros((W * H, 3), np.float32)
        steps = 0
        for b, e in wd.owned_pixel_ranges(W, H, stride, world, rank):
            sol, s = orc.solve_uniform(prob, st, pixel_begin=b, pixel_end=e, nthreads=1)
            film[b:e] = sol[b:e]
            steps += s
        film = wd.allreduce_film(film)
        t = torch.tensor([steps], dtype=torch.int64)
        dist.all_reduce(t)
        # gradient slab: rank 0 has a valid batch, rank 1 does not (contributes zeros and valid = 0)
        slab = torch.zeros(12)
        if rank == 0:
            slab[:8] = torch.arange(8, dtype=torch.float32)
            slab[8] = 1.0
        wd.allreduce_gradient_slab(slab)
        if rank == 0:
            full, full_steps = orc.solve_uniform(prob, st, nthreads=1)
            q.put(dict(ok=bool(np.array_equal(film.view(np.uint32), full.view(np.uint32))), steps=(int(t.item()), int(full_steps)),
                       slab=slab.tolist(), rows=[wd.owned_rows(H, stride, world, r).tolist() for r in range(world)]))
    finally:
        dist.destroy_process_group()


def test_two_rank_sharding_and_reductions_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = q.get(timeout=240)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res["ok"], "sum of the ranks' tiles differs from the single-process film"
    assert res["steps"][0] == res["steps"][1]
    rows = np.array(res["rows"])
    assert np.all(rows.sum(axis=0) == 1)  # every row owned exactly once
    assert res["slab"][:8] == list(map(float, range(8))) and res["slab"][8] == 1.0


def test_shard_rule_covers_every_row_once():
    sys.path.insert(0, ROOT)
    from elaina_b200 import dist as wd
    for world in (1, 2, 4, 8):
        for H in (1, 7, 64, 1023):
            own = np.array([wd.owned_rows(H, 4, world, r) for r in range(world)])
            assert np.all(own.sum(axis=0) == 1)
            total = sum(e - b for r in range(world) for b, e in wd.owned_pixel_ranges(5, H, 4, world, r))
            assert total == 5 * H
    assert wd.shard_kwargs(0, 1) == {}
    assert wd.shard_kwargs(3, 8) == dict(pixel_stride_rows=4, num_shards=8, shard=3)
