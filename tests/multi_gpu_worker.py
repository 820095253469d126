"""Worker of tests/test_gpu_multi.py: one process per GPU (torchrun), NCCL inside libwost.so (csrc/comm.cu).
Writes <out>/rank<r>.npz with what the test compares against the single-GPU results."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def cases():
    """the shared problem definitions (also imported by the test for the single-GPU side)"""
    from elaina_b200 import scenes
    from helpers import mixed_problem_2d
    dv, dp = scenes.star_polygon(600)
    uni = dict(dim=2, dverts=dv, dprims=dp, color_dirichlet=scenes.gray_pairs(scenes.harmonic_2d(dv, 3)), grid_scale=1.2, grid_pos=(0.0, 0.0), grid_up=(0.0, 1.0),
               aabb_min=(-1.5, -1.5), aabb_max=(1.5, 1.5))
    return uni, mixed_problem_2d(200)


UNI_SET = dict(frame_size=(64, 50), max_walking_depth=32, samples_per_pixel=6, epsilon_shell=3e-3, seed=9, spp_batch=1)  # 50 rows: last tile is partial
GUIDED_SET = dict(frame_size=(48, 48), max_walking_depth=16, samples_per_pixel=8, epsilon_shell=5e-3, seed=4, train_spp_count=4, max_guided_depth_training=4,
                  max_guided_depth_guiding=4, uniform_fraction_training=0.5, uniform_fraction_guiding=0.5, train_min_batch=128, train_batch_size=1024)
NET = {"encoding": {"otype": "DenseGrid", "n_levels": 4, "n_features_per_level": 4, "base_resolution": 4, "per_level_scale": 1.405}}


def net_cfg():
    from elaina_b200 import api
    return dict(NET, network=api.DEFAULT_NETWORK["network"], optimizer=api.DEFAULT_NETWORK["optimizer"])


def problem_of(api, spec):
    g_d = api.Scene(spec["dim"], spec["dverts"], spec["dprims"])
    g_n = api.Scene(spec["dim"], spec["nverts"], spec["nprims"]) if spec.get("nverts") is not None else None
    kw = {k: spec[k] for k in ("color_dirichlet", "color_neumann", "grid_scale", "grid_pos", "grid_up", "aabb_min", "aabb_max") if k in spec}
    return api.Problem(spec["dim"], dirichlet=g_d, neumann=g_n, **kw)


def train_samples(n=128 * 16, seed=21):
    rng = np.random.RandomState(seed)
    pos = (-1.4 + rng.rand(n, 2) * 2.8).astype(np.float32)

    def dirs():
        d = rng.randn(n, 2).astype(np.float32)
        return d / np.linalg.norm(d, axis=1, keepdims=True)
    return pos, dirs(), dirs(), (0.05 + rng.rand(n) * 0.4).astype(np.float32), (rng.rand(n) * 2.0).astype(np.float32), (rng.rand(n) < 0.3).astype(np.uint8)


def init_params(n):
    return (np.random.RandomState(1).randn(n) * 0.2).astype(np.float32)


def main():
    out_dir = sys.argv[1]
    rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
    import torch
    import torch.distributed as dist
    from elaina_b200 import api
    from elaina_b200 import dist as wd
    torch.cuda.set_device(local)
    api.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    comm = wd.Comm.from_torch(torch.device("cuda", local))
    res = {"nccl_version": np.int64(comm.nccl_version)}
    uni, mixed = cases()
    # (1) uniform: shard by interleaved row tiles, gather to rank 0, then all-gather
    prob = problem_of(api, uni)
    it = api.UniformIntegrator(prob, api.make_settings(**UNI_SET, **comm.shard_kwargs(4)))
    it.solve()
    res["uni_steps"] = comm.allreduce_scalars([it.last_walk_steps], "sum")
    res["uni_own"] = it.film()
    comm.gather_film(it, 0, 0)
    res["uni_root_gather"] = it.film()
    it.solve()
    comm.gather_film(it, 0, -1)
    res["uni_all_gather"] = it.film()
    # (2) one optimiser step on explicit samples: every rank takes its 1/N of the batch, gradient all-reduced inside the library
    gp = problem_of(api, mixed)
    gi = api.GuidedIntegrator(gp, api.make_settings(**GUIDED_SET, **comm.shard_kwargs(4)), net_cfg())
    comm.attach(gi)
    gi.set_params(init_params(gi.n_params))
    s = train_samples()
    n = len(s[0]) // world
    loss, grad = gi.train_batch(*[a[rank * n:(rank + 1) * n] for a in s])
    res["step_loss"], res["step_grad"], res["step_params"] = np.float32(loss), grad, gi.get_params()
    # (3) a whole guided solve: replicas must stay identical, film gathered
    gi.resetNetwork(net_cfg())
    comm.attach(gi)
    gi.solve()
    res["solve_params"] = gi.get_params()
    res["solve_stats"] = np.array([gi.train_stats["steps"], gi.train_stats["samples"]], np.int64)
    comm.gather_film(gi, 0, 0)
    res["solve_film"] = gi.film()
    np.savez(os.path.join(out_dir, "rank%d.npz" % rank), **res)
    dist.barrier()
    comm.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
