"""Multi-GPU plumbing for the walker: one process per GPU, torch.distributed for the collectives.

Walks shard by evaluation point: rank r owns the rows y with (y // stride_rows) % world == r
(`wost_settings.pixel_stride_rows / num_shards / shard`, applied inside wost_integrator_create). Geometry, boundary
data, source grid and guiding network are replicated. The Philox stream is keyed by (seed, pixel, sample), so the
union of the ranks' films equals the single-GPU film bit for bit. Collectives:
  * final film: one all-reduce(sum) of the ranks' films -- tiles are disjoint and every rank writes zeros elsewhere;
  * guided training: one all-reduce(sum) of the fp32 gradient slab per batch, issued on the integrator's own CUDA
    stream between the backward kernel and the optimiser kernel (wost_set_allreduce). The slab's last words carry the
    number of ranks whose batch was valid; the optimiser divides by it, so a rank with too few samples contributes 0.
The reference has no multi-GPU path at all (core/device/context.cpp:23-24 hard-codes device 0).
"""
import os

import numpy as np


def env_rank_world():
    return int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))


def shard_kwargs(rank, world, stride_rows=4):
    """keyword arguments for api.make_settings selecting this rank's interleaved row tiles"""
    if world <= 1:
        return {}
    return dict(pixel_stride_rows=int(stride_rows), num_shards=int(world), shard=int(rank))


def owned_rows(height, stride_rows, world, rank):
    """bool[height]: the rows rank `rank` evaluates -- the rule wost_integrator_create applies (integrator.cu)"""
    y = np.arange(height)
    if world <= 1 or stride_rows <= 0:
        return np.ones(height, bool)
    return (y // stride_rows) % world == rank


def owned_pixel_ranges(width, height, stride_rows, world, rank):
    """contiguous [begin, end) pixel ranges of this rank's tiles (for callers that take pixel ranges, e.g. the oracle)"""
    rows = owned_rows(height, stride_rows, world, rank)
    out, y = [], 0
    while y < height:
        if rows[y]:
            y0 = y
            while y < height and rows[y]:
                y += 1
            out.append((y0 * width, y * width))
        else:
            y += 1
    return out


def wrap_device_ptr(ptr, n_floats, device):
    """float32 torch view over device memory owned by libwost.so"""
    import torch

    class _Holder:
        pass
    h = _Holder()
    h.__cuda_array_interface__ = {"shape": (int(n_floats),), "typestr": "<f4", "data": (int(ptr), False), "version": 2}
    return torch.as_tensor(h, device=device)


def allreduce_film(film, group=None):
    """sum the ranks' films (torch tensor on any device, or numpy array through a CPU tensor); returns the same type"""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return film
    if isinstance(film, np.ndarray):
        t = torch.from_numpy(np.ascontiguousarray(film))
        dist.all_reduce(t, group=group)
        return t.numpy()
    dist.all_reduce(film, group=group)
    return film


def allreduce_gradient_slab(slab, group=None):
    """SUM all-reduce of a gradient slab [params..., valid_count, pad...]; the optimiser divides by slab[n_params]."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(slab, group=group)
    return slab


def install_gradient_allreduce(integrator, device, group=None):
    """register the NCCL all-reduce of the guiding-network gradients on a GuidedIntegrator (no-op for world size 1)"""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return False

    def fn(ptr, n, stream):
        t = wrap_device_ptr(ptr, n, device)
        with torch.cuda.stream(torch.cuda.ExternalStream(int(stream), device=device)):
            dist.all_reduce(t, group=group)

    integrator.set_allreduce(fn)
    return True
