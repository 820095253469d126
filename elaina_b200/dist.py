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

Two routes to the same collectives:
  * `Comm` (default in bench.py): NCCL inside libwost.so (csrc/comm.cu, include/wost.h wost_comm_*): the gradient all-reduce
    is an ncclAllReduce the library enqueues on the integrator's own stream -- no Python in the per-step path -- and the film
    is GATHERED (packed owned rows, grouped send/recv to the root or one all-gather), not all-reduced. elaina-exec uses the
    same entry points under torchrun / mpirun without any Python. torch.distributed only carries the 128-byte NCCL id and
    the timing barriers.
  * `install_gradient_allreduce` / `allreduce_film`: the round-1 route through torch.distributed (kept for hosts that
    already own a process group, and as the cross-check of the native route in tests/test_gpu_multi.py).
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


class Comm:
    """NCCL communicator owned by libwost.so (include/wost.h: wost_comm_*). One per process, one process per GPU."""

    def __init__(self, rank, world, unique_id):
        import ctypes
        from . import _lib
        self._lib = _lib
        self.rank, self.world = int(rank), int(world)
        self.h = ctypes.c_void_p()
        buf = (ctypes.c_ubyte * 128).from_buffer_copy(bytes(unique_id))
        _lib.check(_lib.load().wost_comm_create(self.rank, self.world, buf, ctypes.byref(self.h)))

    @staticmethod
    def unique_id():
        import ctypes
        from . import _lib
        buf = (ctypes.c_ubyte * 128)()
        _lib.check(_lib.load().wost_comm_unique_id(buf))
        return bytes(buf)

    @classmethod
    def from_torch(cls, device, group=None):
        """share rank 0's NCCL id through an initialised torch.distributed group (any backend) -> Comm, or None for world size 1"""
        import torch
        import torch.distributed as dist
        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
            return None
        rank, world = dist.get_rank(group), dist.get_world_size(group)
        dev = device if dist.get_backend(group) == "nccl" else torch.device("cpu")
        t = torch.zeros(128, dtype=torch.uint8, device=dev)
        if rank == 0:
            t.copy_(torch.frombuffer(bytearray(cls.unique_id()), dtype=torch.uint8))
        dist.broadcast(t, src=0, group=group)
        return cls(rank, world, bytes(t.cpu().numpy().tobytes()))

    @property
    def nccl_version(self):
        import ctypes
        v = ctypes.c_int()
        self._lib.check(self._lib.load().wost_comm_info(self.h, None, None, ctypes.byref(v)))
        return v.value

    def shard_kwargs(self, stride_rows=4):
        return shard_kwargs(self.rank, self.world, stride_rows)

    def attach(self, integrator):
        """guided integrator: the library all-reduces the gradient slab itself on the integrator's stream"""
        self._lib.check(self._lib.load().wost_integrator_set_comm(integrator.h, self.h))

    def gather_film(self, integrator, channel=0, root=0):
        """final grid gather: rank `root` (every rank if root < 0) ends up with the complete device film"""
        self._lib.check(self._lib.load().wost_film_gather(integrator.h, self.h, int(channel), int(root)))

    def allreduce_scalars(self, vals, op="sum"):
        import ctypes
        a = np.ascontiguousarray(vals, np.float64).copy()
        self._lib.check(self._lib.load().wost_comm_allreduce_f64(self.h, a.ctypes.data_as(ctypes.c_void_p), int(a.size), 1 if op == "max" else 0))
        return a

    def close(self):
        if self.h:
            self._lib.load().wost_comm_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
