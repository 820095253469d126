"""Host-side mirror of the reference's walker interface, bound to the C ABI of include/wost.h.

Names and argument meaning follow the reference so that the parity tests read like its own code would:
  Scene               <-> lbvh::scene<DIM> + build_bvh()              (ext/lbvh/include/snch_lbvh/scene.cuh)
  Problem             <-> Problem<DIM>::loadConfig                     (core/problem.cu:151-213)
  UniformIntegrator   <-> UniformIntegrator<DIM>                       (integrator/uniform/integrator.h:54-131)
     .solve() / .renderDirichletSDF() / .renderSilhouetteSDF() / .renderSource() / .film(channel)
Every call goes through libwost.so; there is no Python or CPU implementation of the hot path.
"""
import ctypes
import json
import os

import numpy as np

from . import _lib
from ._lib import ProblemDesc, Settings, c_float, c_int, c_u32, c_u64, c_void_p, check
from . import scenes as _scenes

CH_SOLUTION, CH_DIRICHLET_SDF, CH_NEUMANN_SDF, CH_SOURCE = 0, 1, 2, 3
CHANNELS = {"SOLUTION": 0, "DIRICHLET_SDF": 1, "NEUMANN_SDF": 2, "SOURCE": 3}
UNIFORM, GUIDED = 0, 1


def _p(a):
    return a.ctypes.data_as(c_void_p) if a is not None else None


def set_device(device):
    check(_lib.load().wost_set_device(int(device)))


class Scene:
    """One mesh (line segments in 2D, triangles in 3D) with its SNCH-LBVH on the GPU."""

    def __init__(self, dim, verts, prims):
        L = _lib.load()
        self.dim = int(dim)
        self.verts = np.ascontiguousarray(verts, np.float32).reshape(-1, self.dim)
        self.prims = np.ascontiguousarray(prims, np.int32).reshape(-1, self.dim)
        h = c_void_p()
        check(L.wost_scene_create(self.dim, _p(self.verts), len(self.verts), _p(self.prims), len(self.prims), ctypes.byref(h)))
        self.h = h

    @classmethod
    def from_obj(cls, path, dim):
        v, p = _scenes.load_obj(path, dim)
        return cls(dim, v, p)

    def close(self):
        if getattr(self, "h", None):
            _lib.load().wost_scene_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def num_prims(self):
        return len(self.prims)

    @property
    def num_nodes(self):
        return 2 * len(self.prims) - 1

    @property
    def used_morton64(self):
        d, n, nn, u = c_int(), c_int(), c_int(), c_int()
        check(_lib.load().wost_scene_info(self.h, ctypes.byref(d), ctypes.byref(n), ctypes.byref(nn), ctypes.byref(u)))
        return bool(u.value)

    def dump_bvh(self):
        n, d = self.num_nodes, self.dim
        nodes = np.zeros((n, 4), np.uint32)
        aabbs = np.zeros((n, 2 * d), np.float32)
        cones = np.zeros((n, d + 2), np.float32)
        check(_lib.load().wost_scene_dump_bvh(self.h, _p(nodes), _p(aabbs), _p(cones)))
        return nodes, aabbs, cones

    def dump_sorted(self):
        m = np.zeros(self.num_prims, np.uint32)
        o = np.zeros(self.num_prims, np.uint32)
        check(_lib.load().wost_scene_dump_sorted(self.h, _p(m), _p(o)))
        return m, o

    # ---- queries (host buffers through the C ABI)
    def nearest(self, pts):
        pts = np.ascontiguousarray(pts, np.float32).reshape(-1, self.dim)
        idx = np.zeros(len(pts), np.uint32)
        dist = np.zeros(len(pts), np.float32)
        check(_lib.load().wost_lbvh_nearest(self.h, _p(pts), len(pts), _p(idx), _p(dist)))
        return idx, dist

    def silhouette(self, pts, flip=False):
        pts = np.ascontiguousarray(pts, np.float32).reshape(-1, self.dim)
        dist = np.zeros(len(pts), np.float32)
        check(_lib.load().wost_lbvh_silhouette(self.h, _p(pts), len(pts), int(flip), _p(dist)))
        return dist

    def ray(self, org, dr, max_dist):
        org = np.ascontiguousarray(org, np.float32).reshape(-1, self.dim)
        dr = np.ascontiguousarray(dr, np.float32).reshape(-1, self.dim)
        n = len(org)
        md = np.ascontiguousarray(np.broadcast_to(np.asarray(max_dist, np.float32), (n,)))
        hit = np.zeros(n, np.int32)
        t = np.zeros(n, np.float32)
        uv = np.zeros((n, 2), np.float32)
        obj = np.zeros(n, np.uint32)
        check(_lib.load().wost_lbvh_ray(self.h, _p(org), _p(dr), _p(md), n, _p(hit), _p(t), _p(uv), _p(obj)))
        return hit, t, uv, obj

    def ray_any(self, org, dr, max_dist):
        org = np.ascontiguousarray(org, np.float32).reshape(-1, self.dim)
        dr = np.ascontiguousarray(dr, np.float32).reshape(-1, self.dim)
        n = len(org)
        md = np.ascontiguousarray(np.broadcast_to(np.asarray(max_dist, np.float32), (n,)))
        hit = np.zeros(n, np.int32)
        check(_lib.load().wost_lbvh_ray_any(self.h, _p(org), _p(dr), _p(md), n, _p(hit)))
        return hit

    def sample_in_sphere(self, ctr, radius, u):
        ctr = np.ascontiguousarray(ctr, np.float32).reshape(-1, self.dim)
        n = len(ctr)
        radius = np.ascontiguousarray(np.broadcast_to(np.asarray(radius, np.float32), (n,)))
        u = np.ascontiguousarray(u, np.float32)
        obj = np.zeros(n, np.int32)
        pdf = np.zeros(n, np.float32)
        check(_lib.load().wost_lbvh_sample_sphere(self.h, _p(ctr), _p(radius), _p(u), n, _p(obj), _p(pdf)))
        return obj, pdf


class Problem:
    """Problem<DIM>: meshes + boundary data + source + evaluation grid (core/problem.h, core/problem.cu)."""

    def __init__(self, dim, dirichlet=None, neumann=None, color_dirichlet=None, color_neumann=None,
                 dirichlet_intensity=1.0, neumann_intensity=1.0, source_intensity=1.0,
                 source=None, source_origin=(0, 0, 0), source_voxel=1.0,
                 grid_scale=1.0, grid_pos=None, grid_up=None, grid_right=None, mask=None,
                 aabb_min=None, aabb_max=None):
        L = _lib.load()
        self.dim = dim
        self.dirichlet, self.neumann = dirichlet, neumann
        d = ProblemDesc()
        keep = []
        d.dim = dim
        d.dirichlet = dirichlet.h if dirichlet is not None else None
        d.neumann = neumann.h if neumann is not None else None
        if color_dirichlet is not None:
            a = np.ascontiguousarray(color_dirichlet, np.float32).reshape(-1, 6)
            keep.append(a)
            d.color_dirichlet, d.n_color_dirichlet = a.ctypes.data, len(a)
        if color_neumann is not None:
            a = np.ascontiguousarray(color_neumann, np.float32).reshape(-1, 6)
            keep.append(a)
            d.color_neumann, d.n_color_neumann = a.ctypes.data, len(a)
        d.dirichlet_intensity, d.neumann_intensity, d.source_intensity = dirichlet_intensity, neumann_intensity, source_intensity
        if source is not None:
            src = np.ascontiguousarray(source, np.float32)
            keep.append(src)
            nz, ny, nx = src.shape[:3]
            d.source = src.ctypes.data
            d.source_dims = (c_int * 3)(nx, ny, nz)
            so = list(source_origin) + [0.0, 0.0, 0.0]
            d.source_origin = (c_float * 3)(*so[:3])
            d.source_voxel = source_voxel

        def v3(x, default):
            x = list(x if x is not None else default) + [0.0, 0.0, 0.0]
            return (c_float * 3)(*[float(t) for t in x[:3]])

        d.grid_scale = grid_scale
        d.grid_pos = v3(grid_pos, (0, 0, 0))
        d.grid_up = v3(grid_up, (0, 1, 0) if dim == 2 else (0, 0, 1))
        d.grid_right = v3(grid_right, (1, 0, 0))
        if mask is not None:
            m = np.ascontiguousarray(mask, np.uint8)
            keep.append(m)
            d.mask = m.ctypes.data
            d.mask_h, d.mask_w = m.shape
        d.aabb_min = v3(aabb_min, (0, 0, 0))
        d.aabb_max = v3(aabb_max, (1, 1, 1))
        h = c_void_p()
        check(L.wost_problem_create(ctypes.byref(d), ctypes.byref(h)))
        self.h = h

    def close(self):
        if getattr(self, "h", None):
            _lib.load().wost_problem_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def make_settings(frame_size=(1024, 1024), max_walking_depth=64, samples_per_pixel=1, epsilon_shell=1.0, seed=0,
                  spp_batch=0, walk_mode=0, pixel_begin=0, pixel_end=0, pixel_stride_rows=0, num_shards=1, shard=0,
                  train_spp_count=0, max_guided_depth_training=0, max_guided_depth_guiding=0,
                  uniform_fraction_training=1.0, uniform_fraction_guiding=1.0, debug_pixel=0):
    s = Settings()
    s.frame_w, s.frame_h = int(frame_size[0]), int(frame_size[1])
    s.max_walking_depth, s.samples_per_pixel = int(max_walking_depth), int(samples_per_pixel)
    s.epsilon_shell = float(epsilon_shell)
    s.debug_pixel = debug_pixel
    s.save_spp_metrics_duration = s.save_spp_metrics_until = s.save_time_metrics_duration = -1
    s.train_spp_count = train_spp_count
    s.max_guided_depth_training, s.max_guided_depth_guiding = max_guided_depth_training, max_guided_depth_guiding
    s.uniform_fraction_training, s.uniform_fraction_guiding = uniform_fraction_training, uniform_fraction_guiding
    s.seed, s.walk_mode, s.spp_batch = int(seed), int(walk_mode), int(spp_batch)
    s.pixel_begin, s.pixel_end = int(pixel_begin), int(pixel_end)
    s.pixel_stride_rows, s.num_shards, s.shard = int(pixel_stride_rows), int(num_shards), int(shard)
    return s


class _Integrator:
    kind = UNIFORM

    def __init__(self, problem, settings):
        L = _lib.load()
        self.problem = problem
        self.settings = settings
        h = c_void_p()
        check(L.wost_integrator_create(problem.h, ctypes.byref(settings), self.kind, ctypes.byref(h)))
        self.h = h
        self.last_walk_steps = 0
        self.last_device_ms = 0.0

    def close(self):
        if getattr(self, "h", None):
            _lib.load().wost_integrator_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def solve(self):
        """-> wall-clock ms (like the reference's solve()); walk steps / device ms kept on the object."""
        ms, steps, dms = c_u64(), c_u64(), c_float()
        check(_lib.load().wost_solve(self.h, ctypes.byref(ms), ctypes.byref(steps), ctypes.byref(dms)))
        self.last_walk_steps, self.last_device_ms = steps.value, dms.value
        return ms.value

    def renderDirichletSDF(self):
        check(_lib.load().wost_render_channel(self.h, CH_DIRICHLET_SDF))

    def renderSilhouetteSDF(self):
        check(_lib.load().wost_render_channel(self.h, CH_NEUMANN_SDF))

    def renderSource(self):
        check(_lib.load().wost_render_channel(self.h, CH_SOURCE))

    def film(self, channel=CH_SOLUTION):
        """(H, W, 5) float32: r, g, b, a, weight -- util/film.h layout, index x + y*W, not flipped."""
        w, h = self.settings.frame_w, self.settings.frame_h
        out = np.zeros((h, w, 5), np.float32)
        check(_lib.load().wost_film_download(self.h, channel, _p(out)))
        return out

    def film_device_ptr(self, channel=CH_SOLUTION):
        p = c_void_p()
        check(_lib.load().wost_film_device_ptr(self.h, channel, ctypes.byref(p)))
        return p.value

    def solution(self):
        f = self.film(CH_SOLUTION)
        return f[..., :3] / np.maximum(f[..., 4:5], 1e-30)

    @property
    def launch_count(self):
        n = c_u64()
        check(_lib.load().wost_last_launch_count(self.h, ctypes.byref(n)))
        return n.value


    @property
    def step_kernel_time(self):
        """-> (launches of the fused walk-step kernel, their summed device ms) of the last solve"""
        n, ms = c_u64(), c_float()
        check(_lib.load().wost_last_step_kernel_time(self.h, ctypes.byref(n), ctypes.byref(ms)))
        return n.value, ms.value


class UniformIntegrator(_Integrator):
    kind = UNIFORM
