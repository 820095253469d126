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

    def dump_traversal_layout(self):
        """(records[num_prims-1, F], leaf_verts[num_prims, dim, dim]) of the B200 traversal layout (csrc/lbvh.cuh: TNode)"""
        L = _lib.load()
        nf = ctypes.c_int()
        check(L.wost_scene_dump_traversal_layout(self.h, None, ctypes.byref(nf), None))
        rec = np.zeros((max(0, self.num_prims - 1), nf.value), np.float32)
        lv = np.zeros((self.num_prims, self.dim, self.dim), np.float32)
        check(L.wost_scene_dump_traversal_layout(self.h, _p(rec), ctypes.byref(nf), _p(lv)))
        return rec, lv

    def dump_sorted(self):
        m = np.zeros(self.num_prims, np.uint32)
        o = np.zeros(self.num_prims, np.uint32)
        check(_lib.load().wost_scene_dump_sorted(self.h, _p(m), _p(o)))
        return m, o

    # ---- queries (host buffers through the C ABI)
    def traversal_stats(self, reset=True):
        """(full reference-order traversals, restricted second passes) of the closest-point probes on this scene"""
        a, b = ctypes.c_uint64(), ctypes.c_uint64()
        check(_lib.load().wost_scene_traversal_stats(self.h, ctypes.byref(a), ctypes.byref(b), 1 if reset else 0))
        return a.value, b.value

    def nearest(self, pts):
        pts = np.ascontiguousarray(pts, np.float32).reshape(-1, self.dim)
        idx = np.zeros(len(pts), np.uint32)
        dist = np.zeros(len(pts), np.float32)
        check(_lib.load().wost_lbvh_nearest(self.h, _p(pts), len(pts), _p(idx), _p(dist)))
        return idx, dist

    def nearest_static(self, pts):
        """wost_lbvh_nearest_static: same answer as nearest(), staged through the library's fixed device array (no allocation;
        at most MAX_STATIC_POINTS points per call, one call at a time per GPU)."""
        pts = np.ascontiguousarray(pts, np.float32).reshape(-1, self.dim)
        idx = np.zeros(len(pts), np.uint32)
        dist = np.zeros(len(pts), np.float32)
        check(_lib.load().wost_lbvh_nearest_static(self.h, _p(pts), len(pts), _p(idx), _p(dist)))
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
                  uniform_fraction_training=1.0, uniform_fraction_guiding=1.0, debug_pixel=0,
                  mlp_impl=0, max_train_depth=0, train_batch_size=0, train_min_batch=0, train_batches_per_frame=0):
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
    s.mlp_impl, s.max_train_depth = int(mlp_impl), int(max_train_depth)
    s.train_batch_size, s.train_min_batch, s.train_batches_per_frame = int(train_batch_size), int(train_min_batch), int(train_batches_per_frame)
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

    def set_frame_callback(self, fn):
        """Progressive metrics (uniform/integrator.cu:577-609): fn(kind, sample_id, elapsed_ms, film[H, W, 5]) is called during
        solve() for every sample index the settings' saveSppMetricsDuration / saveTimeMetricsDuration select (kind 0 / 1)."""
        if fn is None:
            self._frame_cb = None
            check(_lib.load().wost_set_frame_callback(self.h, None, None))
            return

        def tramp(user, kind, sample_id, ms, film, w, h):
            try:
                fn(kind, sample_id, ms, np.ctypeslib.as_array(film, shape=(h, w, 5)).copy())
            except Exception:
                import traceback
                traceback.print_exc()
        self._frame_cb = _lib.FRAME_FN(tramp)
        check(_lib.load().wost_set_frame_callback(self.h, ctypes.cast(self._frame_cb, c_void_p), None))

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

    def film(self, channel=CH_SOLUTION, out=None):
        """(H, W, 5) float32: r, g, b, a, weight -- util/film.h layout, index x + y*W, not flipped.
        out: optional preallocated C-contiguous float32 array of that shape (e.g. a view of pinned host memory) to download into."""
        w, h = self.settings.frame_w, self.settings.frame_h
        if out is None:
            out = np.empty((h, w, 5), np.float32)
        assert out.dtype == np.float32 and out.flags["C_CONTIGUOUS"] and out.size == h * w * 5
        check(_lib.load().wost_film_download(self.h, channel, _p(out)))
        return out.reshape(h, w, 5)

    def film_stderr(self):
        """(H, W, 3) measured standard error of the solution of the last solve (wost_film_stderr); exact per-pixel sample
        standard error when the integrator was created with spp_batch == samples_per_pixel."""
        out = np.zeros((self.settings.frame_h, self.settings.frame_w, 3), np.float32)
        check(_lib.load().wost_film_stderr(self.h, _p(out)))
        return out

    def traversal_stats(self, reset=True):
        """(full reference-order traversals, restricted second passes) of this integrator's closest-point queries"""
        a, b = ctypes.c_uint64(), ctypes.c_uint64()
        check(_lib.load().wost_traversal_stats(self.h, ctypes.byref(a), ctypes.byref(b), 1 if reset else 0))
        return a.value, b.value

    def film_device_ptr(self, channel=CH_SOLUTION):
        p = c_void_p()
        check(_lib.load().wost_film_device_ptr(self.h, channel, ctypes.byref(p)))
        return p.value

    def solution(self):
        f = self.film(CH_SOLUTION)
        return f[..., :3] / np.maximum(f[..., 4:5], 1e-30)

    @property
    def l2_window_bytes(self):
        """bytes of top-level LBVH records pinned in L2 for this integrator's stream (0: the tree fits the L2)"""
        n = c_u64()
        check(_lib.load().wost_l2_window_bytes(self.h, ctypes.byref(n)))
        return n.value

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


def settings_from_json(setting, **extra):
    """IntegratorSettings / GuidedIntegratorSettings JSON block (integrator/uniform/integrator.h:39-48,
    integrator/guided/integrator.h:75-89) -> Settings. All reference keys are required, like the reference's
    NLOHMANN_DEFINE_TYPE_INTRUSIVE deserialisation; guided keys only when present."""
    kw = dict(frame_size=tuple(setting["frameSize"]), max_walking_depth=setting["maxWalkingDepth"],
              samples_per_pixel=setting["samplesPerPixel"], epsilon_shell=setting["epsilonShell"],
              debug_pixel=setting.get("debugPixel", 0))
    if "trainSppCount" in setting:
        kw.update(train_spp_count=setting["trainSppCount"],
                  max_guided_depth_training=setting["maxGuidedDepthInTrainingPhase"],
                  max_guided_depth_guiding=setting["maxGuidedDepthInGuidingPhase"],
                  uniform_fraction_training=setting["uniformFractionInTrainingPhase"],
                  uniform_fraction_guiding=setting["uniformFractionInGuidingPhase"])
    kw.update(extra)
    s = make_settings(**kw)
    for key, attr in (("saveSppMetricsDuration", "save_spp_metrics_duration"), ("saveSppMetricsUntil", "save_spp_metrics_until"),
                      ("saveTimeMetricsDuration", "save_time_metrics_duration")):
        if key in setting:
            setattr(s, attr, int(setting[key]))
    return s


DEFAULT_NETWORK = {  # the "network" block of data/ladybug/n.json:49-81
    "encoding": {"otype": "DenseGrid", "n_levels": 8, "n_features_per_level": 4, "base_resolution": 8,
                 "per_level_scale": 1.4049999713897705, "interpolation": "Linear"},
    "loss": {"otype": "L2"},
    "network": {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 64, "n_hidden_layers": 3},
    "optimizer": {"otype": "Ema", "decay": 0.95, "nested": {"otype": "Adam", "learning_rate": 8e-3, "beta1": 0.9, "beta2": 0.99,
                                                           "epsilon": 1e-15, "l2_reg": 1e-6, "adabound": False}},
}


class GuidedIntegrator(_Integrator):
    """GuidedIntegrator<DIM> (integrator/guided/integrator.h:95-253): solve() trains the guiding network for
    trainSppCount samples per pixel, then keeps guiding with the learned mixture."""
    kind = GUIDED

    def __init__(self, problem, settings, network=None):
        super().__init__(problem, settings)
        self.dim = problem.dim
        self._cb = None
        self.resetNetwork(network if network is not None else DEFAULT_NETWORK)

    def resetNetwork(self, config):
        text = config if isinstance(config, str) else json.dumps(config)
        check(_lib.load().wost_reset_network(self.h, text.encode()))
        n, m, k = c_int(), c_int(), c_int()
        check(_lib.load().wost_guided_network_info(self.h, ctypes.byref(n), ctypes.byref(m), ctypes.byref(k)))
        self.n_params, self.n_mlp_params, self.encoded_width = n.value, m.value, k.value

    def resetTraining(self):
        check(_lib.load().wost_reset_training(self.h))

    def queryNetwork(self, p):
        p = np.ascontiguousarray(p, np.float32).reshape(self.dim)
        out = np.zeros(8 * (self.dim + 2) + 1, np.float32)
        check(_lib.load().wost_query_network(self.h, _p(p), _p(out)))
        return out

    def get_params(self, inference=False):
        out = np.zeros(self.n_params, np.float32)
        check(_lib.load().wost_guided_get_params(self.h, 1 if inference else 0, _p(out)))
        return out

    def set_params(self, params):
        params = np.ascontiguousarray(params, np.float32)
        assert len(params) == self.n_params
        check(_lib.load().wost_guided_set_params(self.h, _p(params)))

    def infer(self, pts, impl=0, inference_weights=True):
        """raw network outputs (n, 48) at world-space points; impl 0 = tcgen05 tile kernel, 1 = SIMT kernel"""
        pts = np.ascontiguousarray(pts, np.float32).reshape(-1, self.dim)
        out = np.zeros((len(pts), 48), np.float32)
        check(_lib.load().wost_guided_infer(self.h, _p(pts), len(pts), int(impl), 1 if inference_weights else 0, _p(out)))
        return out

    def train_batch(self, pos, dirs, nrm, dir_pdf, li, on_neumann, want_grad=True):
        """one optimiser step on explicit samples -> (mean loss, loss-scaled gradient slab | None)"""
        d = self.dim
        pos = np.ascontiguousarray(pos, np.float32).reshape(-1, d)
        dirs = np.ascontiguousarray(dirs, np.float32).reshape(-1, d)
        nrm = np.ascontiguousarray(nrm, np.float32).reshape(-1, d)
        dir_pdf = np.ascontiguousarray(dir_pdf, np.float32)
        li = np.ascontiguousarray(li, np.float32)
        on = np.ascontiguousarray(on_neumann, np.uint8)
        grad = np.zeros(self.n_params, np.float32) if want_grad else None
        loss = c_float()
        check(_lib.load().wost_guided_train_batch(self.h, _p(pos), _p(dirs), _p(nrm), _p(dir_pdf), _p(li), _p(on), len(pos), ctypes.byref(loss), _p(grad)))
        return loss.value, grad

    def set_allreduce(self, fn):
        """fn(device_ptr, n_floats, stream_ptr) -> None: enqueue a SUM all-reduce of the gradient slab on that stream"""
        if fn is None:
            self._cb = None
            check(_lib.load().wost_set_allreduce(self.h, None, None))
            return

        def tramp(user, buf, n, stream):
            try:
                fn(buf, n, stream)
                return 0
            except Exception:
                import traceback
                traceback.print_exc()
                return 1
        self._cb = _lib.ALLREDUCE_FN(tramp)
        check(_lib.load().wost_set_allreduce(self.h, ctypes.cast(self._cb, c_void_p), None))

    def last_training_samples(self):
        """the compacted training set of the last training frame -> dict(pos, dir, nrm, pdf, li, onN) like the oracle's emit_samples"""
        n = c_u64()
        L = _lib.load()
        check(L.wost_guided_last_samples(self.h, None, 0, ctypes.byref(n)))
        d = self.dim
        rows = np.zeros((n.value, 3 * d + 3), np.float32)
        if n.value:
            check(L.wost_guided_last_samples(self.h, _p(rows), n.value, ctypes.byref(n)))
        return dict(pos=rows[:, 0:d], dir=rows[:, d:2 * d], nrm=rows[:, 2 * d:3 * d], pdf=rows[:, 3 * d], li=rows[:, 3 * d + 1], onN=rows[:, 3 * d + 2] > 0.5)

    @property
    def profile(self):
        """device ms of the last solve by kernel class (needs WOST_PROFILE=1 in the environment)"""
        out = np.zeros(8, np.float32)
        check(_lib.load().wost_guided_last_profile(self.h, _p(out)))
        return dict(arrive=float(out[0]), bounce=float(out[1]), finish=float(out[2]), gaps=float(out[3]), train=float(out[4]))

    @property
    def train_stats(self):
        a, b, l = c_u64(), c_u64(), c_float()
        check(_lib.load().wost_guided_last_stats(self.h, ctypes.byref(a), ctypes.byref(b), ctypes.byref(l)))
        return dict(samples=a.value, steps=b.value, loss=l.value)


def vmm_eval(dim, raw, dirs, nrm, on_neumann, seed=0):
    """device probe of VMM<DIM,8>: (pdf, grad wrt activated lobe params, one sampled direction per row)"""
    raw = np.ascontiguousarray(raw, np.float32)
    dirs = np.ascontiguousarray(dirs, np.float32).reshape(-1, dim)
    nrm = np.ascontiguousarray(nrm, np.float32).reshape(-1, dim)
    on = np.ascontiguousarray(on_neumann, np.uint8)
    n = len(dirs)
    pdf = np.zeros(n, np.float32)
    grad = np.zeros((n, 8 * (dim + 2)), np.float32)
    smp = np.zeros((n, dim), np.float32)
    check(_lib.load().wost_vmm_eval(dim, _p(raw), _p(dirs), _p(nrm), _p(on), n, int(seed), _p(pdf), _p(grad), _p(smp)))
    return pdf, grad, smp
