"""ctypes bindings for the TEST-ONLY oracle libraries.

oracle/libwost_oracle.so  -- our scalar CPU restatement (always present after `make -C oracle`)
oracle/_ref/libref_lbvh.so -- the reference's own LBVH headers host-compiled (present when built in the
                              authoring container; travels to the GPU box prebuilt)
Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs import this module.
"""
import ctypes
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
ORACLE_SO = os.path.join(ORACLE_DIR, "libwost_oracle.so")
REF_SO = os.path.join(ORACLE_DIR, "_ref", "libref_lbvh.so")

c_void_p, c_int, c_float, c_u32, c_u64 = ctypes.c_void_p, ctypes.c_int, ctypes.c_float, ctypes.c_uint32, ctypes.c_uint64


def _p(a):
    return a.ctypes.data_as(c_void_p) if a is not None else None


def build_oracle(force=False):
    if force or not os.path.exists(ORACLE_SO):
        subprocess.check_call(["make", "-C", ORACLE_DIR, "libwost_oracle.so"], stdout=subprocess.DEVNULL)
    if os.path.isdir("/root/reference/ext/lbvh/include") and (force or not os.path.exists(REF_SO)):
        subprocess.check_call(["make", "-C", ORACLE_DIR, "ref"], stdout=subprocess.DEVNULL)


class ProblemDesc(ctypes.Structure):
    _fields_ = [
        ("dim", c_int),
        ("dirichlet", c_void_p),
        ("neumann", c_void_p),
        ("color_dirichlet", c_void_p),
        ("color_neumann", c_void_p),
        ("dirichlet_intensity", c_float),
        ("neumann_intensity", c_float),
        ("source_intensity", c_float),
        ("source", c_void_p),
        ("source_dims", c_int * 3),
        ("source_origin", c_float * 3),
        ("source_voxel", c_float),
        ("grid_scale", c_float),
        ("grid_pos", c_float * 3),
        ("grid_up", c_float * 3),
        ("grid_right", c_float * 3),
        ("mask", c_void_p),
        ("aabb_min", c_float * 3),
        ("aabb_max", c_float * 3),
    ]


class Settings(ctypes.Structure):
    _fields_ = [
        ("frame_w", c_int),
        ("frame_h", c_int),
        ("max_walking_depth", c_int),
        ("samples_per_pixel", c_int),
        ("epsilon_shell", c_float),
        ("seed", c_u32),
    ]


class GuidedSettings(ctypes.Structure):
    _fields_ = [("max_guided_depth", c_int), ("uniform_fraction", c_float), ("max_train_depth", c_int)]


class OracleNet:
    """DenseGrid + MLP of the oracle (fp32 parameter slab [W0|W1|W2|W3|grid])."""

    def __init__(self, orc, dim, n_levels, base_res, per_level_scale, aabb_min, aabb_max):
        self.orc, self.dim, self.n_levels = orc, dim, n_levels
        lo = np.zeros(3, np.float32)
        hi = np.ones(3, np.float32)
        lo[:dim] = aabb_min[:dim]
        hi[:dim] = aabb_max[:dim]
        self.h = c_void_p(orc.lib.orc_net_create(dim, n_levels, base_res, float(per_level_scale), _p(lo), _p(hi)))
        nm, k0 = c_int(), c_int()
        self.n_params = orc.lib.orc_net_num_params(self.h, ctypes.byref(nm), ctypes.byref(k0))
        self.n_mlp, self.k0 = nm.value, k0.value

    def set_params(self, params):
        params = np.ascontiguousarray(params, np.float32)
        assert len(params) == self.n_params
        self.params = params
        self.orc.lib.orc_net_set_params(self.h, _p(params))

    def forward(self, pts, want_acts=False, want_cells=False):
        pts = np.ascontiguousarray(pts, np.float32).reshape(-1, self.dim)
        n = len(pts)
        out = np.zeros((n, 48), np.float32)
        acts = np.zeros((n, 4, 64), np.float32) if want_acts else None
        C = 1 << self.dim
        ci = np.zeros((n, self.n_levels, C), np.uint32) if want_cells else None
        cw = np.zeros((n, self.n_levels, C), np.float32) if want_cells else None
        self.orc.lib.orc_net_forward(self.h, _p(pts), n, _p(out), _p(acts), _p(ci), _p(cw))
        return out, acts, ci, cw

    def close(self):
        if self.h:
            self.orc.lib.orc_net_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class _GeomLib:
    """Common surface of the oracle and the host-compiled reference (prefix 'orc_' / 'ref_')."""

    def __init__(self, path, prefix):
        self.lib = ctypes.CDLL(path)
        self.prefix = prefix
        f = self._f
        f("scene_create").restype = c_void_p
        f("scene_create").argtypes = [c_int, c_void_p, c_int, c_void_p, c_int]
        f("scene_destroy").argtypes = [c_void_p]
        f("num_nodes").argtypes = [c_void_p]
        f("dump_bvh").argtypes = [c_void_p] * 4
        f("dump_silhouette_assignment").argtypes = [c_void_p] * 2
        f("num_silhouettes").argtypes = [c_void_p]
        f("dump_silhouettes").argtypes = [c_void_p] * 2
        f("query_nearest").argtypes = [c_void_p, c_void_p, c_int, c_void_p, c_void_p]
        f("query_silhouette").argtypes = [c_void_p, c_void_p, c_int, c_int, c_void_p]
        f("query_ray").argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p]
        f("query_ray_any").argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p]
        f("sample_in_sphere").argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p]

    def _f(self, name):
        return getattr(self.lib, self.prefix + name)

    def scene(self, dim, verts, prims):
        return GeomScene(self, dim, verts, prims)


class GeomScene:
    def __init__(self, g, dim, verts, prims):
        self.g, self.dim = g, dim
        self.verts = np.ascontiguousarray(verts, np.float32).reshape(-1, dim)
        self.prims = np.ascontiguousarray(prims, np.int32).reshape(-1, dim)
        self.h = c_void_p(g._f("scene_create")(dim, _p(self.verts), len(self.verts), _p(self.prims), len(self.prims)))

    def close(self):
        if self.h:
            self.g._f("scene_destroy")(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def num_nodes(self):
        return self.g._f("num_nodes")(self.h)

    def dump_bvh(self):
        n, d = self.num_nodes, self.dim
        nodes = np.zeros((n, 4), np.uint32)
        aabbs = np.zeros((n, 2 * d), np.float32)
        cones = np.zeros((n, d + 2), np.float32)
        self.g._f("dump_bvh")(self.h, _p(nodes), _p(aabbs), _p(cones))
        return nodes, aabbs, cones

    def silhouette_assignment(self):
        out = np.zeros((len(self.prims), self.dim), np.int32)
        self.g._f("dump_silhouette_assignment")(self.h, _p(out))
        return out

    def silhouettes(self):
        n = self.g._f("num_silhouettes")(self.h)
        out = np.zeros((n, 4), np.int32)
        self.g._f("dump_silhouettes")(self.h, _p(out))
        return out

    def nearest(self, pts):
        pts = np.ascontiguousarray(pts, np.float32).reshape(-1, self.dim)
        idx = np.zeros(len(pts), np.uint32)
        dist = np.zeros(len(pts), np.float32)
        self.g._f("query_nearest")(self.h, _p(pts), len(pts), _p(idx), _p(dist))
        return idx, dist

    def silhouette(self, pts, flip=False):
        pts = np.ascontiguousarray(pts, np.float32).reshape(-1, self.dim)
        dist = np.zeros(len(pts), np.float32)
        self.g._f("query_silhouette")(self.h, _p(pts), len(pts), int(flip), _p(dist))
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
        self.g._f("query_ray")(self.h, _p(org), _p(dr), _p(md), n, _p(hit), _p(t), _p(uv), _p(obj))
        return hit, t, uv, obj

    def ray_any(self, org, dr, max_dist):
        org = np.ascontiguousarray(org, np.float32).reshape(-1, self.dim)
        dr = np.ascontiguousarray(dr, np.float32).reshape(-1, self.dim)
        n = len(org)
        md = np.ascontiguousarray(np.broadcast_to(np.asarray(max_dist, np.float32), (n,)))
        hit = np.zeros(n, np.int32)
        self.g._f("query_ray_any")(self.h, _p(org), _p(dr), _p(md), n, _p(hit))
        return hit

    def sample_in_sphere(self, ctr, radius, u):
        ctr = np.ascontiguousarray(ctr, np.float32).reshape(-1, self.dim)
        n = len(ctr)
        radius = np.ascontiguousarray(np.broadcast_to(np.asarray(radius, np.float32), (n,)))
        u = np.ascontiguousarray(u, np.float32)
        obj = np.zeros(n, np.int32)
        pdf = np.zeros(n, np.float32)
        self.g._f("sample_in_sphere")(self.h, _p(ctr), _p(radius), _p(u), n, _p(obj), _p(pdf))
        return obj, pdf


class Oracle(_GeomLib):
    def __init__(self):
        build_oracle()
        super().__init__(ORACLE_SO, "orc_")
        L = self.lib
        L.orc_set_math_mode.argtypes = [c_int]
        L.orc_used_morton64.argtypes = [c_void_p]
        L.orc_dump_sorted.argtypes = [c_void_p] * 3
        L.orc_solve_uniform.argtypes = [ctypes.POINTER(ProblemDesc), ctypes.POINTER(Settings), c_int, c_int, c_void_p, c_void_p, c_int]
        L.orc_eval_point.argtypes = [ctypes.POINTER(ProblemDesc), ctypes.POINTER(Settings), c_int, c_void_p]
        L.orc_detmath_eval.argtypes = [c_int, c_void_p, c_void_p, c_int, c_void_p]
        L.orc_philox.argtypes = [c_u32] * 6 + [c_void_p]

    def set_math_mode(self, mode):
        self.lib.orc_set_math_mode(int(mode))

    def set_rng_mode(self, mode):
        """0 = Philox (product stream), 1 = the reference's per-pixel PCG32 (uniform walker only)"""
        self.lib.orc_set_rng_mode.argtypes = [c_int]
        self.lib.orc_set_rng_mode(int(mode))

    def used_morton64(self, scene):
        return bool(self.lib.orc_used_morton64(scene.h))

    def dump_sorted(self, scene):
        n = len(scene.prims)
        m = np.zeros(n, np.uint32)
        o = np.zeros(n, np.uint32)
        self.lib.orc_dump_sorted(scene.h, _p(m), _p(o))
        return m, o

    def detmath(self, fn, x, y=None):
        x = np.ascontiguousarray(x, np.float32)
        y = np.ascontiguousarray(y if y is not None else np.zeros_like(x), np.float32)
        out = np.zeros_like(x)
        self.lib.orc_detmath_eval(fn, _p(x), _p(y), len(x), _p(out))
        return out

    def philox(self, ctr, key):
        out = np.zeros(4, np.uint32)
        self.lib.orc_philox(*[int(c) for c in ctr], *[int(k) for k in key], _p(out))
        return out

    def solve_uniform(self, problem, settings, pixel_begin=0, pixel_end=None, nthreads=0):
        """problem: dict from tests/scenes.py (see make_problem_desc)."""
        desc, keep = make_problem_desc(problem)
        st = Settings(settings["frame_w"], settings["frame_h"], settings["max_walking_depth"],
                      settings["samples_per_pixel"], settings["epsilon_shell"], settings.get("seed", 0))
        npx = st.frame_w * st.frame_h
        if pixel_end is None:
            pixel_end = npx
        sol = np.zeros((npx, 3), np.float32)
        steps = c_u64(0)
        self.lib.orc_solve_uniform(ctypes.byref(desc), ctypes.byref(st), pixel_begin, pixel_end, _p(sol), ctypes.byref(steps), nthreads)
        del keep
        return sol, steps.value

    # ---- guided integrator (oracle/wost_oracle_guided.inl)
    def _bind_guided(self):
        if getattr(self, "_guided_bound", False):
            return
        L = self.lib
        L.orc_net_create.restype = c_void_p
        L.orc_net_create.argtypes = [c_int, c_int, c_int, c_float, c_void_p, c_void_p]
        L.orc_net_destroy.argtypes = [c_void_p]
        L.orc_net_num_params.argtypes = [c_void_p, c_void_p, c_void_p]
        L.orc_net_set_params.argtypes = [c_void_p, c_void_p]
        L.orc_net_forward.argtypes = [c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p]
        L.orc_vmm_eval.argtypes = [c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_u32, c_void_p, c_void_p, c_void_p]
        L.orc_kl_grad.argtypes = [c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_float, c_void_p, c_void_p]
        for name in ("orc_eval_poly_large0",):
            getattr(L, name).restype = c_float
            getattr(L, name).argtypes = [c_float]
        L.orc_log_bessel.restype = c_float
        L.orc_log_bessel.argtypes = [c_float, c_int]
        for name in ("orc_vonmises_log_pdf", "orc_vonmises_pdf", "orc_vonmises_dlog_dkappa", "orc_vmf_pdf"):
            getattr(L, name).restype = c_float
            getattr(L, name).argtypes = [c_float, c_float]
        L.orc_solve_guided.restype = c_int
        L.orc_solve_guided.argtypes = [ctypes.POINTER(ProblemDesc), ctypes.POINTER(Settings), ctypes.POINTER(GuidedSettings), c_void_p, c_int, c_int,
                                       c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int]
        self._guided_bound = True

    def net(self, dim, n_levels, base_res, per_level_scale, aabb_min, aabb_max):
        self._bind_guided()
        return OracleNet(self, dim, n_levels, base_res, per_level_scale, aabb_min, aabb_max)

    def vmm_eval(self, dim, raw, dirs, nrm, on_neumann, seed=0):
        self._bind_guided()
        raw = np.ascontiguousarray(raw, np.float32)
        dirs = np.ascontiguousarray(dirs, np.float32).reshape(-1, dim)
        nrm = np.ascontiguousarray(nrm, np.float32).reshape(-1, dim)
        on = np.ascontiguousarray(on_neumann, np.uint8)
        n = len(dirs)
        pdf = np.zeros(n, np.float32)
        grad = np.zeros((n, 8 * (dim + 2)), np.float32)
        smp = np.zeros((n, dim), np.float32)
        self.lib.orc_vmm_eval(dim, _p(raw), _p(dirs), _p(nrm), _p(on), n, seed, _p(pdf), _p(grad), _p(smp))
        return pdf, grad, smp

    def kl_grad(self, dim, raw48, dirs, nrm, on_neumann, dir_pdf, li, scale):
        self._bind_guided()
        raw48 = np.ascontiguousarray(raw48, np.float32).reshape(-1, 48)
        n = len(raw48)
        dirs = np.ascontiguousarray(dirs, np.float32).reshape(-1, dim)
        nrm = np.ascontiguousarray(nrm, np.float32).reshape(-1, dim)
        on = np.ascontiguousarray(on_neumann, np.uint8)
        dir_pdf = np.ascontiguousarray(dir_pdf, np.float32)
        li = np.ascontiguousarray(li, np.float32)
        dl = np.zeros((n, 48), np.float32)
        loss = np.zeros(n, np.float32)
        self.lib.orc_kl_grad(dim, _p(raw48), _p(dirs), _p(nrm), _p(on), _p(dir_pdf), _p(li), n, float(scale), _p(dl), _p(loss))
        return dl, loss

    def solve_guided(self, problem, settings, net, max_guided_depth, uniform_fraction, max_train_depth=3, sample0=0, nsamples=None,
                     emit_samples=False, pixel_begin=0, pixel_end=None, nthreads=0):
        """-> (solution (npix,3) = sum of the samples' contributions / samples_per_pixel, steps, samples dict | None)"""
        self._bind_guided()
        desc, keep = make_problem_desc(problem)
        st = Settings(settings["frame_w"], settings["frame_h"], settings["max_walking_depth"],
                      settings["samples_per_pixel"], settings["epsilon_shell"], settings.get("seed", 0))
        gs = GuidedSettings(int(max_guided_depth), float(uniform_fraction), int(max_train_depth))
        npx = st.frame_w * st.frame_h
        if pixel_end is None:
            pixel_end = npx
        if nsamples is None:
            nsamples = st.samples_per_pixel - sample0
        sol = np.zeros((npx, 3), np.float32)
        steps = c_u64(0)
        cap = 4 * npx * nsamples if emit_samples else 0
        so = np.zeros((max(cap, 1), 11), np.float32)
        son = np.zeros(max(cap, 1), np.uint8)
        n = self.lib.orc_solve_guided(ctypes.byref(desc), ctypes.byref(st), ctypes.byref(gs), net.h if net is not None else None, pixel_begin, pixel_end,
                                      sample0, nsamples, _p(sol), ctypes.byref(steps), _p(so) if emit_samples else None,
                                      _p(son) if emit_samples else None, cap, nthreads)
        del keep
        samples = None
        if emit_samples:
            dim = problem["dim"]
            so, son = so[:n], son[:n]
            samples = dict(pos=so[:, 0:dim].copy(), dir=so[:, 3:3 + dim].copy(), nrm=so[:, 6:6 + dim].copy(), pdf=so[:, 9].copy(), li=so[:, 10].copy(), onN=son.copy())
        return sol, steps.value, samples

    def eval_points(self, problem, settings, pixels):
        desc, keep = make_problem_desc(problem)
        st = Settings(settings["frame_w"], settings["frame_h"], settings["max_walking_depth"],
                      settings["samples_per_pixel"], settings["epsilon_shell"], settings.get("seed", 0))
        out = np.zeros((len(pixels), problem["dim"]), np.float32)
        tmp = np.zeros(3, np.float32)
        for i, px in enumerate(pixels):
            self.lib.orc_eval_point(ctypes.byref(desc), ctypes.byref(st), int(px), _p(tmp))
            out[i] = tmp[: problem["dim"]]
        return out


def make_problem_desc(problem):
    """problem keys: dim, dirichlet (GeomScene|None), neumann, color_dirichlet (nv,6)|None, color_neumann,
    dirichlet_intensity, neumann_intensity, source_intensity, source (nz,ny,nx,3)|None, source_origin, source_voxel,
    grid_scale, grid_pos, grid_up, grid_right, mask, aabb_min, aabb_max"""
    d = ProblemDesc()
    keep = []
    dim = problem["dim"]
    d.dim = dim
    d.dirichlet = problem["dirichlet"].h if problem.get("dirichlet") is not None else None
    d.neumann = problem["neumann"].h if problem.get("neumann") is not None else None
    for key in ("color_dirichlet", "color_neumann"):
        a = problem.get(key)
        if a is not None:
            a = np.ascontiguousarray(a, np.float32)
            keep.append(a)
            setattr(d, key, a.ctypes.data)
    d.dirichlet_intensity = problem.get("dirichlet_intensity", 1.0)
    d.neumann_intensity = problem.get("neumann_intensity", 1.0)
    d.source_intensity = problem.get("source_intensity", 1.0)
    src = problem.get("source")
    if src is not None:
        src = np.ascontiguousarray(src, np.float32)
        keep.append(src)
        d.source = src.ctypes.data
        nz, ny, nx = src.shape[:3]
        d.source_dims = (c_int * 3)(nx, ny, nz)
        so = list(problem.get("source_origin", (0, 0, 0))) + [0, 0, 0]
        d.source_origin = (c_float * 3)(*so[:3])
        d.source_voxel = problem.get("source_voxel", 1.0)
    d.grid_scale = problem.get("grid_scale", 1.0)

    def v3(x, default):
        x = list(x if x is not None else default) + [0.0, 0.0, 0.0]
        return (c_float * 3)(*x[:3])

    d.grid_pos = v3(problem.get("grid_pos"), (0, 0, 0))
    d.grid_up = v3(problem.get("grid_up"), (0, 1, 0) if dim == 2 else (0, 0, 1))
    d.grid_right = v3(problem.get("grid_right"), (1, 0, 0))
    m = problem.get("mask")
    if m is not None:
        m = np.ascontiguousarray(m, np.uint8)
        keep.append(m)
        d.mask = m.ctypes.data
    d.aabb_min = v3(problem.get("aabb_min"), (0, 0, 0))
    d.aabb_max = v3(problem.get("aabb_max"), (1, 1, 1))
    return d, keep


class RefLBVH(_GeomLib):
    def __init__(self):
        build_oracle()
        if not os.path.exists(REF_SO):
            raise FileNotFoundError(REF_SO)
        super().__init__(REF_SO, "ref_")


def have_ref():
    build_oracle()
    return os.path.exists(REF_SO)
