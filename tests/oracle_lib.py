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
