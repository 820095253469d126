"""Shared builders for GPU parity tests: the same scene goes to the oracle (CPU) and to libwost.so (GPU)."""
import numpy as np

from elaina_b200 import scenes


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def random_queries(verts, n, seed, margin=0.2):
    rng = np.random.RandomState(seed)
    dim = verts.shape[1]
    lo, hi = verts.min(0), verts.max(0)
    ext = hi - lo
    q = (lo - margin * ext + rng.rand(n, dim) * (1 + 2 * margin) * ext).astype(np.float32)
    d = rng.randn(n, dim).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    md = (rng.rand(n) * 2.0 * np.linalg.norm(ext)).astype(np.float32)
    u = rng.rand(n).astype(np.float32)
    rad = (rng.rand(n) * 0.6 * np.linalg.norm(ext)).astype(np.float32)
    return q, d, md, u, rad


def square_boundary(lo=-1.5, hi=1.5):
    v = np.array([[lo, lo], [hi, lo], [hi, hi], [lo, hi]], np.float32)
    p = np.array([[0, 1], [1, 2], [2, 3], [3, 0]], np.int32)
    return v, p


def mixed_problem_2d(nseg=400, seed=3):
    """Star-shaped Dirichlet curve inside a square Neumann box, gray harmonic boundary data + random Neumann flux."""
    dv, dp = scenes.star_polygon(nseg, 0.8, 0.25, 5)
    nv, npr = square_boundary(-1.4, 1.4)
    cd = scenes.gray_pairs(scenes.harmonic_2d(dv, 2))
    rng = np.random.RandomState(seed)
    cn = rng.rand(len(nv), 6).astype(np.float32) * 0.2
    return dict(dim=2, dverts=dv, dprims=dp, nverts=nv, nprims=npr, color_dirichlet=cd, color_neumann=cn,
                grid_scale=1.3, grid_pos=(0.0, 0.0), grid_up=(0.0, 1.0), aabb_min=(-1.5, -1.5), aabb_max=(1.5, 1.5))


def build_both(oracle, spec):
    """-> (oracle problem dict, product Problem) for a spec dict."""
    from elaina_b200 import api
    dim = spec["dim"]
    o_d = oracle.scene(dim, spec["dverts"], spec["dprims"]) if spec.get("dverts") is not None else None
    o_n = oracle.scene(dim, spec["nverts"], spec["nprims"]) if spec.get("nverts") is not None else None
    common = {k: spec[k] for k in ("color_dirichlet", "color_neumann", "dirichlet_intensity", "neumann_intensity", "source_intensity",
                                   "source", "source_origin", "source_voxel", "grid_scale", "grid_pos", "grid_up", "grid_right",
                                   "mask", "aabb_min", "aabb_max") if k in spec}
    oprob = dict(dim=dim, dirichlet=o_d, neumann=o_n, **common)
    g_d = api.Scene(dim, spec["dverts"], spec["dprims"]) if spec.get("dverts") is not None else None
    g_n = api.Scene(dim, spec["nverts"], spec["nprims"]) if spec.get("nverts") is not None else None
    gprob = api.Problem(dim, dirichlet=g_d, neumann=g_n, **common)
    return oprob, gprob


# ---- analytic Poisson / Neumann problems. Conventions of the reference's estimators
# (integrator/uniform/integrator.cu:234-445, util/green.h): the walker solves  -Laplace(u) = f  with u = g on the Dirichlet
# mesh and adds  -h * G  for the Neumann data h, i.e. h = -du/dn with n the OUTWARD normal of the domain.
def poisson_disk_2d(nseg=512, c=2.0):
    """unit disk, f = c, g = 0.5 x on the circle  ->  u = c (1 - r^2) / 4 + 0.5 x"""
    dv, dp = scenes.star_polygon(nseg, 1.0, 0.0, 5)
    src = np.full((4, 4, 4, 3), c, np.float32)
    spec = dict(dim=2, dverts=dv, dprims=dp, nverts=None, nprims=None, color_dirichlet=scenes.gray_pairs(0.5 * dv[:, 0]), color_neumann=None,
                source=src, source_origin=(-2.0, -2.0, -2.0), source_voxel=1.0 + 1.0 / 3.0,
                grid_scale=0.9, grid_pos=(0.0, 0.0), grid_up=(0.0, 1.0), aabb_min=(-1.5, -1.5), aabb_max=(1.5, 1.5))

    def u(p):
        return c * (1.0 - p[..., 0] ** 2 - p[..., 1] ** 2) / 4.0 + 0.5 * p[..., 0]

    def inside(p):
        return np.hypot(p[..., 0], p[..., 1]) < 0.97
    return spec, u, inside


def poisson_ball_3d(subdiv=4, c=3.0):
    """unit ball, f = c, g = 0.3 z on the sphere  ->  u = c (1 - r^2) / 6 + 0.3 z; evaluated on the slice y = 0.1"""
    dv, dp = scenes.icosphere(subdiv, 1.0, 0.0)
    src = np.full((4, 4, 4, 3), c, np.float32)
    spec = dict(dim=3, dverts=dv, dprims=dp, nverts=None, nprims=None, color_dirichlet=scenes.gray_pairs(0.3 * dv[:, 2]), color_neumann=None,
                source=src, source_origin=(-2.0, -2.0, -2.0), source_voxel=1.0 + 1.0 / 3.0,
                grid_scale=0.9, grid_pos=(0.0, 0.1, 0.0), grid_up=(0.0, 0.0, 1.0), grid_right=(1.0, 0.0, 0.0),
                aabb_min=(-1.5, -1.5, -1.5), aabb_max=(1.5, 1.5, 1.5))

    def u(p):
        return c * (1.0 - (p ** 2).sum(-1)) / 6.0 + 0.3 * p[..., 2]

    def inside(p):
        return np.sqrt((p ** 2).sum(-1)) < 0.93
    return spec, u, inside


def neumann_square_2d():
    """unit square: Dirichlet on x = 0 (g = y) and x = 1 (g = 2 y), Neumann on y = 0 / y = 1 with the flux of u = x y + y
    (du/dn = -(x + 1) at the bottom, +(x + 1) at the top; the estimator's data is h = -du/dn)"""
    dv = np.array([[0, 0], [0, 1], [1, 0], [1, 1]], np.float32)
    dp = np.array([[0, 1], [2, 3]], np.int32)
    nv = np.array([[0, 0], [1, 0], [0, 1], [1, 1]], np.float32)
    npr = np.array([[0, 1], [2, 3]], np.int32)
    g = dv[:, 0] * dv[:, 1] + dv[:, 1]
    h = np.array([+(nv[0, 0] + 1), +(nv[1, 0] + 1), -(nv[2, 0] + 1), -(nv[3, 0] + 1)], np.float32)  # -du/dn: bottom du/dn = -(x+1), top +(x+1)
    spec = dict(dim=2, dverts=dv, dprims=dp, nverts=nv, nprims=npr, color_dirichlet=scenes.gray_pairs(g), color_neumann=scenes.gray_pairs(h),
                grid_scale=0.4, grid_pos=(0.5, 0.5), grid_up=(0.0, 1.0), aabb_min=(-0.1, -0.1), aabb_max=(1.1, 1.1))

    def u(p):
        return p[..., 0] * p[..., 1] + p[..., 1]

    def inside(p):
        return np.ones(p.shape[:-1], bool)
    return spec, u, inside


def neumann_cube_3d():
    """unit cube: Dirichlet on x = 0 (g = z) and x = 1 (g = y + z), Neumann on the other four faces with the flux of
    u = x y + z (du/dn = -x on y = 0, +x on y = 1, -1 on z = 0, +1 on z = 1); evaluated on the slice z = 0.45"""
    v = np.array([[x, y, z] for z in (0, 1) for y in (0, 1) for x in (0, 1)], np.float32)  # index = x + 2y + 4z

    def quad(a, b, c, d):
        return [[a, b, c], [a, c, d]]
    dp = np.array(quad(0, 4, 6, 2) + quad(1, 3, 7, 5), np.int32)                      # x = 0, x = 1
    npr = np.array(quad(0, 1, 5, 4) + quad(2, 6, 7, 3) + quad(0, 2, 3, 1) + quad(4, 5, 7, 6), np.int32)  # y = 0, y = 1, z = 0, z = 1
    g = v[:, 0] * v[:, 1] + v[:, 2]
    # every Neumann face needs its own flux at shared vertices: duplicate the vertices per face
    nv, nidx, h = [], [], []
    dudn = [lambda p: -p[0], lambda p: p[0], lambda p: -1.0, lambda p: 1.0]
    for f in range(4):
        base = len(nv)
        ids = sorted(set(npr[2 * f].tolist() + npr[2 * f + 1].tolist()))
        remap = {i: base + k for k, i in enumerate(ids)}
        for i in ids:
            nv.append(v[i])
            h.append(-dudn[f](v[i]))
        for t in (npr[2 * f], npr[2 * f + 1]):
            nidx.append([remap[int(i)] for i in t])
    spec = dict(dim=3, dverts=v, dprims=dp, nverts=np.asarray(nv, np.float32), nprims=np.asarray(nidx, np.int32),
                color_dirichlet=scenes.gray_pairs(g), color_neumann=scenes.gray_pairs(np.asarray(h, np.float32)),
                grid_scale=0.4, grid_pos=(0.5, 0.5, 0.45), grid_up=(0.0, 1.0, 0.0), grid_right=(1.0, 0.0, 0.0),
                aabb_min=(-0.1, -0.1, -0.1), aabb_max=(1.1, 1.1, 1.1))

    def u(p):
        return p[..., 0] * p[..., 1] + p[..., 2]

    def inside(p):
        return np.ones(p.shape[:-1], bool)
    return spec, u, inside


ANALYTIC = {"poisson_disk_2d": poisson_disk_2d, "poisson_ball_3d": poisson_ball_3d, "neumann_square_2d": neumann_square_2d, "neumann_cube_3d": neumann_cube_3d}
