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
