"""Generates tests/golden/lbvh_golden.npz from the REFERENCE ITSELF (host-compiled snch-lbvh, oracle/_ref).

Run in the authoring container only (needs /root/reference): `python tests/golden/make_lbvh_golden.py`.
The fixture pins, for small deterministic scenes, the reference's sorted Morton order, node topology, boxes,
and the answers of closest-point / closest-silhouette / ray / sphere-sample queries at seeded query points.
The scenes are regenerated from the recipe stored in the fixture (no reference data travels, except the
4-segment square of data/ladybug/boundary.obj whose 4 vertices are stored inline).
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle_lib import RefLBVH  # noqa: E402
from elaina_b200 import scenes  # noqa: E402


def golden_scenes():
    out = {}
    v, p = scenes.load_obj("/root/reference/data/ladybug/boundary.obj", 2)
    out["square"] = (2, v, p)
    v, p = scenes.star_polygon(600)
    out["star600"] = (2, v, p)
    # open polyline soup with repeated/degenerate morton cells (forces the 64-bit key path)
    rng = np.random.RandomState(5)
    base = rng.rand(40, 2).astype(np.float32)
    pts = np.concatenate([base + 1e-4 * rng.rand(40, 2).astype(np.float32) * k for k in range(4)], 0).astype(np.float32)
    idx = np.arange(len(pts) - 1, dtype=np.int32)
    out["soup2d"] = (2, pts, np.stack([idx, idx + 1], 1).astype(np.int32))
    v, p = scenes.icosphere(2, 1.0, 0.15)
    out["ico2"] = (3, v, p)
    # open 3D patch (boundary edges have one face)
    g = np.linspace(-1, 1, 9, dtype=np.float32)
    X, Y = np.meshgrid(g, g, indexing="ij")
    V = np.stack([X.ravel(), Y.ravel(), (0.3 * np.sin(3 * X) * np.cos(2 * Y)).ravel()], 1).astype(np.float32)
    F = []
    for i in range(8):
        for j in range(8):
            a, b, c, d = i * 9 + j, i * 9 + j + 1, (i + 1) * 9 + j, (i + 1) * 9 + j + 1
            F += [[a, c, b], [b, c, d]]
    out["patch3d"] = (3, V, np.asarray(F, np.int32))
    return out


def main():
    R = RefLBVH()
    fix = {}
    for name, (dim, v, p) in golden_scenes().items():
        s = R.scene(dim, v, p)
        nodes, aabbs, cones = s.dump_bvh()
        rng = np.random.RandomState(11)
        lo, hi = v.min(0), v.max(0)
        ext = hi - lo
        q = (lo - 0.2 * ext + rng.rand(256, dim) * 1.4 * ext).astype(np.float32)
        idx, dist = s.nearest(q)
        sil0 = s.silhouette(q, False)
        sil1 = s.silhouette(q, True)
        d = rng.randn(256, dim).astype(np.float32)
        d /= np.linalg.norm(d, axis=1, keepdims=True)
        md = (rng.rand(256) * 2.0 * np.linalg.norm(ext)).astype(np.float32)
        hit, t, uv, obj = s.ray(q, d, md)
        anyhit = s.ray_any(q, d, md)
        u = rng.rand(256).astype(np.float32)
        rad = (rng.rand(256) * 0.6 * np.linalg.norm(ext)).astype(np.float32)
        sobj, spdf = s.sample_in_sphere(q, rad, u)
        pre = name + "/"
        fix.update({pre + "dim": np.int32(dim), pre + "verts": v, pre + "prims": p, pre + "nodes": nodes, pre + "aabbs": aabbs,
                    pre + "cone_half_angle": cones[:, dim], pre + "sil_assign": s.silhouette_assignment(), pre + "sil_rec": s.silhouettes(),
                    pre + "q": q, pre + "idx": idx, pre + "dist": dist, pre + "sil0": sil0, pre + "sil1": sil1,
                    pre + "dir": d, pre + "max_dist": md, pre + "hit": hit, pre + "t": t, pre + "uv": uv, pre + "obj": obj, pre + "any": anyhit,
                    pre + "u": u, pre + "rad": rad, pre + "sobj": sobj, pre + "spdf": spdf})
    np.savez_compressed(os.path.join(HERE, "lbvh_golden.npz"), **fix)
    print("wrote", os.path.join(HERE, "lbvh_golden.npz"), len(fix), "arrays")


if __name__ == "__main__":
    main()
