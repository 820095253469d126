"""Debug helper (GPU box): which closest-point queries differ between the oracle and the product, and why."""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
from oracle_lib import Oracle
from elaina_b200 import api, scenes
from test_gpu_lbvh import _scene_cases
from helpers import bits
o = Oracle(); o.set_math_mode(0)
for name in ["star10k", "ico5", "soup2d_random", "soup3d_random"]:
    dim, v, p = _scene_cases()[name]
    so = o.scene(dim, v, p); sg = api.Scene(dim, v, p)
    rng = np.random.RandomState(5)
    vq = v[rng.randint(0, len(v), 4000)]
    mid = v[p[rng.randint(0, len(p), 4000)]].mean(1).astype(np.float32)
    off = (vq + rng.randn(*vq.shape).astype(np.float32) * 1e-4).astype(np.float32)
    far = (rng.randn(4000, dim) * np.array([10.0, 1e3, 1e5, 1e7])[rng.randint(0, 4, 4000)][:, None]).astype(np.float32)
    for cname, q in (("vq", vq), ("mid", mid), ("off", off), ("far", far)):
        io, do = so.nearest(q); ig, dg = sg.nearest(q)
        bad = np.nonzero(io != ig)[0]
        badd = np.nonzero(bits(do) != bits(dg))[0]
        print(name, cname, "idx mismatches", len(bad), "dist mismatches", len(badd))
        for b in bad[:6]:
            print("   q", q[b], "|q|", np.linalg.norm(q[b]), "oracle", io[b], do[b], "gpu", ig[b], dg[b])
