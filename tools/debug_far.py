"""Debug helper (GPU box): closest-point index mismatches vs query distance (in units of the mesh's max |coordinate|)."""
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
    m = np.abs(v).max()
    for scale in [2, 8, 32, 128, 512, 2048, 8192, 32768, 131072]:
        d = rng.randn(20000, dim); d /= np.linalg.norm(d, axis=1, keepdims=True)
        q = (d * scale * m * (0.5 + rng.rand(20000, 1))).astype(np.float32)
        io, do = so.nearest(q); ig, dg = sg.nearest(q)
        print(name, "scale", scale, "idx mismatches", int((io != ig).sum()), "dist mismatches", int((bits(do) != bits(dg)).sum()), flush=True)
