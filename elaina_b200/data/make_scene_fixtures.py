#!/usr/bin/env python
"""Scene fixtures of BASELINE.json configs[0] / configs[1]: the shipped meshes data/{ladybug,fille}/{model,boundary}.obj
as compressed arrays, so that the GPU box (which has no /root/reference) can run the real geometry.

  python tests/golden/make_scene_fixtures.py      (needs /root/reference; writes tests/golden/scene_<name>.npz)

Vertices are stored as the float32 the reference's loader produces (`line_stream >> float`, i.e. strtof of the OBJ text,
ext/lbvh/include/snch_lbvh/scene_loader.cuh:49-82), primitives as 0-based int32 pairs. DATA only -- no reference source.
The missing color.json of both scenes (SURVEY 7 "Missing inputs") is synthesised at run time by
elaina_b200.scenes.synth_stroke_colors (seeded, per connected stroke), identically for every consumer.
"""
import ctypes
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference/data"
libc = ctypes.CDLL(None)
libc.strtof.restype = ctypes.c_float
libc.strtof.argtypes = [ctypes.c_char_p, ctypes.c_void_p]


def load(path):
    verts, prims = [], []
    for line in open(path, "rb"):
        p = line.split()
        if not p:
            continue
        if p[0] == b"v":
            verts.append([libc.strtof(p[1], None), libc.strtof(p[2], None)])
        elif p[0] == b"l":
            prims.append([int(p[1]) - 1, int(p[2]) - 1])
    return np.asarray(verts, np.float32), np.asarray(prims, np.int32)


def main():
    for name in ("ladybug", "fille"):
        dv, dp = load(os.path.join(REF, name, "model.obj"))
        nv, npr = load(os.path.join(REF, name, "boundary.obj"))
        out = os.path.join(HERE, "scene_%s.npz" % name)
        np.savez_compressed(out, dverts=dv, dprims=dp, nverts=nv, nprims=npr)
        print(name, dv.shape, dp.shape, nv.shape, npr.shape, os.path.getsize(out), "bytes")


if __name__ == "__main__":
    if not os.path.isdir(REF):
        sys.exit("needs /root/reference/data")
    main()
