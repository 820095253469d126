"""Scene ingest + synthetic scene generators (host side, numpy only).

`load_obj` mirrors the reference OBJ reader (ext/lbvh/include/snch_lbvh/scene_loader.cuh:49-82,118-151):
only `v x y z`, `l i j` (2D) and `f i j k` (3D, bare indices) records are honoured, indices are 1-based.
The generators build the synthetic configurations named in BASELINE.json.
"""
import json
import math
import os

import numpy as np


def load_obj(path, dim):
    verts, prims = [], []
    tag = "l" if dim == 2 else "f"
    with open(path, "r") as f:
        for line in f:
            p = line.split()
            if not p:
                continue
            if p[0] == "v":
                verts.append([float(x) for x in p[1:4]][:dim])
            elif p[0] == tag:
                prims.append([int(x.split("/")[0]) - 1 for x in p[1 : 1 + dim]])
    if not verts:
        raise RuntimeError("Could not read any vertex from " + path)
    return np.asarray(verts, np.float32).reshape(-1, dim), np.asarray(prims, np.int32).reshape(-1, dim)


def save_obj(path, verts, prims):
    dim = verts.shape[1]
    with open(path, "w") as f:
        for v in verts:
            f.write("v %r %r %r\n" % (float(v[0]), float(v[1]), float(v[2]) if dim == 3 else 0.0))
        tag = "l" if dim == 2 else "f"
        for p in prims:
            f.write(tag + " " + " ".join(str(int(i) + 1) for i in p) + "\n")


def load_vertex_colors(path):
    """core/problem.cu:63-96: {"ColorConfigurations":[{vertexID (1-based, sorted), leftColor{R,G,B}, rightColor{R,G,B}}]}"""
    with open(path, "r") as f:
        conf = json.load(f)
    items = conf["ColorConfigurations"]
    if not isinstance(items, list):
        raise RuntimeError("The ColorConfigurations item is not an array.")
    out = np.zeros((len(items), 6), np.float32)
    for i, it in enumerate(items):
        if int(it["vertexID"]) != i + 1:
            raise RuntimeError("The configurations should be sorted.")
        out[i, 0:3] = [it["leftColor"][c] for c in "RGB"]
        out[i, 3:6] = [it["rightColor"][c] for c in "RGB"]
    return out


def save_vertex_colors(path, colors):
    items = []
    for i, c in enumerate(np.asarray(colors, np.float32)):
        items.append({"vertexID": i + 1,
                      "leftColor": {"R": float(c[0]), "G": float(c[1]), "B": float(c[2])},
                      "rightColor": {"R": float(c[3]), "G": float(c[4]), "B": float(c[5])}})
    with open(path, "w") as f:
        json.dump({"ColorConfigurations": items}, f)


def load_nvdb_source(path):
    """core/problem.cu:135-149 (nanovdb::io::readGrid): -> (dense[nz, ny, nx, 3] float32, origin (x, y, z), voxel size) of the
    first Vec3f grid of an uncompressed .nvdb file, flattened over its index bounding box by libwost (csrc/nvdb.h)."""
    import ctypes
    from . import _lib
    L = _lib.load()
    dims, origin, voxel = (ctypes.c_int * 3)(), (ctypes.c_float * 3)(), ctypes.c_float()
    _lib.check(L.wost_nvdb_read_vec3f(str(path).encode(), dims, origin, ctypes.byref(voxel), None, 0))
    out = np.zeros((dims[2], dims[1], dims[0], 3), np.float32)
    _lib.check(L.wost_nvdb_read_vec3f(str(path).encode(), dims, origin, ctypes.byref(voxel), out.ctypes.data, out.size))
    return out, (origin[0], origin[1], origin[2]), voxel.value


def synth_stroke_colors(nverts, prims, seed=7):
    """Deterministic stand-in for the missing data/*/color.json (.MISSING_LARGE_BLOBS): piecewise-constant
    left/right colour pairs per connected stroke of the polyline soup (diffusion-curve style)."""
    rng = np.random.RandomState(seed)
    parent = np.arange(nverts)

    def find(a):
        while parent[a] != a:
            parent[a] = parent[parent[a]]
            a = parent[a]
        return a

    for p in np.asarray(prims):
        ra, rb = find(int(p[0])), find(int(p[-1]))
        if ra != rb:
            parent[rb] = ra
        if len(p) == 3:
            rc = find(int(p[1]))
            ra = find(int(p[0]))
            if ra != rc:
                parent[rc] = ra
    roots = np.array([find(i) for i in range(nverts)])
    uniq, inv = np.unique(roots, return_inverse=True)
    table = rng.rand(len(uniq), 6).astype(np.float32)
    return table[inv]


def star_polygon(nseg=10000, r0=1.0, amp=0.3, lobes=5):
    """C3: closed polygon r(theta) = r0 + amp*sin(lobes*theta), CCW, nseg segments."""
    t = np.arange(nseg, dtype=np.float64) * (2.0 * math.pi / nseg)
    r = r0 + amp * np.sin(lobes * t)
    verts = np.stack([r * np.cos(t), r * np.sin(t)], 1).astype(np.float32)
    idx = np.arange(nseg, dtype=np.int32)
    prims = np.stack([idx, (idx + 1) % nseg], 1).astype(np.int32)
    return verts, prims


def stroke_soup(nstrokes=60, nseg=1024, lo=30.0, hi=470.0, seed=11):
    """C1 stand-in (data/ladybug/model.obj is a soup of 61 476 open diffusion-curve strokes inside a 500x500 Neumann
    square): `nstrokes` open polylines of `nseg` segments each, smooth random-walk headings, reflected into [lo, hi]^2."""
    rng = np.random.RandomState(seed)
    verts, prims = [], []
    base = 0
    for _ in range(nstrokes):
        length = rng.uniform(150.0, 500.0)
        step = length / nseg
        k = rng.randn(3) * 6.0 / nseg
        ph = rng.rand(3) * 2.0 * math.pi
        i = np.arange(nseg + 1, dtype=np.float64)
        heading = rng.rand() * 2.0 * math.pi + 1.5 * (np.sin(k[0] * i * 3 + ph[0]) + np.sin(k[1] * i * 5 + ph[1]) + 0.5 * np.sin(k[2] * i * 11 + ph[2]))
        xy = np.cumsum(np.stack([np.cos(heading), np.sin(heading)], 1) * step, 0) + rng.uniform(lo + 60.0, hi - 60.0, 2)
        span = hi - lo  # fold (reflect) the curve back into the box
        xy = np.abs(np.mod(xy - lo, 2.0 * span) - span)
        xy = lo + span - xy
        verts.append(xy)
        idx = np.arange(nseg, dtype=np.int32) + base
        prims.append(np.stack([idx, idx + 1], 1))
        base += nseg + 1
    return np.concatenate(verts, 0).astype(np.float32), np.concatenate(prims, 0).astype(np.int32)


def stroke_colors(nstrokes, nseg, seed=7):
    """Piecewise-constant left/right colour pairs per stroke of `stroke_soup` (stand-in for the missing color.json)."""
    table = np.random.RandomState(seed).rand(nstrokes, 6).astype(np.float32)
    return np.repeat(table, nseg + 1, axis=0)


def harmonic_2d(xy, n=3):
    """Re(z^n): harmonic in the plane."""
    z = xy[..., 0].astype(np.float64) + 1j * xy[..., 1].astype(np.float64)
    return np.real(z ** n)


def icosphere(subdiv=3, radius=1.0, displace=0.0, seed=3):
    """Closed triangle mesh with 20*4^subdiv faces (outward orientation), optional smooth radial displacement."""
    t = (1.0 + math.sqrt(5.0)) / 2.0
    v = np.array([[-1, t, 0], [1, t, 0], [-1, -t, 0], [1, -t, 0], [0, -1, t], [0, 1, t], [0, -1, -t], [0, 1, -t],
                  [t, 0, -1], [t, 0, 1], [-t, 0, -1], [-t, 0, 1]], np.float64)
    v /= np.linalg.norm(v, axis=1, keepdims=True)
    f = np.array([[0, 11, 5], [0, 5, 1], [0, 1, 7], [0, 7, 10], [0, 10, 11], [1, 5, 9], [5, 11, 4], [11, 10, 2],
                  [10, 7, 6], [7, 1, 8], [3, 9, 4], [3, 4, 2], [3, 2, 6], [3, 6, 8], [3, 8, 9], [4, 9, 5], [2, 4, 11],
                  [6, 2, 10], [8, 6, 7], [9, 8, 1]], np.int64)
    for _ in range(subdiv):
        a, b, c = f[:, 0], f[:, 1], f[:, 2]
        edges = np.concatenate([np.stack([a, b], 1), np.stack([b, c], 1), np.stack([c, a], 1)], 0)
        es = np.sort(edges, 1)
        key = es[:, 0] * (len(v) + 1) + es[:, 1]
        uniq, first, inv = np.unique(key, return_index=True, return_inverse=True)
        mids = v[es[first, 0]] + v[es[first, 1]]
        mids /= np.linalg.norm(mids, axis=1, keepdims=True)
        base = len(v)
        v = np.concatenate([v, mids], 0)
        n = len(f)
        ab, bc, ca = base + inv[:n], base + inv[n : 2 * n], base + inv[2 * n :]
        f = np.concatenate([np.stack([a, ab, ca], 1), np.stack([b, bc, ab], 1), np.stack([c, ca, bc], 1), np.stack([ab, bc, ca], 1)], 0)
    if displace != 0.0:
        rng = np.random.RandomState(seed)
        k = rng.randn(4, 3)
        ph = rng.rand(4) * 2 * math.pi
        bump = sum(np.sin(3.0 * (v @ k[i]) + ph[i]) for i in range(4)) / 4.0
        v = v * (1.0 + displace * bump)[:, None]
    return (v * radius).astype(np.float32), f.astype(np.int32)


def harmonic_3d(xyz):
    """x^2 + y^2 - 2 z^2: harmonic in space."""
    p = xyz.astype(np.float64)
    return p[..., 0] ** 2 + p[..., 1] ** 2 - 2.0 * p[..., 2] ** 2


def gray_pairs(values):
    """left = right = (v, v, v) colour pairs from a scalar per vertex."""
    v = np.asarray(values, np.float32)
    return np.repeat(v[:, None], 6, axis=1).astype(np.float32)


def reference_data_dir():
    return "/root/reference/data"


def have_reference_data():
    return os.path.isdir(reference_data_dir())
