"""GPU: LBVH build + queries through the C ABI, bit-exact against the oracle and the reference's golden vectors."""
import numpy as np
import pytest

from conftest import GOLDEN_SCENES
from helpers import bits, random_queries

pytestmark = pytest.mark.gpu


def _api():
    from elaina_b200 import api
    return api


@pytest.mark.parametrize("name", GOLDEN_SCENES)
def test_gpu_matches_reference_golden(golden, name):
    api = _api()
    g = {k.split("/", 1)[1]: golden[k] for k in golden.files if k.startswith(name + "/")}
    dim = int(g["dim"])
    s = api.Scene(dim, g["verts"], g["prims"])
    nodes, aabbs, cones = s.dump_bvh()
    assert np.array_equal(nodes, g["nodes"])
    assert np.array_equal(bits(aabbs), bits(g["aabbs"]))
    idx, dist = s.nearest(g["q"])
    assert np.array_equal(idx, g["idx"])
    assert np.array_equal(bits(dist), bits(g["dist"]))
    assert np.array_equal(bits(s.silhouette(g["q"], False)), bits(g["sil0"]))
    assert np.array_equal(bits(s.silhouette(g["q"], True)), bits(g["sil1"]))
    hit, t, uv, obj = s.ray(g["q"], g["dir"], g["max_dist"])
    assert np.array_equal(hit, g["hit"]) and np.array_equal(obj, g["obj"])
    assert np.array_equal(bits(t), bits(g["t"])) and np.array_equal(bits(uv), bits(g["uv"]))
    assert np.array_equal(s.ray_any(g["q"], g["dir"], g["max_dist"]), g["any"])
    sobj, spdf = s.sample_in_sphere(g["q"], g["rad"], g["u"])
    assert np.array_equal(sobj, g["sobj"])
    np.testing.assert_allclose(spdf, g["spdf"], rtol=2e-5)


def _scene_cases():
    from elaina_b200 import scenes
    cases = {}
    cases["star10k"] = (2,) + scenes.star_polygon(10000)
    cases["ico5"] = (3,) + scenes.icosphere(5, 1.0, 0.2)
    rng = np.random.RandomState(9)
    pts = rng.rand(3000, 2).astype(np.float32)
    idx = rng.randint(0, 3000, (5000, 2)).astype(np.int32)
    idx = idx[idx[:, 0] != idx[:, 1]]
    cases["soup2d_random"] = (2, pts, idx)
    tri = rng.rand(2000, 3).astype(np.float32)
    tidx = rng.randint(0, 2000, (3000, 3)).astype(np.int32)
    tidx = tidx[(tidx[:, 0] != tidx[:, 1]) & (tidx[:, 1] != tidx[:, 2]) & (tidx[:, 0] != tidx[:, 2])]
    cases["soup3d_random"] = (3, tri, tidx)
    from elaina_b200 import workloads
    for shipped in ("ladybug", "fille"):   # BASELINE configs[0]/[1] on their REAL geometry: 61 476 / 66 367 segments, 64-bit Morton path
        dv, dp, _, _ = workloads._shipped(shipped)
        cases[shipped] = (2, dv, dp)
    sv, sp = scenes.star_polygon(3000)
    cases["star_x4096"] = (2, (sv * np.float32(4096.0) + np.float32(30000.0)).astype(np.float32), sp)  # a scene authored on a large canvas
    cases["single_segment"] = (2, np.array([[0, 0], [1, 0.5]], np.float32), np.array([[0, 1]], np.int32))
    cases["two_triangles"] = (3, np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [1, 1, 0.3]], np.float32), np.array([[0, 1, 2], [1, 3, 2]], np.int32))
    return cases


@pytest.mark.parametrize("name", ["star10k", "ico5", "soup2d_random", "soup3d_random", "single_segment", "two_triangles", "ladybug", "fille", "star_x4096"])
def test_gpu_matches_oracle_bit_exact(oracle, name):
    """Build order, topology, boxes AND cones bit-exact (one arithmetic), all five query kinds bit-exact."""
    api = _api()
    dim, v, p = _scene_cases()[name]
    so = oracle.scene(dim, v, p)
    sg = api.Scene(dim, v, p)
    no, ao, co = so.dump_bvh()
    ng, ag, cg = sg.dump_bvh()
    assert np.array_equal(no, ng)
    assert np.array_equal(bits(ao), bits(ag))
    valid = co[:, dim] >= 0
    assert np.array_equal(co[:, dim] < 0, cg[:, dim] < 0)
    assert np.array_equal(bits(co[valid]), bits(cg[valid]))
    mo, oo = oracle.dump_sorted(so)
    mg, og = sg.dump_sorted()
    assert np.array_equal(mo, mg) and np.array_equal(oo, og)
    assert oracle.used_morton64(so) == sg.used_morton64
    q, d, md, u, rad = random_queries(v, 20000, 4)
    io, do = so.nearest(q)
    ig, dg = sg.nearest(q)
    assert np.array_equal(io, ig) and np.array_equal(bits(do), bits(dg))
    for flip in (False, True):
        assert np.array_equal(bits(so.silhouette(q, flip)), bits(sg.silhouette(q, flip)))
    ho, to, uo, bo = so.ray(q, d, md)
    hg, tg, ug, bg = sg.ray(q, d, md)
    assert np.array_equal(ho, hg) and np.array_equal(bo, bg)
    assert np.array_equal(bits(to), bits(tg)) and np.array_equal(bits(uo), bits(ug))
    assert np.array_equal(so.ray_any(q, d, md), sg.ray_any(q, d, md))
    po, fo = so.sample_in_sphere(q, rad, u)
    pg, fg = sg.sample_in_sphere(q, rad, u)
    assert np.array_equal(po, pg) and np.array_equal(bits(fo), bits(fg))


def test_empty_and_invalid_inputs():
    api = _api()
    from elaina_b200._lib import WostError
    v = np.array([[0, 0], [1, 0]], np.float32)
    s = api.Scene(2, v, np.array([[0, 1]], np.int32))
    idx, dist = s.nearest(np.zeros((0, 2), np.float32))
    assert len(idx) == 0
    with pytest.raises(WostError):
        api.Scene(2, v, np.array([[0, 5]], np.int32))  # index out of range


def test_full_size_query_properties():
    """BASELINE-size check by properties: 1024^2 closest-point queries on a 10k-segment polygon.
    d(x) is 1-Lipschitz, d == 0 on the vertices, and idx reproduces d by direct evaluation."""
    api = _api()
    from elaina_b200 import scenes
    v, p = scenes.star_polygon(10000)
    s = api.Scene(2, v, p)
    g = np.linspace(-1.5, 1.5, 1024, dtype=np.float32)
    X, Y = np.meshgrid(g, g)
    q = np.stack([X.ravel(), Y.ravel()], 1)
    idx, dist = s.nearest(q)
    D = dist.reshape(1024, 1024)
    h = g[1] - g[0]
    assert np.all(np.abs(np.diff(D, axis=0)) <= h * 1.0001 + 1e-6)
    assert np.all(np.abs(np.diff(D, axis=1)) <= h * 1.0001 + 1e-6)
    a, b = v[p[idx, 0]].astype(np.float64), v[p[idx, 1]].astype(np.float64)
    ab = b - a
    t = np.clip(((q - a) * ab).sum(1) / (ab * ab).sum(1), 0, 1)
    direct = np.linalg.norm(q - (a + t[:, None] * ab), axis=1)
    np.testing.assert_allclose(dist, direct, rtol=1e-4, atol=1e-6)
    _, dv = s.nearest(v)
    assert np.all(dv <= 1e-6)


@pytest.mark.parametrize("name", ["star10k", "ico5", "soup2d_random", "soup3d_random"])
def test_nearest_tie_breaks_and_far_queries(oracle, name):
    """The B200 traversal order must reproduce the reference's winner among EXACT ties (queries on vertices, on edge
    midpoints, on the medial axis of a symmetric shape) and for far-away / huge-coordinate queries."""
    api = _api()
    dim, v, p = _scene_cases()[name]
    so = oracle.scene(dim, v, p)
    sg = api.Scene(dim, v, p)
    rng = np.random.RandomState(5)
    vq = v[rng.randint(0, len(v), 4000)]
    mid = v[p[rng.randint(0, len(p), 4000)]].mean(1).astype(np.float32)
    off = (vq + rng.randn(*vq.shape).astype(np.float32) * 1e-4).astype(np.float32)
    far = (rng.randn(4000, dim) * np.array([10.0, 1e3, 1e5, 1e7])[rng.randint(0, 4, 4000)][:, None]).astype(np.float32)
    centre = np.zeros((1, dim), np.float32) + v.mean(0).astype(np.float32)
    for q in (vq, mid, off, far, centre):
        io, do = so.nearest(q)
        ig, dg = sg.nearest(q)
        assert np.array_equal(io, ig)
        assert np.array_equal(bits(do), bits(dg))


def test_survey_model_obj_rows_on_gpu():
    """SURVEY.md section 8(c), data/ladybug/model.obj rows (values from the host-compiled reference LBVH)."""
    api = _api()
    from elaina_b200 import workloads
    dv, dp, _, _ = workloads._shipped("ladybug")
    s = api.Scene(2, dv, dp)
    assert s.num_nodes == 122951 and s.used_morton64
    q = np.array([(100, 250), (180, 220), (260, 190), (340, 160), (420, 130)], np.float32)
    idx, dist = s.nearest(q)
    assert idx.tolist() == [338, 44560, 18206, 39513, 58322]
    np.testing.assert_array_equal(dist, np.array([1.48958969, 2.60965014, 18.6593742, 5.0058713, 1.48767483], np.float32))
    np.testing.assert_array_equal(s.silhouette(q), np.array([14.9877558, 4.2426405, 19.1049728, 14.5602198, 2.23606801], np.float32))
    hit, t, uv, obj = s.ray(q, np.tile(np.array([0.6, 0.8], np.float32), (5, 1)), np.full(5, 1e9, np.float32))
    assert hit.tolist() == [1] * 5 and obj.tolist() == [9655, 44578, 5780, 37537, 56845]
    np.testing.assert_array_equal(t, np.array([10.078681, 3.18717504, 260.930939, 18.5931568, 57.7269135], np.float32))


def test_large_canvas_scene_stays_on_the_fast_path(oracle):
    """round 1 sent every 2D query with |coordinate| > 2000 to a full traversal in the reference's visit
    order. A scene authored on a 4096-unit canvas (coordinates 25 000-35 000) must be answered by the B200 order -- at most a
    restricted second pass for near-ties -- and still bit-exactly."""
    api = _api()
    dim, v, p = _scene_cases()["star_x4096"]
    so, sg = oracle.scene(dim, v, p), api.Scene(dim, v, p)
    rng = np.random.RandomState(3)
    q = (30000.0 + (rng.rand(20000, 2) * 2 - 1) * 6000.0).astype(np.float32)
    on_vertex = v[rng.randint(0, len(v), 2000)]
    sg.traversal_stats(reset=True)
    for pts in (q, on_vertex):
        io, do = so.nearest(pts)
        ig, dg = sg.nearest(pts)
        assert np.array_equal(io, ig) and np.array_equal(bits(do), bits(dg))
    full, restricted = sg.traversal_stats()
    assert full == 0
    # queries whose closest point is a polygon VERTEX tie exactly between its two segments and take the restricted second pass
    # (the subtree around that vertex, in the reference's order): about 44 % of the queries around this smooth polygon
    assert restricted < 0.6 * (len(q) + len(on_vertex)), restricted
    print("restricted second passes: %d of %d queries" % (restricted, len(q) + len(on_vertex)))


@pytest.mark.gpu
def test_device_edge_adjacency_equals_host_path(monkeypatch):
    """Row N1: the triangle-mesh silhouette adjacency is built on the device (sort by edge key); the host hash-map path
    (the reference's sequential semantics, scene.cuh:1135-1229) is kept behind WOST_HOST_SILHOUETTES as a cross-check.
    Non-manifold soup (edges shared by > 2 triangles, both orientations) + a closed mesh; cones and silhouette distances equal."""
    from elaina_b200 import api, scenes
    rng = np.random.RandomState(5)
    tri = rng.rand(300, 3).astype(np.float32)
    tidx = rng.randint(0, 300, (4000, 3)).astype(np.int32)   # few vertices, many triangles: high edge valence
    tidx = tidx[(tidx[:, 0] != tidx[:, 1]) & (tidx[:, 1] != tidx[:, 2]) & (tidx[:, 0] != tidx[:, 2])]
    for verts, prims in ((tri, tidx), scenes.icosphere(5, 1.0, 0.2)):
        q = (rng.rand(2000, 3).astype(np.float32) * 2.4 - 1.2)
        monkeypatch.delenv("WOST_HOST_SILHOUETTES", raising=False)
        dev = api.Scene(3, verts, prims)
        monkeypatch.setenv("WOST_HOST_SILHOUETTES", "1")
        host = api.Scene(3, verts, prims)
        nd, bd, cd = dev.dump_bvh()
        nh, bh, ch = host.dump_bvh()
        assert np.array_equal(nd, nh) and np.array_equal(bits(bd), bits(bh)) and np.array_equal(bits(cd), bits(ch))
        for flip in (False, True):
            assert np.array_equal(bits(dev.silhouette(q, flip)), bits(host.silhouette(q, flip)))


@pytest.mark.parametrize("name", ["star10k", "ico5", "soup2d_random", "soup3d_random", "two_triangles", "star_x4096"])
def test_oriented_boxes_contain_their_subtrees(name):
    """The traversal records carry one oriented box per child (csrc/lbvh.cuh: Obb) that only prunes: every vertex below a
    child must project inside the child's intervals (fp64 check of the stored fp32 frame), the frame must be orthonormal
    to rounding, and the intervals must be tight (exact extent of the projections + the 2^-20 max|coordinate| widening)."""
    api = _api()
    dim, v, p = _scene_cases()[name]
    s = api.Scene(dim, v, p)
    rec, lv = s.dump_traversal_layout()
    n = len(p)
    assert rec.shape == (n - 1, 24 if dim == 2 else 40)
    refs = rec.view(np.uint32)[:, (8, 9) if dim == 2 else (12, 13)]
    LEAF = np.uint32(0x80000000)
    # leaf range of every child by a post-order walk over the child references
    first = np.zeros((n - 1, 2), np.int64)
    last = np.zeros((n - 1, 2), np.int64)
    order, stack = [], [0]
    while stack:
        i = stack.pop()
        order.append(i)
        for r in refs[i]:
            if not (r & LEAF):
                stack.append(int(r))
    for i in reversed(order):
        for side in range(2):
            r = refs[i, side]
            if r & LEAF:
                first[i, side] = last[i, side] = int(r & ~LEAF)
            else:
                first[i, side], last[i, side] = first[int(r)].min(), last[int(r)].max()
    assert first[0].min() == 0 and last[0].max() == n - 1
    M = float(np.abs(v[np.unique(p)]).max())
    w = M * 2.0 ** -20
    lv64 = lv.astype(np.float64)
    rng = np.random.RandomState(3)
    nodes = np.arange(n - 1) if n - 1 <= 4000 else np.unique(np.concatenate([np.arange(64), rng.randint(0, n - 1, 4000)]))
    for i in nodes:
        for side in range(2):
            o = rec[i, (10 if dim == 2 else 16) + side * (6 if dim == 2 else 12):][: 6 if dim == 2 else 12].astype(np.float64)
            pts = lv64[first[i, side]: last[i, side] + 1].reshape(-1, dim)
            if dim == 2:
                axes = np.array([[o[0], o[1]], [-o[1], o[0]]])
                lo, hi = o[[2, 4]], o[[3, 5]]
            else:
                nrm, t = o[0:3], o[3:6]
                axes = np.stack([nrm, t, np.cross(nrm, t)])
                lo, hi = o[6:9], o[9:12]
            np.testing.assert_allclose(axes @ axes.T, np.eye(dim), atol=2e-6)
            pr = pts @ axes.T
            assert np.all(pr >= lo - 1e-30) and np.all(pr <= hi + 1e-30), (name, i, side)
            # tight: the interval is the extent of the projections plus the widening (and fp32 rounding of the projection)
            slack = w + 4.0 * M * 2.0 ** -23
            assert np.all(pr.min(0) - lo <= 2.0 * slack) and np.all(hi - pr.max(0) <= 2.0 * slack), (name, i, side)


def _degenerate_cases():
    rng = np.random.RandomState(21)
    cases = {}
    # 2D: zero-length segments (both endpoints equal), duplicated segments, everything on one line, coordinates of very different scale
    pts = rng.rand(400, 2).astype(np.float32)
    idx = rng.randint(0, 400, (900, 2)).astype(np.int32)
    idx[:60, 1] = idx[:60, 0]                      # zero-length
    idx[60:120] = idx[120:180]                     # duplicates
    cases["degenerate2d"] = (2, pts, idx)
    line = np.stack([np.linspace(-3, 3, 300), np.full(300, 0.25)], 1).astype(np.float32)
    cases["collinear2d"] = (2, line, np.stack([np.arange(299), np.arange(1, 300)], 1).astype(np.int32))
    big = (rng.rand(500, 2) * np.array([1e5, 1e-2])).astype(np.float32)
    bidx = rng.randint(0, 500, (800, 2)).astype(np.int32)
    cases["anisotropic2d"] = (2, big, bidx[bidx[:, 0] != bidx[:, 1]])
    # 3D: needle and sliver triangles, a flat (coplanar) sheet, duplicated triangles
    v = rng.rand(600, 3).astype(np.float32)
    v[:200, 2] = 0.5                               # coplanar sheet
    t = rng.randint(0, 200, (500, 3)).astype(np.int32)
    t = t[(t[:, 0] != t[:, 1]) & (t[:, 1] != t[:, 2]) & (t[:, 0] != t[:, 2])]
    needle = np.stack([np.arange(200, 590), np.arange(201, 591), np.arange(202, 592)], 1).astype(np.int32)
    v[200:] = np.cumsum(rng.randn(400, 3).astype(np.float32) * np.array([1e-1, 1e-5, 1e-5], np.float32), 0)  # nearly collinear chain -> slivers
    cases["slivers3d"] = (3, v, np.concatenate([t, needle, t[:50]], 0))
    return cases


@pytest.mark.parametrize("name", ["degenerate2d", "collinear2d", "anisotropic2d", "slivers3d"])
def test_degenerate_meshes_match_oracle(oracle, name):
    """Edge cases of the oriented-box records (zero normals, zero-length axes, flat and needle-shaped subtrees, duplicated
    primitives = exact ties everywhere): closest-point and silhouette results stay bit-identical to the oracle, and every
    vertex stays inside the oriented boxes of its ancestors (the build never produces a NaN frame)."""
    api = _api()
    dim, v, p = _degenerate_cases()[name]
    so, sg = oracle.scene(dim, v, p), api.Scene(dim, v, p)
    rec, _ = sg.dump_traversal_layout()
    assert np.isfinite(rec[:, (10 if dim == 2 else 16):(22 if dim == 2 else 40)]).all()
    q, d, md, u, rad = random_queries(v, 20000, 6, margin=0.5)
    io, do = so.nearest(q)
    ig, dg = sg.nearest(q)
    assert np.array_equal(bits(do), bits(dg))
    assert np.array_equal(io, ig)
    for flip in (False, True):
        assert np.array_equal(bits(so.silhouette(q, flip)), bits(sg.silhouette(q, flip)))
    # on the primitives themselves (distance 0, many exact ties)
    on = v[p[:, 0]][:3000]
    io, do = so.nearest(on)
    ig, dg = sg.nearest(on)
    assert np.array_equal(bits(do), bits(dg)) and np.array_equal(io, ig)


def _caterpillar_2d(log2n=17):
    """A polyline soup whose LBVH is as deep as 2^log2n primitives allow. Twenty tiny segments sit in grid cells whose Morton code
    has exactly ONE bit set (cell 2^j on one axis, 0 on the other): sorted, every split peels off one of them -> 20 levels. The
    segments with indices 0, 1, 2, 4, ..., 2^(log2n-1) share cell (0, 0), hence one Morton code, so the 64-bit (code << 32 | index)
    keys continue the chain through the index bits -> log2n more levels. Everything else fills distinct cells of the far quadrant."""
    n = 1 << log2n
    cell = np.zeros((n, 2), np.int64)
    chain_idx = [0] + [1 << k for k in range(log2n)]
    single_bit = [(1 << j, 0) for j in range(10)] + [(0, 1 << j) for j in range(10)]
    free = [i for i in range(3, 3 + 4 * len(single_bit)) if i & (i - 1)][:len(single_bit)]   # indices that are not powers of two
    used = set(chain_idx) | set(free)
    for i, c in zip(free, single_bit):
        cell[i] = c
    rest = np.array([i for i in range(n) if i not in used])
    k = np.arange(len(rest))
    cell[rest, 0] = 512 + k % 512
    cell[rest, 1] = 512 + (k // 512) % 512
    ctr = (cell + 0.5) / 1024.0
    h = 1.0 / 8192.0
    a = ctr + np.array([-h, -h * 0.5])
    b = ctr + np.array([h, h * 0.5])
    # two anchors pin the whole box to [0, 1]^2 so that cell = floor(1024 * centroid) exactly
    a[rest[-1]], b[rest[-1]] = (1.0 - 2 * h, 1.0 - h), (1.0, 1.0)
    a[0], b[0] = (0.0, 0.0), (2 * h, h)
    verts = np.concatenate([a, b], 0).astype(np.float32)
    prims = np.stack([np.arange(n), np.arange(n) + n], 1).astype(np.int32)
    return verts, prims


def _max_leaf_depth(nodes):
    parent = nodes[:, 0].astype(np.int64)
    depth = np.zeros(len(nodes), np.int64)
    n_internal = (len(nodes) - 1) // 2
    # parents are not ordered by index in a Karras tree: relax until stable (depth <= 64 rounds)
    for _ in range(80):
        new = np.where(parent == 0xFFFFFFFF, 0, depth[np.minimum(parent, len(nodes) - 1)] + 1)
        new[0] = 0
        if np.array_equal(new, depth):
            break
        depth = new
    return int(depth[n_internal:].max())


def test_deep_caterpillar_tree_uses_the_spilled_stack_without_overflow(oracle):
    """the traversal stack holds 64 entries in total (the reference's stack[64], core/query.cuh:250) whatever share of it
    lives in shared memory -- 20 levels in the query kernels, 12 in the 2D walker. A degenerate tree (clustered Morton codes +
    duplicated codes) of depth >= 36 forces queries deep into the thread-local spill area behind BOTH shares: closest point,
    silhouette and ray results stay bit-identical to the oracle, no overflow is reported, and a short walker solve (12 shared
    levels + spill) reproduces the oracle's film. (Depths of 57-64 need > 2^26 primitives with this builder: out of a test's reach;
    the capacity arithmetic itself is a constexpr, kStackCapacity in csrc/lbvh.cuh.)"""
    api = _api()
    v, p = _caterpillar_2d(18)
    so, sg = oracle.scene(2, v, p), api.Scene(2, v, p)
    assert sg.used_morton64
    no, _, _ = so.dump_bvh()
    ng, _, _ = sg.dump_bvh()
    assert np.array_equal(no, ng)
    assert _max_leaf_depth(ng) >= 36
    rng = np.random.RandomState(4)
    q = np.concatenate([rng.rand(4000, 2) * 0.02, rng.rand(4000, 2), rng.rand(2000, 2) * 4 - 2]).astype(np.float32)
    io, do = so.nearest(q)
    ig, dg = sg.nearest(q)            # WOST_ERR_OVERFLOW would raise here
    assert np.array_equal(io, ig) and np.array_equal(bits(do), bits(dg))
    for flip in (False, True):
        assert np.array_equal(bits(so.silhouette(q, flip)), bits(sg.silhouette(q, flip)))
    d = rng.randn(len(q), 2).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    md = np.full(len(q), 3.0, np.float32)
    ho, to, uvo, oo = so.ray(q, d, md)
    hg, tg, uvg, og = sg.ray(q, d, md)
    assert np.array_equal(ho, hg) and np.array_equal(oo, og) and np.array_equal(bits(to), bits(tg))
    # the 2D persistent walker keeps 12 levels in shared memory: evaluation points right on top of the deep chain
    from elaina_b200 import scenes
    kw = dict(color_dirichlet=scenes.gray_pairs(rng.rand(len(v)).astype(np.float32)), color_neumann=None, grid_scale=0.01, grid_pos=(0.01, 0.01), grid_up=(0.0, 1.0),
              aabb_min=(-0.5, -0.5), aabb_max=(1.5, 1.5))
    oprob = dict(dim=2, dirichlet=so, neumann=None, **kw)
    gprob = api.Problem(2, dirichlet=sg, neumann=None, **kw)
    st = dict(frame_w=24, frame_h=24, max_walking_depth=12, samples_per_pixel=2, epsilon_shell=1e-4, seed=3)
    sol_o, steps_o = oracle.solve_uniform(oprob, st)
    it = api.UniformIntegrator(gprob, api.make_settings((24, 24), 12, 2, 1e-4, seed=3, spp_batch=1))
    it.solve()                        # raises on a stack overflow
    assert it.last_walk_steps == steps_o
    assert np.array_equal(bits(it.film()[..., :3].reshape(-1, 3)), bits(sol_o))
