"""Design prototype (CPU, statistics only): how many node opens / leaf evaluations does a nearest-first closest-point
query take on the LBVH with (a) AABB pruning only and (b) AABB + one oriented slab per child (mean normal of the subtree)?
Queries are the steps >= 1 of walk-on-spheres walks from random interior points (the first step is cached per pixel).

  python tools/proto_slab.py c3|c4-small [nwalks]
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from elaina_b200.workloads import workload  # noqa: E402
from oracle_lib import Oracle  # noqa: E402


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "c3"
    nwalks = int(sys.argv[2]) if len(sys.argv) > 2 else 200
    wl = workload(name)
    D = wl["dim"]
    orc = Oracle()
    sc = orc.scene(D, wl["dverts"], wl["dprims"])
    nodes, aabbs, _ = sc.dump_bvh()
    N = len(wl["dprims"])
    V = wl["dverts"].astype(np.float64)
    P = wl["dprims"]
    left, right, obj = nodes[:, 1].astype(np.int64), nodes[:, 2].astype(np.int64), nodes[:, 3].astype(np.int64)
    nn = len(nodes)
    is_leaf = np.arange(nn) >= N - 1
    up, lo = aabbs[:, :D].astype(np.float64), aabbs[:, D:].astype(np.float64)
    # primitive normals (unnormalised = weighted by length / 2*area)
    pv = V[P]  # (N, D, D)
    if D == 2:
        e = pv[:, 1] - pv[:, 0]
        pn = np.stack([-e[:, 1], e[:, 0]], 1)
    else:
        pn = np.cross(pv[:, 1] - pv[:, 0], pv[:, 2] - pv[:, 0])
    # post-order: leaf ranges + summed normals
    order = []
    stack = [0]
    while stack:
        n = stack.pop()
        order.append(n)
        if not is_leaf[n]:
            stack.append(left[n])
            stack.append(right[n])
    first = np.zeros(nn, np.int64)
    last = np.zeros(nn, np.int64)
    nsum = np.zeros((nn, D))
    for n in reversed(order):
        if is_leaf[n]:
            first[n] = last[n] = n - (N - 1)
            nsum[n] = pn[obj[n]]
        else:
            l, r = left[n], right[n]
            first[n], last[n] = min(first[l], first[r]), max(last[l], last[r])
            a, b = nsum[l], nsum[r]
            nsum[n] = a + (b if np.dot(a, b) >= 0 else -b)
    leaf_obj = obj[N - 1:]
    leaf_verts = pv[leaf_obj]  # (N, D, D) in Morton order
    nrm = nsum / np.maximum(np.linalg.norm(nsum, axis=1, keepdims=True), 1e-300)
    slo = np.zeros(nn)
    shi = np.zeros(nn)
    OBB = os.environ.get("OBB", "0") == "1"
    # tangent frame: D-1 extra axes orthogonal to the normal
    tax = np.zeros((nn, D - 1, D))
    tlo = np.zeros((nn, D - 1))
    thi = np.zeros((nn, D - 1))
    for n in range(nn):
        vs = leaf_verts[first[n]:last[n] + 1].reshape(-1, D)
        pr = vs @ nrm[n]
        slo[n], shi[n] = pr.min(), pr.max()
        if OBB:
            if D == 2:
                tax[n, 0] = [-nrm[n, 1], nrm[n, 0]]
            else:
                a = np.eye(3)[np.argmin(np.abs(nrm[n]))]
                t1 = np.cross(nrm[n], a); t1 /= np.linalg.norm(t1)
                tax[n, 0] = t1
                tax[n, 1] = np.cross(nrm[n], t1)
            pt = vs @ tax[n].T
            tlo[n], thi[n] = pt.min(0), pt.max(0)
    print("tree: %d prims, slabs built" % N, flush=True)

    def box_d2(n, q):
        d = np.minimum(up[n], np.maximum(lo[n], q)) - q
        return float(d @ d)

    def slab_d2(n, q):
        t = float(nrm[n] @ q)
        d = max(0.0, t - shi[n], slo[n] - t)
        r = d * d
        if OBB:
            tt = tax[n] @ q
            dd = np.maximum(0.0, np.maximum(tt - thi[n], tlo[n] - tt))
            r += float(dd @ dd)
        return r

    def prim_d(pos, q):
        p = leaf_verts[pos]
        if D == 2:
            a, b = p[0], p[1]
            u, v = b - a, q - a
            c1 = u @ v
            if c1 <= 0:
                return np.linalg.norm(q - a)
            c2 = u @ u
            if c2 <= c1:
                return np.linalg.norm(q - b)
            return np.linalg.norm(q - (a + u * (c1 / c2)))
        a, b, c = p
        ab, ac, ap = b - a, c - a, q - a
        d1, d2 = ab @ ap, ac @ ap
        if d1 <= 0 and d2 <= 0:
            return np.linalg.norm(ap)
        bp = q - b
        d3, d4 = ab @ bp, ac @ bp
        if d3 >= 0 and d4 <= d3:
            return np.linalg.norm(bp)
        vc = d1 * d4 - d3 * d2
        if vc <= 0 and d1 >= 0 and d3 <= 0:
            return np.linalg.norm(q - (a + ab * (d1 / (d1 - d3))))
        cp = q - c
        d5, d6 = ab @ cp, ac @ cp
        if d6 >= 0 and d5 <= d6:
            return np.linalg.norm(cp)
        vb = d5 * d2 - d1 * d6
        if vb <= 0 and d2 >= 0 and d6 <= 0:
            return np.linalg.norm(q - (a + ac * (d2 / (d2 - d6))))
        va = d3 * d6 - d5 * d4
        if va <= 0 and (d4 - d3) >= 0 and (d5 - d6) >= 0:
            w = (d4 - d3) / ((d4 - d3) + (d5 - d6))
            return np.linalg.norm(q - (b + (c - b) * w))
        den = 1.0 / (va + vb + vc)
        return np.linalg.norm(q - (a + ab * (vb * den) + ac * (vc * den)))

    def query(q, bound, use_slab):
        best, win = bound, -1
        opens = evals = 0
        st = []
        cur = 0
        while True:
            opens += 1
            kids = []
            for ch in (left[cur], right[cur]):
                d2 = box_d2(ch, q)
                if use_slab:
                    d2 = max(d2, slab_d2(ch, q))
                kids.append((d2, ch))
            kids.sort()
            nxt = []
            for d2, ch in kids:
                if d2 > best * best:
                    continue
                if is_leaf[ch]:
                    evals += 1
                    d = prim_d(ch - (N - 1), q)
                    if d <= best:
                        best, win = d, ch - (N - 1)
                else:
                    nxt.append((d2, ch))
            nxt = [(d2, ch) for d2, ch in nxt if d2 <= best * best]
            if len(nxt) == 2:
                st.append(nxt[1])
                cur = nxt[0][1]
                continue
            if len(nxt) == 1:
                cur = nxt[0][1]
                continue
            cur = -1
            while st:
                d2, ch = st.pop()
                if d2 <= best * best:
                    cur = ch
                    break
            if cur < 0:
                return best, win, opens, evals

    def query_deferred(q, bound, K):
        """nearest-first traversal with OBB pruning where leaf distances are evaluated only every K trips (warp-level leaf rounds)
        or when the lane has nothing else to do; returns (best, opens, evals, idle_trips)"""
        best = bound
        opens = evals = idle = 0
        st = []
        cur = 0
        pend = []
        trip = 0
        while True:
            trip += 1
            if trip % K == 0 and pend:
                for pos in pend:
                    evals += 1
                    d = prim_d(pos, q)
                    if d <= best:
                        best = d
                pend = []
            if cur < 0:
                # pop
                while st:
                    d2, ch = st.pop()
                    if d2 <= best * best:
                        cur = ch
                        break
                if cur < 0:
                    if pend:
                        idle += 1   # drained: wait for the next leaf round
                        continue
                    return best, opens, evals, idle
            opens += 1
            kids = []
            for ch in (left[cur], right[cur]):
                d2 = max(box_d2(ch, q), slab_d2(ch, q))
                kids.append((d2, ch))
            kids.sort()
            nxt = []
            for d2, ch in kids:
                if d2 > best * best:
                    continue
                if is_leaf[ch]:
                    pend.append(ch - (N - 1))
                else:
                    nxt.append((d2, ch))
            if len(nxt) == 2:
                st.append(nxt[1])
                cur = nxt[0][1]
            elif len(nxt) == 1:
                cur = nxt[0][1]
            else:
                cur = -1

    rng = np.random.RandomState(5)
    W, H = wl["frame"]
    mask = wl.get("mask")
    pix = np.flatnonzero(mask.reshape(-1)) if mask is not None else np.arange(W * H)
    eps = wl["eps"]
    tot = {False: [0, 0], True: [0, 0]}
    DEFER = [1, 2, 3, 4, 6] if os.environ.get('OBB') == '1' else []
    dstat = {}
    nq = 0
    hist = []
    for w in range(nwalks):
        p = pix[rng.randint(len(pix))]
        px, py = p % W, p // W
        s = wl["grid_scale"]
        if D == 2:
            x = np.array([wl["grid_pos"][0] + s * (2.0 * (px + 0.5) / W - 1.0), wl["grid_pos"][1] + s * (2.0 * (py + 0.5) / H - 1.0)])
        else:
            x = np.array([s * (2.0 * (px + 0.5) / W - 1.0), s * (2.0 * (py + 0.5) / H - 1.0), 0.0])
        R, win, _, _ = query(x, np.inf, False)
        for depth in range(wl["depth"]):
            if R < eps:
                break
            v = rng.randn(D)
            v /= np.linalg.norm(v)
            x = x + R * v
            bound = prim_d(win, x) * 1.00001
            r0 = None
            for use_slab in (False, True):
                r, wn, o, e = query(x, bound, use_slab)
                tot[use_slab][0] += o
                tot[use_slab][1] += e
                if r0 is None:
                    r0 = (r, wn, o, e)
                else:
                    assert abs(r - r0[0]) < 1e-9, (r, r0)
                    hist.append((r0[0], r0[2], r0[3], o, e))
            for K in DEFER:
                rb, o_, e_, i_ = query_deferred(x, bound, K)
                assert abs(rb - r0[0]) < 1e-9
                dstat.setdefault(K, [0, 0, 0])
                dstat[K][0] += o_; dstat[K][1] += e_; dstat[K][2] += i_
            R, win = r0[0], r0[1]
            nq += 1
    print("%s: %d queries" % (name, nq))
    for k in (False, True):
        print("  slab=%s: node opens/query %.1f, leaf evals/query %.1f" % (k, tot[k][0] / nq, tot[k][1] / nq))
    for K, v in dstat.items():
        print('  deferred leaves, round every %d trips: opens/query %.1f evals/query %.1f idle trips/query %.2f' % (K, v[0] / nq, v[1] / nq, v[2] / nq))
    h = np.array(hist)
    qs = np.quantile(h[:, 0], [0.25, 0.5, 0.75])
    for a, b in ((0, qs[0]), (qs[0], qs[1]), (qs[1], qs[2]), (qs[2], 1e30)):
        m = (h[:, 0] >= a) & (h[:, 0] < b)
        print("  d in [%.3g, %.3g): AABB opens %.1f evals %.1f | +slab opens %.1f evals %.1f" % (a, b, h[m, 1].mean(), h[m, 2].mean(), h[m, 3].mean(), h[m, 4].mean()))


if __name__ == "__main__":
    main()
