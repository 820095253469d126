// SNCH-LBVH device representation and traversal (sm_100a).
//
// What is kept from the reference (ext/lbvh/include/snch_lbvh): the tree itself -- Morton order, Karras
// topology, boxes, cones -- and the exact visit order of each query, so that indices and distances are
// bit-identical (ties at shared polyline vertices are resolved by visit order, core/query.cuh:279-314).
// What is re-designed for B200:
//   * one fused record per INTERNAL node holding both children's boxes (and cones, for silhouette queries)
//     plus both child references: a visit is 3-6 independent 128-bit loads from ONE 48..96 B record instead of
//     the reference's 5-7 dependent scattered loads (node 16 B -> 2 child boxes -> 2 child nodes -> object -> vertices);
//   * leaves reference Morton-ordered inline primitive / silhouette records (no index indirection);
//   * the traversal stack lives in shared memory, one bank-conflict-free column per lane ([level][lane]),
//     not in a 512 B local-memory array (core/query.cuh:250).
#pragma once
#include "geom.cuh"

namespace wost
{
    // The boxes inside TNode are the EXACT extents of the children (no padding). The reference's padded boxes are
    // rebuilt per visit with one add per coordinate (pad_box): bit-identical to its leaf boxes and to its merged
    // internal boxes because x -> fl(x + c) is monotone, so max_i fl(a_i + c) == fl(max_i a_i + c).
    template <int D>
    struct TNode; // nearest / ray / sphere-descent record
    // Each child also carries an ORIENTED box (Obb): the frame follows the mean normal of the primitives below the child
    // (2D: normal n and tangent (-n.y, n.x); 3D: n, an in-plane axis t, and n x t), with the exact interval of the
    // subtree's vertices along every axis, widened by 2^-20 max|coordinate| against rounding. It is a pruning bound only:
    // a smooth curve / surface patch seen from a distance d has an AABB that reaches much closer to the query than any of
    // its primitives (the AABB of a diagonal patch of size L protrudes by ~L/2), so on tessellated boundaries a closest-
    // point query opens 3-8x more nodes and evaluates 15-40x more leaves than it has to (tools/proto_slab.py: C3 57 -> 19
    // node opens and 36 -> 2.4 leaf distances per query; c4-small 302 -> 40 and 128 -> 3). The AABB extents stay because the
    // reference's sibling rule is defined on them (nearest_trip); results are unchanged bit for bit -- the oriented box only
    // removes work whose result could not have been used.
    template <int D>
    struct Obb;
    template <>
    struct Obb<2>
    {
        float ax[2];       // unit normal n; the tangent is (-n.y, n.x)
        float lo[2], hi[2]; // intervals along n and t
    };
    template <>
    struct Obb<3>
    {
        float n[3], t[3];   // unit normal and unit in-plane axis; the third axis is n x t
        float lo[3], hi[3]; // intervals along n, t, n x t
    };
    template <>
    struct __align__(16) TNode<2>
    {
        float lb[4]; // up.x up.y lo.x lo.y
        float rb[4];
        uint32_t lref, rref; // child refs: internal index, or leaf position | kLeafFlag
        float lo6[6];        // Obb<2> of the left child: ax.x ax.y lo.n hi.n lo.t hi.t
        float ro6[6];
        float pad[2];
    };
    template <>
    struct __align__(16) TNode<3>
    {
        float lb[6]; // up.xyz lo.xyz
        float rb[6];
        uint32_t lref, rref, pad0, pad1;
        float lo12[12]; // Obb<3> of the left child: n.xyz t.xyz lo.ntb hi.ntb
        float ro12[12];
    };
    static_assert(sizeof(TNode<2>) == 96 && sizeof(TNode<3>) == 160, "TNode layout");

    template <int D>
    struct CNode; // cones of both children (silhouette queries read TNode + CNode)
    template <>
    struct __align__(16) CNode<2>
    {
        float lc[4]; // axis.x axis.y half_angle radius
        float rc[4];
    };
    template <>
    struct __align__(16) CNode<3>
    {
        float lc[5]; // axis.xyz half_angle radius
        float rc[5];
        float pad[2];
    };
    static_assert(sizeof(CNode<2>) == 32 && sizeof(CNode<3>) == 48, "CNode layout");

    template <int D>
    struct LeafPrim;
    template <>
    struct __align__(16) LeafPrim<2>
    {
        float p[4]; // pa.xy pb.xy
    };
    template <>
    struct __align__(16) LeafPrim<3>
    {
        float p[9]; // pa pb pc
        uint32_t obj;
        float pad[2];
    };
    static_assert(sizeof(LeafPrim<2>) == 16 && sizeof(LeafPrim<3>) == 48, "LeafPrim layout");

    template <int D>
    struct LeafSil;
    template <>
    struct __align__(16) LeafSil<2>
    {
        float p[2], n0[2], n1[2];
        uint32_t flags, pad;
    };
    template <>
    struct __align__(16) LeafSil<3>
    {
        float pa[3], pb[3], n0[3], n1[3];
        uint32_t flags, pad[3];
    };
    static_assert(sizeof(LeafSil<2>) == 32 && sizeof(LeafSil<3>) == 64, "LeafSil layout");

    template <int D>
    struct DevBVH
    {
        int nprims;
        const TNode<D> *tnodes;     // nprims-1
        const CNode<D> *cnodes;     // nprims-1
        const LeafPrim<D> *leaf_prim; // nprims (Morton order)
        const uint32_t *leaf_obj;   // nprims: original object index of leaf position
        const uint32_t *leaf_rank;  // nprims: visit rank of the leaf in the reference's closest-point order (tie-break)
        float max_abs_coord;        // max |coordinate| of the mesh (rounding slack of the closest-point pruning)
        const LeafSil<D> *leaf_sil; // nprims*D
        float root_box[2 * D];      // up, lo
        float root_cone[D + 2];
    };

    // ---------------------------------------------------------------- 128-bit loads
    WOST_DEV float4 ldg4(const void *p) { return __ldg(reinterpret_cast<const float4 *>(p)); }

    template <int D>
    struct NodeView
    {
        Box<D> lb, rb; // the reference's (padded) boxes
        uint32_t lref, rref;
    };
    template <int D>
    struct NodeViewE
    {
        Box<D> le, re; // exact extents
        uint32_t lref, rref;
    };
    // scene.cuh:429-430 (2D: +-bvh_offset) / aabb.cuh:21-50,93-101 (3D: +-epsilon<float>())
    template <int D>
    WOST_DEV Box<D> pad_box(const Box<D> &e)
    {
        const float pad = (D == 2) ? kBvhOffset : kEps;
        Box<D> b;
#pragma unroll
        for (int k = 0; k < D; ++k)
        {
            b.up[k] = e.up[k] + pad;
            b.lo[k] = e.lo[k] - pad;
        }
        return b;
    }
    WOST_DEV NodeViewE<2> load_tnode_e(const TNode<2> *n)
    {
        const float4 a = ldg4(n), b = ldg4(reinterpret_cast<const char *>(n) + 16), c = ldg4(reinterpret_cast<const char *>(n) + 32);
        NodeViewE<2> v;
        v.le.up[0] = a.x, v.le.up[1] = a.y, v.le.lo[0] = a.z, v.le.lo[1] = a.w;
        v.re.up[0] = b.x, v.re.up[1] = b.y, v.re.lo[0] = b.z, v.re.lo[1] = b.w;
        v.lref = __float_as_uint(c.x);
        v.rref = __float_as_uint(c.y);
        return v;
    }
    WOST_DEV NodeViewE<3> load_tnode_e(const TNode<3> *n)
    {
        const char *q = reinterpret_cast<const char *>(n);
        const float4 a = ldg4(q), b = ldg4(q + 16), c = ldg4(q + 32), d = ldg4(q + 48);
        NodeViewE<3> v;
        v.le.up[0] = a.x, v.le.up[1] = a.y, v.le.up[2] = a.z, v.le.lo[0] = a.w;
        v.le.lo[1] = b.x, v.le.lo[2] = b.y, v.re.up[0] = b.z, v.re.up[1] = b.w;
        v.re.up[2] = c.x, v.re.lo[0] = c.y, v.re.lo[1] = c.z, v.re.lo[2] = c.w;
        v.lref = __float_as_uint(d.x);
        v.rref = __float_as_uint(d.y);
        return v;
    }
    // ---- oriented boxes (pruning bound of the closest-point query)
    WOST_HOSTDEV void obb_project(const Obb<2> &o, const Vec<2> &p, float pr[2])
    {
        pr[0] = o.ax[0] * p[0] + o.ax[1] * p[1];
        pr[1] = o.ax[0] * p[1] - o.ax[1] * p[0];
    }
    WOST_HOSTDEV void obb_project(const Obb<3> &o, const Vec<3> &p, float pr[3])
    {
        const float bx = o.n[1] * o.t[2] - o.n[2] * o.t[1], by = o.n[2] * o.t[0] - o.n[0] * o.t[2], bz = o.n[0] * o.t[1] - o.n[1] * o.t[0];
        pr[0] = o.n[0] * p[0] + o.n[1] * p[1] + o.n[2] * p[2];
        pr[1] = o.t[0] * p[0] + o.t[1] * p[1] + o.t[2] * p[2];
        pr[2] = bx * p[0] + by * p[1] + bz * p[2];
    }
    // squared distance from q to the oriented box, scaled down by 2^-18 (the frame is orthonormal to ~2^-23 only): a lower
    // bound of the fp32 distance to every primitive below the child, up to the rounding the caller's slack covers
    template <int D>
    WOST_HOSTDEV float obb_mindist2(const Obb<D> &o, const Vec<D> &q)
    {
        float pr[D];
        obb_project(o, q, pr);
        float s = 0.0f;
#pragma unroll
        for (int k = 0; k < D; ++k)
        {
            const float d = fmax_nn(fmax_nn(pr[k] - o.hi[k], o.lo[k] - pr[k]), 0.0f);
            s = (k == 0) ? d * d : s + d * d;
        }
        return s * 0.99999618530273438f; // 1 - 2^-18
    }
    WOST_DEV Obb<2> load_obb(const TNode<2> *n, int side)
    {
        const char *q = reinterpret_cast<const char *>(n);
        Obb<2> o;
        if (side == 0)
        {
            const float4 c = ldg4(q + 32), d = ldg4(q + 48);
            o.ax[0] = c.z, o.ax[1] = c.w, o.lo[0] = d.x, o.hi[0] = d.y, o.lo[1] = d.z, o.hi[1] = d.w;
        }
        else
        {
            const float4 e = ldg4(q + 64), f = ldg4(q + 80);
            o.ax[0] = e.x, o.ax[1] = e.y, o.lo[0] = e.z, o.hi[0] = e.w, o.lo[1] = f.x, o.hi[1] = f.y;
        }
        return o;
    }
    WOST_DEV Obb<3> load_obb(const TNode<3> *n, int side)
    {
        const char *q = reinterpret_cast<const char *>(n) + 64 + side * 48;
        const float4 a = ldg4(q), b = ldg4(q + 16), c = ldg4(q + 32);
        Obb<3> o;
        o.n[0] = a.x, o.n[1] = a.y, o.n[2] = a.z, o.t[0] = a.w;
        o.t[1] = b.x, o.t[2] = b.y, o.lo[0] = b.z, o.lo[1] = b.w;
        o.lo[2] = c.x, o.hi[0] = c.y, o.hi[1] = c.z, o.hi[2] = c.w;
        return o;
    }
    template <int D>
    WOST_DEV NodeView<D> load_tnode(const TNode<D> *n)
    {
        const NodeViewE<D> e = load_tnode_e(n);
        NodeView<D> v;
        v.lb = pad_box(e.le);
        v.rb = pad_box(e.re);
        v.lref = e.lref;
        v.rref = e.rref;
        return v;
    }
    WOST_DEV void load_cnode(const CNode<2> *n, Cone<2> &l, Cone<2> &r)
    {
        const float4 a = ldg4(n), b = ldg4(reinterpret_cast<const char *>(n) + 16);
        l.axis[0] = a.x, l.axis[1] = a.y, l.half_angle = a.z, l.radius = a.w;
        r.axis[0] = b.x, r.axis[1] = b.y, r.half_angle = b.z, r.radius = b.w;
    }
    WOST_DEV void load_cnode(const CNode<3> *n, Cone<3> &l, Cone<3> &r)
    {
        const char *q = reinterpret_cast<const char *>(n);
        const float4 a = ldg4(q), b = ldg4(q + 16), c = ldg4(q + 32);
        l.axis[0] = a.x, l.axis[1] = a.y, l.axis[2] = a.z, l.half_angle = a.w;
        l.radius = b.x, r.axis[0] = b.y, r.axis[1] = b.z, r.axis[2] = b.w;
        r.half_angle = c.x, r.radius = c.y;
    }
    WOST_DEV Prim<2> load_prim(const LeafPrim<2> *p)
    {
        const float4 a = ldg4(p);
        Prim<2> r;
        r.p[0][0] = a.x, r.p[0][1] = a.y, r.p[1][0] = a.z, r.p[1][1] = a.w;
        return r;
    }
    WOST_DEV Prim<3> load_prim(const LeafPrim<3> *p)
    {
        const char *q = reinterpret_cast<const char *>(p);
        const float4 a = ldg4(q), b = ldg4(q + 16), c = ldg4(q + 32);
        Prim<3> r;
        r.p[0][0] = a.x, r.p[0][1] = a.y, r.p[0][2] = a.z, r.p[1][0] = a.w;
        r.p[1][1] = b.x, r.p[1][2] = b.y, r.p[2][0] = b.z, r.p[2][1] = b.w;
        r.p[2][2] = c.x;
        return r;
    }
    WOST_DEV SilRec<2> load_sil(const LeafSil<2> *p)
    {
        const float4 a = ldg4(p), b = ldg4(reinterpret_cast<const char *>(p) + 16);
        SilRec<2> s;
        s.p[0] = a.x, s.p[1] = a.y, s.n0[0] = a.z, s.n0[1] = a.w;
        s.n1[0] = b.x, s.n1[1] = b.y;
        s.flags = __float_as_uint(b.z);
        return s;
    }
    WOST_DEV SilRec<3> load_sil(const LeafSil<3> *p)
    {
        const char *q = reinterpret_cast<const char *>(p);
        const float4 a = ldg4(q), b = ldg4(q + 16), c = ldg4(q + 32), d = ldg4(q + 48);
        SilRec<3> s;
        s.pa[0] = a.x, s.pa[1] = a.y, s.pa[2] = a.z, s.pb[0] = a.w;
        s.pb[1] = b.x, s.pb[2] = b.y, s.n0[0] = b.z, s.n0[1] = b.w;
        s.n0[2] = c.x, s.n1[0] = c.y, s.n1[1] = c.z, s.n1[2] = c.w;
        s.flags = __float_as_uint(d.x);
        return s;
    }
    template <int D>
    WOST_DEV Box<D> root_box_of(const DevBVH<D> &b)
    {
        Box<D> r;
#pragma unroll
        for (int k = 0; k < D; ++k)
        {
            r.up[k] = b.root_box[k];
            r.lo[k] = b.root_box[D + k];
        }
        return r;
    }

    // ---------------------------------------------------------------- traversal stack
    // Shared-memory stack, one column per thread: entry(level, tid) = base[level * blockDim + tid].
    // Consecutive lanes hit consecutive 8-byte slots -> conflict-free. The first kSmemLevels levels live in shared memory, deeper ones
    // (real scenes need <= 26 in total) spill to local memory; beyond 64 (the reference's limit) a flag is raised.
    // The TOTAL is 64 entries whatever share of it a kernel keeps in shared memory (20, or 12 in the 2D walker and the bounce
    // kernel): the spill area is sized for the smallest share any kernel may ask for.
    constexpr int kStackCapacity = 64; // the reference's stack[64] (core/query.cuh:250)
    constexpr int kSmemLevels = 20;    // short stack in shared memory (20 KB per 128-thread CTA: 11 CTAs per SM)
    constexpr int kMinSmemLevels = 8;  // fewest shared-memory levels a kernel may configure (WOST_STACK_LEVELS is clamped to it)
    constexpr int kSpillLevels = kStackCapacity - kMinSmemLevels; // thread-local entries behind the shared-memory levels
    struct StackEntry
    {
        uint32_t ref;
        float d;
    };
    // cold path of the stack, kept out of line so that the hot loop carries no local-memory traffic
    static __device__ __noinline__ void spill_store(uint2 *spill, int i, uint2 e) { spill[i] = e; }
    static __device__ __noinline__ uint2 spill_load(const uint2 *spill, int i) { return spill[i]; }
    struct SmemStack
    {
        uint2 *base; // column of this thread (already offset by tid)
        int stride;  // blockDim.x
        int sp;
        int levels; // entries of this stack that live in shared memory (kSmemLevels unless the kernel asked for fewer)
        int *overflow; // [0] stack overflow flag, [1] full reference-order traversals, [2] restricted second passes (statistics)
        uint2 *spill; // kSpillLevels entries of thread-local memory owned by the kernel (kept outside this struct so that the
                      // struct itself -- sp, base -- stays in registers); entries [0, kStackCapacity - levels) are used
        WOST_DEV void init(uint2 *smem, int *overflow_flag, uint2 *spill_mem, int smem_levels = kSmemLevels)
        {
            levels = smem_levels; // callers pass kMinSmemLevels..kSmemLevels
            spill = spill_mem;
            base = smem + threadIdx.x;
            stride = blockDim.x;
            sp = 0;
            overflow = overflow_flag;
        }
        WOST_DEV void push(uint32_t ref, float d)
        {
            const uint2 e = make_uint2(ref, __float_as_uint(d));
            if (__builtin_expect(sp < levels, 1))
                base[sp * stride] = e;
            else if (sp < kStackCapacity)
                spill_store(spill, sp - levels, e);
            else
            {
                if (overflow)
                    *overflow = 1;
                return;
            }
            ++sp;
        }
        WOST_DEV StackEntry pop()
        {
            --sp;
            uint2 e;
            if (__builtin_expect(sp < levels, 1))
                e = base[sp * stride];
            else
                e = spill_load(spill, sp - levels);
            StackEntry r;
            r.ref = e.x;
            r.d = __uint_as_float(e.y);
            return r;
        }
        WOST_DEV bool empty() const { return sp <= 0; }
    };
    inline size_t stack_smem_bytes(int block_threads, int levels = kSmemLevels) { return (size_t)block_threads * levels * sizeof(uint2); }

    // ---------------------------------------------------------------- closest primitive (core/query.cuh:238-318)
    // The closest-point traversal is RESUMABLE: nearest_begin() sets up the per-lane state, nearest_trip() opens one
    // node or evaluates one leaf, and the caller owns the loop. Two kinds of callers:
    //   * query_nearest / query_nearest_ref: warp-synchronous drivers, `while (__any_sync(full, active)) trip();`
    //     -- the warp reconverges once per trip instead of wherever the compiler would place the reconvergence point of
    //     a per-thread loop (a divergent `continue` left 4-8 of 32 lanes active, profiles/r1_notes.md);
    //   * the persistent walk kernel (integrator.cu): lanes whose query has finished run the rest of their walk step
    //     and start the next query while the other lanes keep traversing.
    //
    // Phase 1 = the reference's visit order (core/query.cuh:238-318): L handled before R, leaf children evaluated at
    // once, internal children pushed L then R (R popped first); "<=" keeps the last visited of exact ties; a popped
    // node is skipped when its padded mindist exceeds the running best. With thrW < inf it is "restricted": children
    // whose EXACT-extent mindist exceeds thrW are skipped too -- the second pass of the B200 order, see below.
    //
    // Phase 0 = the B200 order. Same RESULT as phase 1 alone (index and distance bit for bit), ~5x fewer visits on
    // curve-like meshes. It finds the minimum with an order the reference does not use:
    //   * the reference's sibling rule "mindist(child) <= minmaxdist(sibling)" (query.cuh:279,299) is evaluated on the
    //     rebuilt padded boxes exactly as the reference does -- it is local to a node, hence independent of visit order,
    //     and it defines the candidate set S of leaves the reference can ever reach;
    //   * pruning by the running best is free to be aggressive: children (internal nodes and leaves alike) are visited
    //     nearest first and dropped as soon as the mindist of their EXACT extent exceeds (sqrt(best) + slack)^2 (the
    //     reference only tests internal nodes, at pop time, on padded boxes, in fixed L/R order). slack = 2^-19 * max
    //     |coordinate| (>= 16 ulp) covers the rounding of the leaf distance itself, so no (near-)minimal leaf is dropped;
    //   * the sqrt + re-square of the reference (dist = length(..); dist *= dist) is only done for leaves whose squared
    //     length can still win (x -> fl(fl(sqrt x)^2) is monotone and within 3 ulp of x);
    //   * one stack entry type for nodes and leaves: per trip a lane either opens a node or evaluates a leaf.
    // Which leaf the reference returns among EXACT ties (shared vertices / edges):
    //   * 2D (boxes padded by 1e-3, scene.cuh:12): the reference's pop-time test can never drop a tied minimum, so it
    //     returns the last visited one of its static order == the highest leaf_rank. No second pass.
    //   * 3D (boxes padded by one float epsilon, aabb.cuh:21-50): the pop-time test compares the padded mindist with
    //     the re-squared best and drops tied / 1-ulp-near minima depending on visit order. Whenever phase 0 saw more
    //     than one leaf within 8 ulp of the minimum, the lane continues in phase 1 restricted to thrW = the final
    //     pruning threshold: only the subtrees holding the (near-)tied minima are re-walked, in the reference's order.
    //   * 2D far field (|q| > kFar2D, where the 1e-3 padding no longer dominates rounding): B200 order + near-tie count +
    //     restricted pass, like 3D ("phase 2"). 3D far field (|q| beyond ~1000x the mesh): phase 1 from the start.
    constexpr uint32_t kFullMask = 0xffffffffu;
    constexpr float kFar2D = 2000.0f; // 1e-3 (scene.cuh:12 bvh_offset) / (8 * 2^-24), rounded down: see nearest_begin

    template <int D>
    struct NearestTrav
    {
        uint32_t cur;   // phase 0: node or leaf ref to process; phase 1: node to open (kInvalid: pop first)
        float cur_d;    // phase 0: exact-extent mindist of cur; phase 1: padded mindist of cur
        float best;     // re-squared distance of the winner, as the reference keeps it
        float thr;      // phase 0: (sqrt(best) + slack)^2; phase 1: thrW
        float slack;
        uint32_t win;   // leaf position of the winner
        uint32_t best_rank;
        int n_near;     // 3D: leaves seen within 8 ulp of the running minimum
        int phase;
        bool have_rank;
#ifdef WOST_TRAV_DEBUG
        int dbg_dec; // decade of max |coordinate| of the query: trips are counted per decade in overflow[8 + dec] (16 + dec: queries)
#endif
    };

    template <int D>
    WOST_DEV void nearest_enter_phase1(const DevBVH<D> &bvh, const Vec<D> &q, SmemStack &st, NearestTrav<D> &t, float thrW)
    {
        t.phase = 1;
        t.win = kInvalid;
        t.best = dm_inf();
        t.thr = thrW;
        t.cur = 0;
        t.cur_d = box_mindist(root_box_of(bvh), q);
        st.sp = 0;
    }
    // 2D far field (see nearest_begin): phase 2 = count near-ties, finish with the restricted pass; phase 3 = distance only
    template <int D>
    __device__ __noinline__ int nearest_far_phase(const float *root_box, Vec<D> q, float slack, float need_index_below, bool allow_brute)
    {
        Box<D> r;
#pragma unroll
        for (int k = 0; k < D; ++k)
        {
            r.up[k] = root_box[k];
            r.lo[k] = root_box[D + k];
        }
        const float lim = 2.0f * need_index_below + slack;
        if (!(box_mindist(r, q) > lim * lim)) // never true when need_index_below is inf
            return 2;
        // ultra-far (|q| > ~2^17 x the mesh extent, ~1e8 for a 600-unit scene): even the 2^-21 |q| shell of fp32-indistinguishable
        // candidates covers most of the mesh, a traversal visits ~all nodes one dependent load at a time (13 000 trips measured
        // on data/fille, and the handful of lanes that own such walks ran for seconds after every other warp had finished).
        // Phase 4: the warp scans all leaves cooperatively instead (nearest_brute_warp).
        float ext = 0.0f;
#pragma unroll
        for (int k = 0; k < D; ++k)
            ext = fmax_nn(ext, r.up[k] - r.lo[k]);
        return (allow_brute && slack > ext) ? 4 : 3; // slack is 2^-19 |q| here: the 2^-21 |q| shell is thicker than a quarter of the mesh
    }
    // Phase 4 (distance-only, ultra-far): all 32 lanes evaluate the leaves of the lanes that asked (`mine`), 32 leaves per
    // step, coalesced, no dependent loads; result = exact minimum of the re-squared fp32 leaf distances (lowest leaf position
    // among equals). Warp-synchronous: every lane of the warp calls it. Returns the lanes that were served.
    template <int D>
    __device__ __noinline__ uint2 nearest_brute_scan(const DevBVH<D> *bvh, Vec<D> q, unsigned served) // rare: out of line
    {
        const int lane = threadIdx.x & 31;
        uint2 mine_r = make_uint2(kInvalid, __float_as_uint(dm_inf()));
        for (unsigned m = served; m != 0u; m &= m - 1u)
        {
            const int src = __ffs(m) - 1;
            Vec<D> qs;
#pragma unroll
            for (int k = 0; k < D; ++k)
                qs[k] = __shfl_sync(kFullMask, q[k], src);
            float best = dm_inf();
            uint32_t win = kInvalid;
            for (uint32_t pos = (uint32_t)lane; pos < (uint32_t)bvh->nprims; pos += 32u)
            {
                const float d = dm_sqrt(prim_distance_d2(load_prim(bvh->leaf_prim + pos), qs));
                const float g = d * d;
                if (g < best)
                {
                    best = g;
                    win = pos;
                }
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1)
            {
                const float ob = __shfl_xor_sync(kFullMask, best, o);
                const uint32_t ow = __shfl_xor_sync(kFullMask, win, o);
                if (ob < best || (ob == best && ow < win))
                {
                    best = ob;
                    win = ow;
                }
            }
            if (lane == src)
                mine_r = make_uint2(win, __float_as_uint(best));
        }
        return mine_r;
    }
    template <int D>
    WOST_DEV unsigned nearest_brute_warp(const DevBVH<D> &bvh, const Vec<D> &q, bool mine, NearestTrav<D> &t)
    {
        const unsigned served = __ballot_sync(kFullMask, mine);
        if (__builtin_expect(served != 0u, 0))
        {
            const uint2 r = nearest_brute_scan<D>(&bvh, q, served);
            if (mine)
            {
                t.win = r.x;
                t.best = __uint_as_float(r.y);
            }
        }
        return served;
    }
    // -> true when trips are needed; false when the result is already in t (0 or 1 primitives)
    // bound: an a-priori upper bound of the result (inf when unknown). A walk that moved by s from a point whose
    // closest distance was R knows d(x') <= R + s; seeding the pruning threshold with it keeps the far siblings of the
    // first descent off the stack (they are otherwise pushed, and popped later only to be discarded).
    // need_index_below: the caller only uses the winner's INDEX when the distance is below this value (the walker: the
    // epsilon shell, uniform/integrator.cu:155-183; the probes: always -> inf). 2D far-field queries that provably end
    // above it (twice the value away from the root box) run in "distance-only" mode (phase 3): no near-tie bookkeeping, no
    // second pass, and a pruning slack of 2^-21 |q| that only has to cover the rounding mismatch between a box distance and a
    // leaf distance. The distance is then the exact minimum of the fp32 leaf distances over the reference's candidate set;
    // the reference's own pick in that regime depends on its visit order by a few ulp (its pop-time test misprunes once the
    // 1e-3 padding is below rounding), so this is a documented deviation for walks that have LEFT the scene by > 2000 units
    // (DESIGN.md section 5). Without it such walks -- the reference leaks them through the corners of closed Neumann
    // polygons -- cost a full traversal of every leaf per step.
    template <int D, bool RefOrder>
    WOST_DEV bool nearest_begin(const DevBVH<D> &bvh, const Vec<D> &q, SmemStack &st, NearestTrav<D> &t, float bound = dm_inf(),
                                float need_index_below = dm_inf(), bool allow_brute = false)
    {
        t.win = kInvalid;
        t.best = dm_inf();
        t.phase = 0;
        if (bvh.nprims <= 0)
            return false;
        if (bvh.nprims == 1)
        {
            const float d = prim_distance(load_prim(bvh.leaf_prim), q);
            t.win = 0;
            t.best = d * d;
            return false;
        }
        if (RefOrder)
        {
            nearest_enter_phase1(bvh, q, st, t, dm_inf());
            return true;
        }
        float qmax = 0.0f;
#pragma unroll
        for (int k = 0; k < D; ++k)
            qmax = fmax_nn(qmax, dm_abs(q[k]));
        if (D == 3 && qmax > 1024.0f * bvh.max_abs_coord)
        {
            if (st.overflow)
                atomicAdd(st.overflow + 1, 1); // statistics: full traversals in the reference's order
            nearest_enter_phase1(bvh, q, st, t, dm_inf());
            return true;
        }
        // 2D: the "no second pass" argument above needs the reference's box padding (1e-3) to dominate the rounding of a
        // distance, d * 2^-24 * 8 < 1e-3, i.e. coordinates below kFar2D = 1e-3 / (8 * 2^-24) ~ 2097. Beyond it (large scenes,
        // walks that left the domain) the query still runs in the B200 order, but counts near-ties like the 3D path (phase 2)
        // and finishes with the restricted reference-order pass when there was more than one: the cost then grows smoothly
        // with the distance (the pass covers the subtrees within 2^-19 |q| of the minimum) instead of jumping to a full
        // traversal of every leaf. (Round 1 switched to the full reference-order traversal at |q| > 2000: walks that leak
        // through a corner of a closed Neumann square -- the reference's walks do, tests/test_oracle_b1_pin.py -- and wander
        // off to 1e3..1e8 made data/fille 9x slower per walk-step than data/ladybug.)
#ifdef WOST_TRAV_DEBUG
        t.dbg_dec = qmax < 600.0f ? 0 : (qmax < 2000.0f ? 1 : (qmax < 1e4f ? 2 : (qmax < 1e5f ? 3 : (qmax < 1e6f ? 4 : (qmax < 1e7f ? 5 : (qmax < 1e8f ? 6 : 7))))));
        if (st.overflow)
            atomicAdd(st.overflow + 16 + t.dbg_dec, 1);
#endif
        t.slack = fmax_nn(qmax, bvh.max_abs_coord) * 1.9073486328125e-06f; // 2^-19
        if (D == 2 && __builtin_expect(qmax > kFar2D, 0))
        {
            t.phase = nearest_far_phase<D>(bvh.root_box, q, t.slack, need_index_below, allow_brute); // rare: out of line (registers of the hot loop)
            if (t.phase >= 3)
                t.slack *= 0.25f; // 2^-21 |q|
        }
        {
            const float u = bound + t.slack; // inf stays inf
            t.thr = u * u;
        }
        t.best_rank = 0;
        t.have_rank = false;
        t.n_near = 0;
        t.cur = 0;
        t.cur_d = 0.0f;
        st.sp = 0;
        return true;
    }
    template <int D>
    WOST_DEV float nearest_distance(const NearestTrav<D> &t) { return dm_sqrt(t.best); }

    // one trip of phase 1 -> still active
    template <int D, bool RefOrder>
    WOST_DEV bool nearest_trip_ref(const DevBVH<D> &bvh, const Vec<D> &q, SmemStack &st, NearestTrav<D> &t)
    {
        if (t.cur == kInvalid)
        {
            if (st.empty())
                return false;
            const StackEntry e = st.pop();
            t.cur = e.ref;
            t.cur_d = e.d;
        }
        const uint32_t node = t.cur;
        t.cur = kInvalid;
        if (t.cur_d > t.best)
            return true;
        const TNode<D> *rec = bvh.tnodes + node;
        const NodeViewE<D> n = load_tnode_e(rec);
        const Box<D> lp = pad_box(n.le), rp = pad_box(n.re);
        const float Lmin = box_mindist(lp, q), Rmin = box_mindist(rp, q);
        const float Lmm = box_minmaxdist(lp, q), Rmm = box_minmaxdist(rp, q);
        bool goL = Lmin <= Rmm, goR = Rmin <= Lmm;
        if (!RefOrder)
        {
            // restricted pass: only subtrees that can hold a leaf within thrW matter (exact extent, then oriented box)
            goL = goL && box_mindist(n.le, q) <= t.thr;
            goR = goR && box_mindist(n.re, q) <= t.thr;
            if (goL)
                goL = obb_mindist2(load_obb(rec, 0), q) <= t.thr;
            if (goR)
                goR = obb_mindist2(load_obb(rec, 1), q) <= t.thr;
        }
#pragma unroll
        for (int side = 0; side < 2; ++side)
        {
            const uint32_t ref = side == 0 ? n.lref : n.rref;
            if ((side == 0 ? goL : goR) && (ref & kLeafFlag))
            {
                const uint32_t pos = ref & ~kLeafFlag;
                float d = prim_distance(load_prim(bvh.leaf_prim + pos), q);
                d *= d;
                if (d <= t.best)
                {
                    t.best = d;
                    t.win = pos;
                }
            }
        }
        const bool inL = goL && !(n.lref & kLeafFlag), inR = goR && !(n.rref & kLeafFlag);
        if (inL && inR)
            st.push(n.lref, Lmin); // R would be pushed last and popped next: keep it in registers
        if (inR)
        {
            t.cur = n.rref;
            t.cur_d = Rmin;
        }
        else if (inL)
        {
            t.cur = n.lref;
            t.cur_d = Lmin;
        }
        return true;
    }

    // distance to one leaf primitive + the bookkeeping of the running minimum (phase 0)
    template <int D>
    WOST_DEV void nearest_eval_leaf(const DevBVH<D> &bvh, const Vec<D> &q, NearestTrav<D> &t, uint32_t pos)
    {
        const float d2 = prim_distance_d2(load_prim(bvh.leaf_prim + pos), q);
        if (d2 <= t.best * 1.000002f)
        {
            const float d = dm_sqrt(d2);
            const float g = d * d;
            bool take = g < t.best;
            if (D == 3 || t.phase == 2)
            {
                if (take)
                    t.n_near = (t.best <= g * 1.000001f) ? t.n_near + 1 : 1;
                else if (g <= t.best * 1.000001f)
                    ++t.n_near;
            }
            if (take)
                t.have_rank = false;
            else if (D == 2 && g == t.best && t.win != kInvalid)
            {
                // exact tie: the reference keeps the last visited leaf of its own (static) order
                const uint32_t rk = __ldg(bvh.leaf_rank + pos);
                if (!t.have_rank)
                {
                    t.best_rank = __ldg(bvh.leaf_rank + t.win);
                    t.have_rank = true;
                }
                take = rk > t.best_rank;
                if (take)
                    t.best_rank = rk;
            }
            if (take)
            {
                t.best = g;
                t.win = pos;
                const float u = d + t.slack;
                t.thr = u * u;
            }
        }
    }

    // one trip -> still active
    // Phase 0 opens ONE internal node per trip and evaluates its leaf children on the spot (nearer one first): leaves never
    // travel through `cur` or the stack, so a query takes ~30% fewer trips and the stack only ever holds internal nodes.
    template <int D, bool RefOrder>
    WOST_DEV bool nearest_trip(const DevBVH<D> &bvh, const Vec<D> &q, SmemStack &st, NearestTrav<D> &t)
    {
#ifdef WOST_TRAV_DEBUG
        if (st.overflow)
            atomicAdd(st.overflow + 8 + t.dbg_dec, 1);
#endif
        if (RefOrder || __builtin_expect(t.phase == 1, 0)) // phase 1: far field / 3D tie resolution, kept off the hot path
            return nearest_trip_ref<D, RefOrder>(bvh, q, st, t);
        bool have = false;
        {
            const TNode<D> *rec = bvh.tnodes + t.cur;
            const NodeViewE<D> n = load_tnode_e(rec);
            float Le = box_mindist(n.le, q), Re = box_mindist(n.re, q);
            bool goL = Le <= t.thr, goR = Re <= t.thr;
            // children that pass on their AABB are tested against their oriented box as well (pruning only, see Obb). Not in the
            // distance-only far-field phases (>= 3): their slack is 2^-21 |q|, below the rounding of a projection onto the frame
            // (~2^-20 |q|), and that far from the mesh every box looks like a point anyway.
            const bool use_obb = D == 3 || t.phase < 3;
            if (goL && use_obb)
            {
                Le = fmax_nn(Le, obb_mindist2(load_obb(rec, 0), q));
                goL = Le <= t.thr;
            }
            if (goR && use_obb)
            {
                Re = fmax_nn(Re, obb_mindist2(load_obb(rec, 1), q));
                goR = Re <= t.thr;
            }
            if (goL || goR)
            {
                // the reference's sibling rule on the padded boxes: child X is visited iff mindist(X) <= minmaxdist(Y).
                // mindist(B) <= minmaxdist(B) holds for every box also in fp32 (same subtractions, same summation order,
                // termwise <=, monotone rounding), so the child whose padded mindist does not exceed its sibling's passes
                // without evaluating the rule: at most ONE minmaxdist per node, that of the nearer child's box.
                const Box<D> lp = pad_box(n.le), rp = pad_box(n.re);
                const float Lp = box_mindist(lp, q), Rp = box_mindist(rp, q);
                const bool testL = goL && Lp > Rp, testR = goR && Rp > Lp;
                if (testL || testR)
                {
                    Box<D> nb;
#pragma unroll
                    for (int k = 0; k < D; ++k)
                    {
                        nb.up[k] = testL ? rp.up[k] : lp.up[k];
                        nb.lo[k] = testL ? rp.lo[k] : lp.lo[k];
                    }
                    const float mm = box_minmaxdist(nb, q);
                    if (testL)
                        goL = Lp <= mm;
                    else
                        goR = Rp <= mm;
                }
                const bool lnear = Le <= Re;
                // leaf children: evaluate now, nearer first (the second one only if it can still matter)
                const bool leafL = goL && (n.lref & kLeafFlag), leafR = goR && (n.rref & kLeafFlag);
                if (leafL || leafR)
                {
                    const bool firstL = leafL && (lnear || !leafR);
                    const uint32_t p1 = (firstL ? n.lref : n.rref) & ~kLeafFlag;
                    nearest_eval_leaf<D>(bvh, q, t, p1);
                    if (leafL && leafR)
                    {
                        const float d2nd = firstL ? Re : Le;
                        if (d2nd <= t.thr)
                            nearest_eval_leaf<D>(bvh, q, t, (firstL ? n.rref : n.lref) & ~kLeafFlag);
                    }
                    if (leafL)
                        goL = false;
                    if (leafR)
                        goR = false;
                    // the threshold may have shrunk: re-test the internal sibling
                    goL = goL && Le <= t.thr;
                    goR = goR && Re <= t.thr;
                }
                if (goL && goR)
                {
                    st.push(lnear ? n.rref : n.lref, lnear ? Re : Le);
                    t.cur = lnear ? n.lref : n.rref;
                    t.cur_d = lnear ? Le : Re;
                    have = true;
                }
                else if (goL || goR)
                {
                    t.cur = goL ? n.lref : n.rref;
                    t.cur_d = goL ? Le : Re;
                    have = true;
                }
            }
        }
        if (have)
            return true;
        while (!st.empty())
        {
            const StackEntry e = st.pop();
            if (e.d <= t.thr)
            {
                t.cur = e.ref;
                t.cur_d = e.d;
                return true;
            }
        }
        if ((D == 3 || t.phase == 2) && __builtin_expect(t.n_near > 1, 0))
        {
            if (st.overflow)
                atomicAdd(st.overflow + 2, 1); // statistics: restricted second passes
            nearest_enter_phase1(bvh, q, st, t, t.thr);
            return true;
        }
        return false;
    }

    // Warp-synchronous drivers: the 32 lanes of a warp call these together; lanes without a query pass valid = false.
    // Returns the LEAF POSITION of the winner (original index = leaf_obj[pos]) and the distance.
    template <int D, bool RefOrder>
    WOST_DEV void query_nearest_sync(const DevBVH<D> &bvh, const Vec<D> &q, SmemStack &st, uint32_t &leaf_pos, float &dist, bool valid,
                                     float need_index_below = dm_inf())
    {
        NearestTrav<D> t;
        bool active = false;
        if (valid)
            active = nearest_begin<D, RefOrder>(bvh, q, st, t, dm_inf(), need_index_below, true);
        if (!RefOrder && D == 2 && (nearest_brute_warp<D>(bvh, q, active && t.phase == 4, t) >> (threadIdx.x & 31) & 1u))
            active = false;
        while (__any_sync(kFullMask, active))
        {
            if (active)
                active = nearest_trip<D, RefOrder>(bvh, q, st, t);
        }
        if (valid)
        {
            leaf_pos = t.win;
            dist = nearest_distance(t);
        }
    }
    template <int D>
    WOST_DEV void query_nearest(const DevBVH<D> &bvh, const Vec<D> &q, SmemStack &st, uint32_t &leaf_pos, float &dist, bool valid = true)
    {
        query_nearest_sync<D, false>(bvh, q, st, leaf_pos, dist, valid);
    }
    template <int D>
    WOST_DEV void query_nearest_ref(const DevBVH<D> &bvh, const Vec<D> &q, SmemStack &st, uint32_t &leaf_pos, float &dist, bool valid = true)
    {
        query_nearest_sync<D, true>(bvh, q, st, leaf_pos, dist, valid);
    }

    // ---------------------------------------------------------------- closest silhouette (core/query.cuh:325-423)
    template <int D>
    WOST_DEV bool leaf_sil_distance(const DevBVH<D> &bvh, uint32_t pos, const Vec<D> &q, float max_r2, float &dist, bool flip)
    {
        // scene.cuh:517-541 / 978-1003
        float maxd = max_r2;
        bool ret = false;
#pragma unroll
        for (int j = 0; j < D; ++j)
        {
            const SilRec<D> s = load_sil(bvh.leaf_sil + (size_t)pos * D + j);
            if (!(s.flags & 1u))
                continue;
            if (sil_closest(s, q, maxd, dist, flip))
            {
                ret = true;
                maxd = dist * dist;
            }
        }
        return ret;
    }
    // bound: the caller only uses min(bound, result) (the walker: R_B = min(R_D, R_N), integrator.cu:186-195), so the search
    // starts with best = bound: subtrees and leaf children farther than that are skipped before their cone test (acos / asin /
    // atan2 per child). Silhouettes closer than the bound are found exactly as without it -- the result is the minimum over the
    // candidates that pass the (node-local) cone tests, independent of the visit order -- and when there is none the function
    // returns the bound itself. bound = inf is the reference's query.
    template <int D>
    WOST_DEV float query_silhouette(const DevBVH<D> &bvh, const Vec<D> &q, bool flip, SmemStack &st, float bound = dm_inf())
    {
        float best = bound;
        if (bvh.nprims <= 0)
            return best;
        if (bvh.nprims == 1)
        {
            float d = dm_inf();
            if (bvh.root_cone[D] >= 0.0f && leaf_sil_distance(bvh, 0, q, dm_inf(), d, flip) && d <= best)
                best = d;
            return best;
        }
        st.sp = 0;
        uint32_t node = 0;
        float node_d = box_mindist(root_box_of(bvh), q);
        bool have = true;
        while (true)
        {
            if (!have)
            {
                if (st.empty())
                    break;
                const StackEntry e = st.pop();
                node = e.ref;
                node_d = e.d;
            }
            have = false;
            if (node_d > best * best)
                continue;
            const TNode<D> *rec = bvh.tnodes + node;
            const NodeView<D> n = load_tnode(rec);
            Cone<D> lc, rc;
            load_cnode(bvh.cnodes + node, lc, rc);
            const float Lmin = box_mindist(n.lb, q), Rmin = box_mindist(n.rb, q);
            const float best2 = best * best;
            // a child whose oriented box (a lower bound of the distance to every vertex and edge below it) lies beyond the running
            // minimum cannot improve it: skipped before its cone test (pruning only, like the closest-point query)
            const bool hitL = lc.half_angle >= 0.0f && Lmin <= best2 && obb_mindist2(load_obb(rec, 0), q) <= best2 && cone_overlap(lc, q, n.lb, Lmin);
            const bool hitR = rc.half_angle >= 0.0f && Rmin <= best2 && obb_mindist2(load_obb(rec, 1), q) <= best2 && cone_overlap(rc, q, n.rb, Rmin);
            bool pushL = false;
            if (hitL)
            {
                if (n.lref & kLeafFlag)
                {
                    float d = dm_inf();
                    const bool found = leaf_sil_distance(bvh, n.lref & ~kLeafFlag, q, best * best, d, flip);
                    if (found && d <= best)
                        best = d;
                }
                else
                    pushL = true;
            }
            if (hitR)
            {
                if (n.rref & kLeafFlag)
                {
                    float d = dm_inf();
                    const bool found = leaf_sil_distance(bvh, n.rref & ~kLeafFlag, q, best * best, d, flip);
                    if (found && d <= best)
                        best = d;
                }
                else
                {
                    if (pushL)
                        st.push(n.lref, Lmin);
                    pushL = false;
                    node = n.rref;
                    node_d = Rmin;
                    have = true;
                }
            }
            if (pushL)
            {
                node = n.lref;
                node_d = Lmin;
                have = true;
            }
        }
        return best;
    }

    // ---------------------------------------------------------------- rays (core/query.cuh:79-169)
    struct RayResult
    {
        bool hit;
        float t, u, v;
        uint32_t leaf_pos;
    };
    template <int D, bool TestOnly>
    WOST_DEV RayResult query_ray(const DevBVH<D> &bvh, const Ray<D> &r, float max_dist, SmemStack &st)
    {
        RayResult res;
        res.hit = false, res.t = dm_inf(), res.u = 0.0f, res.v = 0.0f, res.leaf_pos = kInvalid;
        if (bvh.nprims <= 0)
            return res;
        float best = dm_inf();
        st.sp = 0;
        uint32_t ref = (bvh.nprims == 1) ? kLeafFlag : 0u;
        float ref_d = dm_inf();
        bool have = true;
        while (true)
        {
            if (!have)
            {
                if (st.empty())
                    break;
                const StackEntry e = st.pop();
                ref = e.ref;
                ref_d = e.d;
            }
            have = false;
            if (ref_d > best)
                continue;
            if (ref & kLeafFlag)
            {
                const uint32_t pos = ref & ~kLeafFlag;
                const Hit h = ray_prim(load_prim(bvh.leaf_prim + pos), r);
                if (h.hit && h.t < max_dist && h.t < best)
                {
                    best = h.t;
                    res.hit = true;
                    if (TestOnly)
                        return res;
                    res.u = h.u;
                    res.v = h.v;
                    res.leaf_pos = pos;
                }
            }
            else
            {
                const NodeView<D> n = load_tnode(bvh.tnodes + ref);
                float Ld = 0.0f, Rd = 0.0f;
                const bool Lh = ray_box(r, n.lb, max_dist, &Ld);
                const bool Rh = ray_box(r, n.rb, max_dist, &Rd);
                if (Lh && Rh)
                {
                    uint32_t closer = n.lref, other = n.rref;
                    if (Rd < Ld)
                    {
                        const float t = Ld;
                        Ld = Rd;
                        Rd = t;
                        closer = n.rref;
                        other = n.lref;
                    }
                    st.push(other, Rd);
                    ref = closer;
                    ref_d = Ld;
                    have = true;
                }
                else if (Lh)
                {
                    ref = n.lref;
                    ref_d = Ld;
                    have = true;
                }
                else if (Rh)
                {
                    ref = n.rref;
                    ref_d = Rd;
                    have = true;
                }
            }
        }
        res.t = best;
        return res;
    }

    // ---------------------------------------------------------------- Green-weighted descent (core/sample.cuh:23-92)
    template <int D>
    WOST_DEV void sample_in_sphere(const DevBVH<D> &bvh, const Vec<D> &c, float radius, float u, int &leaf_pos, float &pdf_out)
    {
        leaf_pos = -1;
        pdf_out = 0.0f;
        if (bvh.nprims <= 0)
            return;
        uint32_t ref = (bvh.nprims == 1) ? kLeafFlag : 0u;
        float pdf = 1.0f;
        for (;;)
        {
            if (ref & kLeafFlag)
            {
                const uint32_t pos = ref & ~kLeafFlag;
                const Prim<D> pr = load_prim(bvh.leaf_prim + pos);
                if (sphere_prim(pr, c, radius))
                {
                    leaf_pos = (int)pos;
                    pdf_out = dm_div(pdf, prim_measure(pr));
                }
                return;
            }
            const NodeView<D> n = load_tnode(bvh.tnodes + ref);
            const float Lw = sphere_box(c, radius, n.lb) ? green_weight(c, box_centroid(n.lb)) : 0.0f;
            const float Rw = sphere_box(c, radius, n.rb) ? green_weight(c, box_centroid(n.rb)) : 0.0f;
            const float tw = Lw + Rw;
            if (!(tw > 0))
                return;
            const float Lp = dm_div(Lw, tw);
            if (u < Lp)
            {
                u = dm_div(u, Lp);
                ref = n.lref;
                pdf = Lp * pdf;
            }
            else
            {
                const float Rp = 1.0f - Lp;
                u = dm_div(u - Lp, Rp);
                ref = n.rref;
                pdf = Rp * pdf;
            }
        }
    }
} // namespace wost
