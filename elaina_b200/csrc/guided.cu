// Guided walk-on-stars integrator (replaces integrator/guided/integrator.cu:110-1094).
//
// Per samples-per-pixel batch:
//   k_g_first            evaluation points -> closest point / epsilon-shell / boundary + source + Neumann contributions of
//                        depth 0 -> compacted queue of out-shell items                 (K3 + K4..K7 of the reference)
//   k_g_bounce (x G)     ONE kernel per guided bounce d < maxGuidedDepth, 128 items per CTA tile:
//                        DenseGrid gather -> 4-layer MLP on tcgen05/TMEM -> VMM decode -> selection -> guided or
//                        uniform direction with the MIS pdf -> Neumann reflection fold -> ray -> walk record ->
//                        the SAME thread continues with the closest-point query and contributions of depth d+1 and
//                        appends the survivor to the next queue      (K10..K15 + K4..K7; 2 host syncs + 7 launches
//                        + a 132-164 B/row network output round trip per depth in the reference)
//   k_g_finish           walks that outlive the guided depths run to completion with uniform steps, in registers
//   k_g_count/scatter    walk records -> compacted training set, deterministic order       (generate_training_data)
//   k_g_train (x <= 5)   forward with stored activations, KL gradient, backward, weight gradients accumulated in TMEM
//   k_g_adam_ema         Adam + EMA + fp16 working copies in one pass                       (adam_step + ema_step)
// No host synchronisation inside a solve: batch sizes live in a device control block (TrainCtl).
#include "integrator.cuh"
#include "mlp_tc.cuh"
#include "vmm.cuh"
#include "walk.cuh"

#include <cub/device/device_scan.cuh>

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>

#include "json.h"

namespace wost
{
    constexpr int kGBlock = 128; // == kTileRows: thread r of the CTA owns row r of the MLP tile

    struct GQueue
    {
        float4 *q0, *q1, *q2;
        float *q3;
    };
    template <int D>
    struct GItem
    {
        WalkState<D> w;
        float R_B;
        float T[3]; // contributions of the walk so far
    };
    template <int D>
    WOST_DEV void g_load(const GQueue &q, size_t i, const uint32_t *active, GItem<D> &it)
    {
        const float4 a = q.q0[i], b = q.q1[i], c = q.q2[i];
        it.w.x[0] = a.x, it.w.x[1] = a.y;
        if (D == 3)
            it.w.x[D - 1] = a.z;
        it.w.thp = a.w;
        it.w.nN[0] = b.x, it.w.nN[1] = b.y;
        if (D == 3)
            it.w.nN[D - 1] = b.z;
        it.R_B = b.w;
        it.w.ai = __float_as_uint(c.x);
        it.w.pixel = __ldg(active + it.w.ai);
        const uint32_t sf = __float_as_uint(c.y);
        it.w.sample = sf & 0x7fffffffu;
        it.w.onN = (sf >> 31) != 0;
        it.T[0] = c.z, it.T[1] = c.w;
        it.T[2] = q.q3[i];
    }
    template <int D>
    WOST_DEV void g_store(const GQueue &q, size_t i, const GItem<D> &it)
    {
        q.q0[i] = make_float4(it.w.x[0], it.w.x[1], D == 3 ? it.w.x[D - 1] : 0.0f, it.w.thp);
        q.q1[i] = make_float4(it.w.nN[0], it.w.nN[1], D == 3 ? it.w.nN[D - 1] : 0.0f, it.R_B);
        q.q2[i] = make_float4(__uint_as_float(it.w.ai), __uint_as_float(it.w.sample | (it.w.onN ? 0x80000000u : 0u)), it.T[0], it.T[1]);
        q.q3[i] = it.T[2];
    }

    // record / training-set field offsets (in units of [kMaxTrainDepth][n_active] resp. [cap] planes)
    template <int D>
    struct RecLayout
    {
        static constexpr int kPos = 0, kDir = D, kNrm = 2 * D, kPdf = 3 * D, kThp = 3 * D + 1, kBase = 3 * D + 2, kPlanes = 3 * D + 5;
        static constexpr int sPos = 0, sDir = D, sNrm = 2 * D, sPdf = 3 * D, sLi = 3 * D + 1, sPlanes = 3 * D + 2;
    };

    struct GArgs
    {
        GQueue in, out;
        const int *in_count; // items of this bounce (device)
        int *out_count;
        int *cursor; // tile / item claim cursor
        unsigned long long *cursor64; // claim cursor of the first arrival (n_active * batch_samples may exceed 2^31)
        unsigned long long *steps;
        const uint32_t *active;
        const uint2 *first; // first-step cache (k_first_hit), indexed like `active`
        int n_active;
        uint32_t sample0;
        int batch_samples;
        uint32_t depth;
        float4 *acc; // [slot][active rank]
        int slot0;   // accumulator slot of sample0 (training batches of one sample rotate through the slots)
        int max_depth;
        int *overflow;
        // guiding
        int training;        // records are written for depth < max_train_depth
        int max_train_depth;
        float uniform_fraction;
        float *rec;
        uint8_t *rec_onN;
        int *rec_count;
        float *total;
        const float *precomputed; // validation mode: raw outputs [item][48] from k_g_mlp_simt, else nullptr
        int refill;               // lanes that must be waiting before the warp runs a tail phase (see default_refill_threshold)
    };

    // closest-point query + separate / boundary / source / Neumann stages of the point the walk has just reached.
    // Warp-synchronous: every lane calls it; lanes without a walk pass valid = false. Returns true when the walk goes on.
    template <int D>
    WOST_DEV bool g_arrive(const WalkCtx<D> &c, GItem<D> &it, uint32_t depth, SmemStack &st, float4 *acc_base, uint32_t sample0, int n_active, bool valid)
    {
        float R_D = dm_inf();
        uint32_t pos = kInvalid;
        if (c.dir.nprims > 0)
            query_nearest_sync<D, false>(c.dir, it.w.x, st, pos, R_D, valid, c.eps);
        if (!valid)
            return false;
        PixelAccT<true> acc;
        acc.init(reinterpret_cast<float *>(acc_base + (size_t)(it.w.sample - sample0) * n_active + it.w.ai));
        acc.tot[0] = it.T[0], acc.tot[1] = it.T[1], acc.tot[2] = it.T[2];
        const bool alive = walk_step_contrib<D>(c, it.w, depth, st, acc, R_D, pos, 1.0f, it.R_B);
        acc.flush();
        it.T[0] = acc.tot[0], it.T[1] = acc.tot[1], it.T[2] = acc.tot[2];
        return alive;
    }
    // GuidedPixelStateBuffer::incrementDepth (guided.h:24-44): the vertex the direction was sampled at
    template <int D>
    WOST_DEV void g_record(const GArgs &a, const GItem<D> &it, const Vec<D> &dir, float pdf, uint32_t depth)
    {
        using L = RecLayout<D>;
        const size_t plane = (size_t)kMaxTrainDepth * a.n_active, o = (size_t)depth * a.n_active + it.w.ai;
#pragma unroll
        for (int k = 0; k < D; ++k)
        {
            a.rec[(L::kPos + k) * plane + o] = it.w.x[k];
            a.rec[(L::kDir + k) * plane + o] = dir[k];
            a.rec[(L::kNrm + k) * plane + o] = it.w.nN[k];
        }
        a.rec[L::kPdf * plane + o] = pdf;
        a.rec[L::kThp * plane + o] = it.w.thp;
#pragma unroll
        for (int ch = 0; ch < 3; ++ch)
            a.rec[(L::kBase + ch) * plane + o] = it.T[ch];
        a.rec_onN[o] = it.w.onN ? 1 : 0;
        a.rec_count[it.w.ai] = (int)depth + 1;
    }
    template <int D>
    WOST_DEV void g_finalize(const GArgs &a, const GItem<D> &it)
    {
        if (a.training)
        {
#pragma unroll
            for (int ch = 0; ch < 3; ++ch)
                a.total[(size_t)ch * a.n_active + it.w.ai] = it.T[ch];
        }
    }
    // warp-aggregated append of the survivors of this warp
    template <int D>
    WOST_DEV void g_push(const GArgs &a, const GItem<D> &it, bool alive)
    {
        const int lane = threadIdx.x & 31;
        const uint32_t mask = __ballot_sync(kFullMask, alive);
        if (mask)
        {
            int slot0 = 0;
            if (lane == 0)
                slot0 = atomicAdd(a.out_count, __popc(mask));
            slot0 = __shfl_sync(kFullMask, slot0, 0);
            if (alive)
                g_store(a.out, (size_t)slot0 + __popc(mask & ((1u << lane) - 1u)), it);
        }
    }

    // ---- arrival of a walk at a new point: closest-point query + separate / boundary / source / Neumann stages.
    // Persistent lanes with refill, like k_walk_persistent (integrator.cu): a lane owns one item at a time, the warp
    // alternates between traversal trips (every lane with a query in flight opens one node) and a tail phase in which
    // the lanes whose query has finished run the contribution stages, append the survivors to the out-shell queue
    // (one atomic per warp) and claim new items (one atomic per warp). The number of node visits varies ~10x between
    // neighbouring items; lock-step traversal (one item per lane per round) left 7 of 32 lanes busy
    // (profiles/r1_notes.md), refilling keeps the traversal -- the bulk of the instructions -- near full width.
    // First: items are synthesised from (pixel, sample) = generateEvaluationPointsImpl (integrator.cu:131-151).
    // lean 3D variants are compiled for 8 CTAs/SM like k_walk_persistent (integrator.cu: persist_min_ctas); 0 = compiler's budget
    constexpr int g_min_ctas(int D, int F) { return (D == 3 && !(F & kFeatNeumann)) ? 8 : 0; }
    template <int D, bool First, int F>
    __global__ void __launch_bounds__(kGBlock, g_min_ctas(D, F)) k_g_arrive(const __grid_constant__ WalkCtx<D> c, const __grid_constant__ GArgs a)
    {
        extern __shared__ uint2 smem_stack[];
        SmemStack st;
        uint2 spill_mem[kSpillLevels];
        st.init(smem_stack, a.overflow, spill_mem);
        const int lane = threadIdx.x & 31;
        const long long njobs = First ? (long long)a.n_active * a.batch_samples : (long long)*a.in_count;
        GItem<D> it;
        NearestTrav<D> t;
        bool has = false, exhausted = false, trav = false;
        unsigned int my_steps = 0;
        it.w.pixel = 0, it.w.ai = 0, it.w.sample = a.sample0, it.w.thp = 0.0f, it.w.onN = false, it.R_B = 0.0f;
        it.T[0] = it.T[1] = it.T[2] = 0.0f;
#pragma unroll
        for (int k = 0; k < D; ++k)
            it.w.x[k] = it.w.nN[k] = 0.0f;
        t.win = kInvalid, t.best = 0.0f;
        for (;;)
        {
            const unsigned trav_mask = __ballot_sync(kFullMask, trav);
            const unsigned pend_mask = __ballot_sync(kFullMask, !trav && (has || !exhausted));
            if (pend_mask != 0u && (__popc(pend_mask) >= a.refill || trav_mask == 0u))
            {
                bool on = false;
                if (!trav && has)
                {
                    PixelAccT<true> acc;
                    acc.init(reinterpret_cast<float *>(a.acc + (size_t)(it.w.sample - a.sample0 + (uint32_t)a.slot0) * a.n_active + it.w.ai));
                    acc.tot[0] = it.T[0], acc.tot[1] = it.T[1], acc.tot[2] = it.T[2];
                    on = walk_step_contrib<D, F>(c, it.w, a.depth, st, acc, nearest_distance(t), t.win, 1.0f, it.R_B);
                    acc.flush();
                    it.T[0] = acc.tot[0], it.T[1] = acc.tot[1], it.T[2] = acc.tot[2];
                    if (!on)
                        g_finalize<D>(a, it);
                    has = false;
                }
                g_push<D>(a, it, on);
                const bool want = !trav && !exhausted;
                const unsigned want_mask = __ballot_sync(kFullMask, want);
                if (want_mask != 0u)
                {
                    const int leader = __ffs(want_mask) - 1;
                    long long base = 0;
                    if (lane == leader)
                        base = First ? (long long)atomicAdd(reinterpret_cast<unsigned long long *>(a.cursor64), (unsigned long long)__popc(want_mask))
                                     : (long long)atomicAdd(a.cursor, __popc(want_mask));
                    base = __shfl_sync(kFullMask, base, leader);
                    if (want)
                    {
                        const long long i = base + __popc(want_mask & ((1u << lane) - 1u));
                        if (i < njobs)
                        {
                            if (First)
                            {
                                const uint32_t sj = (uint32_t)(i / a.n_active);
                                const long long ai = i - (long long)sj * a.n_active;
                                it.w.ai = (uint32_t)ai;
                                it.w.pixel = __ldg(a.active + ai);
                                it.w.sample = a.sample0 + sj;
                                it.w.x = eval_point(c, it.w.pixel);
                                it.w.thp = 1.0f;
                                it.w.onN = false;
                                it.R_B = 0.0f;
                                it.T[0] = it.T[1] = it.T[2] = 0.0f;
#pragma unroll
                                for (int k = 0; k < D; ++k)
                                    it.w.nN[k] = 0.0f;
                                // the closest-point query of an evaluation point is answered once per solve (k_first_hit)
                                const uint2 f = __ldg(a.first + ai);
                                t.win = f.x;
                                t.best = __uint_as_float(f.y);
                            }
                            else
                                g_load<D>(a.in, (size_t)i, a.active, it);
                            has = true;
                            ++my_steps;
                            if (!First)
                            {
                                // the bounce moved the walk by R_B * dir from a point whose closest distance was R_D; without a
                                // Neumann mesh R_B = max(1e-4, R_D) >= R_D, so d(x') <= 2 R_B (triangle inequality + rounding slack)
                                const float bound = c.neu.nprims == 0 ? it.R_B * 2.0002f : dm_inf();
                                trav = nearest_begin<D, false>(c.dir, it.w.x, st, t, bound, c.eps, (F & kFeatNeumann) != 0);
                            }
                        }
                        else
                            exhausted = true;
                    }
                }
                // walks ~1e8 units away from the scene: the warp scans the leaves together (lbvh.cuh, phase 4)
                if (D == 2 && (F & kFeatNeumann) && (nearest_brute_warp<D>(c.dir, it.w.x, trav && t.phase == 4, t) >> lane & 1u))
                    trav = false;
            }
            else if (trav_mask == 0u)
                break;
            if (trav)
                trav = nearest_trip<D, false>(c.dir, it.w.x, st, t);
        }
        unsigned int ws = my_steps;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1)
            ws += __shfl_xor_sync(kFullMask, ws, o);
        if (lane == 0 && ws)
            atomicAdd(a.steps, (unsigned long long)ws);
    }

    // Direction sampling of one out-shell item given the raw network output `raw` (kOut floats):
    // handleOutShellPointImpl + handleGuidedSamplingImpl / handleUniformSamplingImpl (integrator.cu:496-526, 670-880).
    template <int D>
    WOST_DEV Vec<D> g_sample_direction(const WalkCtx<D> &c, const GArgs &a, const GItem<D> &it, const float *raw, float &pdf)
    {
        const WalkState<D> &w = it.w;
        const float p_sel = g_logistic(raw[kLobes * (D + 2)]);
        const bool inside = aabb_contains<D>(c.aabb_min, c.aabb_max, w.x);
        const Philox4 qs = wost_rng(c.seed, w.pixel, w.sample, a.depth, WOST_RNG_SELECT, 0);
        // uniformSamplingFraction is only ever compared with 0 and 1 by the reference (:514, :1034): 0 = always guided,
        // < 1 = learned selection probability; >= 1 = the guided kernel is skipped -- there the reference silently drops
        // the walks it had queued for guided sampling, here they take the uniform branch.
        bool guided = inside && a.uniform_fraction < 1.0f && (a.uniform_fraction == 0.0f || u32_to_unit_float(qs.x) < p_sel);
        VMM<D> m;
        if (inside)
            m.decode(raw);
        Vec<D> dir;
        float uniformPdf = w.onN ? Dims<D>::hemi_pdf() : Dims<D>::sphere_pdf();
        if (guided)
        {
            const int lobe = m.pick(u32_to_unit_float(qs.y));
            float kappa = m.kappa[0];
            Vec<D> mu = m.mu[0];
#pragma unroll
            for (int i = 1; i < kLobes; ++i)
                if (i == lobe)
                {
                    kappa = m.kappa[i];
                    mu = m.mu[i];
                }
            dir = lobe_sample(kappa, mu, c.seed, w.pixel, w.sample, a.depth);
            float guidedPdf = m.pdf(dir);
            if (w.onN)
            {
                const Vec<D> refl = vreflect(dir, w.nN);
                if (vdot(w.nN, dir) <= 0.0f)
                    dir = refl;
                guidedPdf += m.pdf(refl);
            }
            pdf = p_sel * guidedPdf + (1.0f - p_sel) * uniformPdf;
        }
        else
        {
            dir = uniform_direction<D>(c, w, a.depth, pdf);
            if (inside)
            {
                float guidedPdf = m.pdf(dir);
                if (w.onN)
                    guidedPdf += m.pdf(vreflect(dir, w.nN));
                pdf = p_sel * guidedPdf + (1.0f - p_sel) * pdf;
            }
        }
        return dir;
    }

    constexpr int kBounceStackLevels = 12;
    static_assert(kBounceStackLevels >= kMinSmemLevels && kBounceStackLevels <= kSmemLevels, "spill area is sized for kMinSmemLevels");
    // ---- one guided bounce
    template <int D, int K0>
    __global__ void __launch_bounds__(kGBlock) k_g_bounce(const __grid_constant__ WalkCtx<D> c, const __grid_constant__ GArgs a, const __grid_constant__ GuideNetView net)
    {
        extern __shared__ __align__(1024) char smem_raw[];
        SmemStack st;
        uint2 spill_mem[kSpillLevels];
        // the bounce kernel only traverses for the ray cast against the Neumann mesh: a 12-level shared stack (deeper levels
        // spill) keeps the CTA at 55 KB of shared memory -> 4 CTAs per SM instead of 3 (the kernel is latency-bound)
        st.init(reinterpret_cast<uint2 *>(smem_raw), a.overflow, spill_mem, kBounceStackLevels);
        MlpFwdSmem<K0> sm;
        sm.carve(smem_raw + (size_t)kGBlock * kBounceStackLevels * sizeof(uint2));
        __shared__ int s_tile;
        const bool fused = a.precomputed == nullptr;
        uint32_t tmem_base = 0, phase = 0;
        if (fused)
            tmem_base = mlp_fwd_setup<K0>(sm, net.mlp);
        const int n = *a.in_count;
        for (;;)
        {
            __syncthreads();
            if (threadIdx.x == 0)
                s_tile = atomicAdd(a.cursor, kTileRows);
            __syncthreads();
            const int base = s_tile;
            if (base >= n)
                break;
            const int i = base + threadIdx.x;
            const bool valid = i < n;
            GItem<D> it;
            if (valid)
                g_load<D>(a.in, (size_t)i, a.active, it);
            else
            {
                it.w.pixel = 0, it.w.ai = 0, it.w.sample = a.sample0, it.w.thp = 0.0f, it.w.onN = false, it.R_B = 0.0f;
                it.T[0] = it.T[1] = it.T[2] = 0.0f;
#pragma unroll
                for (int k = 0; k < D; ++k)
                    it.w.x[k] = it.w.nN[k] = 0.0f;
            }
            float raw[kOutPad];
            if (fused)
            {
                float enc[K0];
                encode_point<D, K0>(net, it.w.x, enc);
                mlp_fwd_tile<K0>(sm, tmem_base, phase, enc, raw);
            }
            else if (valid)
            {
                const float4 *src = reinterpret_cast<const float4 *>(a.precomputed + (size_t)i * kOutPad);
#pragma unroll
                for (int k = 0; k < kOutPad / 4; ++k)
                {
                    const float4 v = __ldg(src + k);
                    raw[4 * k] = v.x, raw[4 * k + 1] = v.y, raw[4 * k + 2] = v.z, raw[4 * k + 3] = v.w;
                }
            }
            if (valid)
            {
                float pdf;
                const Vec<D> dir = g_sample_direction<D>(c, a, it, raw, pdf);
                if (a.training && a.depth < (uint32_t)a.max_train_depth)
                    g_record<D>(a, it, dir, pdf, a.depth);
                if ((int)a.depth + 1 < a.max_depth)
                {
                    walk_advance<D>(c, it.w, dir, pdf, it.R_B, st);
                    g_store<D>(a.out, (size_t)i, it); // same slot: k_g_arrive of depth + 1 reads the moved items 1:1
                }
                else
                    g_finalize<D>(a, it);
            }
        }
        if (fused)
            mlp_fwd_teardown<K0>(sm, tmem_base);
    }

    // ---- walks that outlive the guided depths: uniform steps to completion (oneStepWalkImpl, integrator.cu:882-965).
    // Same persistent-lane scheme as k_g_arrive, but a lane keeps its walk until it ends: tail phase = contribution
    // stages of the point just reached, then uniform direction + walk record + move; lanes without a walk claim the
    // next out-shell item of the queue.
    template <int D, int F>
    __global__ void __launch_bounds__(kGBlock, g_min_ctas(D, F)) k_g_finish(const __grid_constant__ WalkCtx<D> c, const __grid_constant__ GArgs a)
    {
        extern __shared__ uint2 smem_stack[];
        SmemStack st;
        uint2 spill_mem[kSpillLevels];
        st.init(smem_stack, a.overflow, spill_mem);
        const int lane = threadIdx.x & 31;
        const int njobs = *a.in_count;
        GItem<D> it;
        NearestTrav<D> t;
        uint32_t depth = a.depth;
        bool has = false, exhausted = false, trav = false, need_contrib = false;
        float rd = dm_inf(); // closest distance at the point the walk is about to leave (upper-bounds the next query)
        bool win_known = false; // t.win is the closest primitive of the point the walk is about to leave
        unsigned int my_steps = 0;
        it.w.pixel = 0, it.w.ai = 0, it.w.sample = a.sample0, it.w.thp = 0.0f, it.w.onN = false, it.R_B = 0.0f;
        it.T[0] = it.T[1] = it.T[2] = 0.0f;
#pragma unroll
        for (int k = 0; k < D; ++k)
            it.w.x[k] = it.w.nN[k] = 0.0f;
        t.win = kInvalid, t.best = 0.0f;
        for (;;)
        {
            const unsigned trav_mask = __ballot_sync(kFullMask, trav);
            const unsigned pend_mask = __ballot_sync(kFullMask, !trav && (has || !exhausted));
            if (pend_mask != 0u && (__popc(pend_mask) >= a.refill || trav_mask == 0u))
            {
                if (!trav && has && need_contrib)
                {
                    PixelAccT<true> acc;
                    acc.init(reinterpret_cast<float *>(a.acc + (size_t)(it.w.sample - a.sample0 + (uint32_t)a.slot0) * a.n_active + it.w.ai));
                    acc.tot[0] = it.T[0], acc.tot[1] = it.T[1], acc.tot[2] = it.T[2];
                    rd = nearest_distance(t);
                    win_known = true;
                    const bool on = walk_step_contrib<D, F>(c, it.w, depth, st, acc, rd, t.win, 1.0f, it.R_B);
                    acc.flush();
                    it.T[0] = acc.tot[0], it.T[1] = acc.tot[1], it.T[2] = acc.tot[2];
                    need_contrib = false;
                    if (!on)
                    {
                        g_finalize<D>(a, it);
                        has = false;
                    }
                }
                const bool want = !trav && !has && !exhausted;
                const unsigned want_mask = __ballot_sync(kFullMask, want);
                if (want_mask != 0u)
                {
                    const int leader = __ffs(want_mask) - 1;
                    int base = 0;
                    if (lane == leader)
                        base = atomicAdd(a.cursor, __popc(want_mask));
                    base = __shfl_sync(kFullMask, base, leader);
                    if (want)
                    {
                        const int i = base + __popc(want_mask & ((1u << lane) - 1u));
                        if (i < njobs)
                        {
                            g_load<D>(a.in, (size_t)i, a.active, it);
                            depth = a.depth;
                            has = true;
                            rd = c.neu.nprims == 0 ? it.R_B : dm_inf(); // R_B = max(1e-4, R_D) when there is no Neumann mesh
                            win_known = false;
                        }
                        else
                            exhausted = true;
                    }
                }
                if (!trav && has)
                {
                    float pdf;
                    const Vec<D> dir = uniform_direction<D, F>(c, it.w, depth, pdf);
                    if (a.training && depth < (uint32_t)a.max_train_depth)
                        g_record<D>(a, it, dir, pdf, depth);
                    if ((int)depth + 1 < a.max_depth)
                    {
                        const Vec<D> x0 = it.w.x;
                        walk_advance<D, F>(c, it.w, dir, pdf, it.R_B, st);
                        ++depth;
                        ++my_steps;
                        need_contrib = true;
                        // bound: distance to the primitive that was closest a step ago when this lane knows it, else R_D + |step|
                        const float bound = (win_known && t.win != kInvalid) ? prim_distance(load_prim(c.dir.leaf_prim + t.win), it.w.x) * 1.00001f
                                                                             : (rd + vlen(vsub(it.w.x, x0))) * 1.00001f;
                        trav = nearest_begin<D, false>(c.dir, it.w.x, st, t, bound, c.eps, (F & kFeatNeumann) != 0);
                    }
                    else
                    {
                        g_finalize<D>(a, it);
                        has = false;
                    }
                }
                if (D == 2 && (F & kFeatNeumann) && (nearest_brute_warp<D>(c.dir, it.w.x, trav && t.phase == 4, t) >> lane & 1u))
                    trav = false;
            }
            else if (trav_mask == 0u)
                break;
            if (trav)
                trav = nearest_trip<D, false>(c.dir, it.w.x, st, t);
        }
        unsigned int ws = my_steps;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1)
            ws += __shfl_xor_sync(kFullMask, ws, o);
        if (lane == 0 && ws)
            atomicAdd(a.steps, (unsigned long long)ws);
    }

    // ---- validation-mode / probe kernels for the network -------------------------------------------------------
    // raw outputs of the MLP at the positions of queue items (or of an explicit point list), one thread per row, SIMT
    template <int D>
    __global__ void k_g_mlp_simt(const __grid_constant__ GuideNetView net, const float4 *q0, const float *pts, const int *count, int n_fixed, float *out)
    {
        const int n = count ? *count : n_fixed;
        for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        {
            Vec<D> x;
            if (q0)
            {
                const float4 a = q0[i];
                x[0] = a.x, x[1] = a.y;
                if (D == 3)
                    x[D - 1] = a.z;
            }
            else
            {
#pragma unroll
                for (int k = 0; k < D; ++k)
                    x[k] = pts[(size_t)i * D + k];
            }
            float enc[kHidden], o[kOutPad];
            encode_point<D, kHidden>(net, x, enc);
            mlp_forward_row_simt(net, enc, o);
            for (int k = 0; k < kOutPad; ++k)
                out[(size_t)i * kOutPad + k] = o[k];
        }
    }
    // the tcgen05 forward pass on an explicit point list (parity probe of mlp_fwd_tile)
    template <int D, int K0>
    __global__ void __launch_bounds__(kGBlock) k_g_infer_tc(const __grid_constant__ GuideNetView net, const float *pts, int n, float *out)
    {
        extern __shared__ __align__(1024) char smem_raw[];
        MlpFwdSmem<K0> sm;
        sm.carve(smem_raw);
        uint32_t phase = 0;
        const uint32_t tmem_base = mlp_fwd_setup<K0>(sm, net.mlp);
        for (int base = blockIdx.x * kTileRows; base < n; base += gridDim.x * kTileRows)
        {
            const int i = base + threadIdx.x;
            Vec<D> x;
#pragma unroll
            for (int k = 0; k < D; ++k)
                x[k] = i < n ? pts[(size_t)i * D + k] : 0.0f;
            float enc[K0], raw[kOutPad];
            encode_point<D, K0>(net, x, enc);
            mlp_fwd_tile<K0>(sm, tmem_base, phase, enc, raw);
            if (i < n)
            {
#pragma unroll
                for (int k = 0; k < kOutPad; ++k)
                    out[(size_t)i * kOutPad + k] = raw[k];
            }
        }
        mlp_fwd_teardown<K0>(sm, tmem_base);
    }

    // ---- walk records -> training set (generate_training_data, guided/train.h:422-471), deterministic order:
    // samples are laid out by (position in the active-pixel list, record depth) through an exclusive scan.
    template <int D>
    WOST_DEV bool g_sample_of_record(const WalkCtx<D> &c, const float *rec, const uint8_t *rec_onN, const float *total, int n_active, uint32_t ai, int d,
                                     Vec<D> &pos, TrainSample<D> &t)
    {
        using L = RecLayout<D>;
        const size_t plane = (size_t)kMaxTrainDepth * n_active, o = (size_t)d * n_active + ai;
#pragma unroll
        for (int k = 0; k < D; ++k)
        {
            pos[k] = rec[(L::kPos + k) * plane + o];
            t.dir[k] = rec[(L::kDir + k) * plane + o];
            t.nN[k] = rec[(L::kNrm + k) * plane + o];
        }
        t.dirPdf = rec[L::kPdf * plane + o];
        const float thp = rec[L::kThp * plane + o];
        t.onN = rec_onN[o] != 0;
        float li = 0.0f;
        bool nan = false;
#pragma unroll
        for (int ch = 0; ch < 3; ++ch)
        {
            const float sol = total[(size_t)ch * n_active + ai] - rec[(L::kBase + ch) * plane + o];
            float v = 0.0f;
            if (dm_abs(thp) > kMEps)
                v = dm_div(sol, thp);
            nan = nan || dm_isnan(v);
            li += dm_abs(v);
        }
        t.Li = dm_div(li, 3.0f);
        bool ok = aabb_contains<D>(c.aabb_min, c.aabb_max, pos);
        bool bad = nan || dm_isnan(t.dirPdf) || t.dirPdf == 0.0f;
#pragma unroll
        for (int k = 0; k < D; ++k)
            bad = bad || dm_isnan(pos[k]) || dm_isnan(t.dir[k]);
        return ok && !bad;
    }
    template <int D>
    __global__ void k_g_count_samples(const __grid_constant__ WalkCtx<D> c, int n_active, const float *rec,
                                      const uint8_t *rec_onN, const int *rec_count, const float *total, int *counts)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x;
        if (i > n_active)
            return;
        int cnt = 0;
        if (i < n_active)
        {
            const int n = min(rec_count[i], kMaxTrainDepth);
            for (int d = 0; d < n; ++d)
            {
                Vec<D> pos;
                TrainSample<D> t;
                cnt += g_sample_of_record<D>(c, rec, rec_onN, total, n_active, (uint32_t)i, d, pos, t) ? 1 : 0;
            }
        }
        counts[i] = cnt; // counts[n_active] = 0 so that the exclusive scan ends with the total
    }
    template <int D>
    __global__ void k_g_scatter_samples(const __grid_constant__ WalkCtx<D> c, int n_active, const float *rec,
                                        const uint8_t *rec_onN, const int *rec_count, const float *total, const int *offsets, int cap, float *smp,
                                        uint8_t *smp_onN, TrainCtl *ctl)
    {
        using L = RecLayout<D>;
        const int i = blockIdx.x * blockDim.x + threadIdx.x;
        if (i == 0)
            ctl->n_samples = min(offsets[n_active], cap);
        if (i >= n_active)
            return;
        const int n = min(rec_count[i], kMaxTrainDepth);
        int o = offsets[i];
        for (int d = 0; d < n; ++d)
        {
            Vec<D> pos;
            TrainSample<D> t;
            if (!g_sample_of_record<D>(c, rec, rec_onN, total, n_active, (uint32_t)i, d, pos, t))
                continue;
            if (o < cap)
            {
#pragma unroll
                for (int k = 0; k < D; ++k)
                {
                    smp[(size_t)(L::sPos + k) * cap + o] = pos[k];
                    smp[(size_t)(L::sDir + k) * cap + o] = t.dir[k];
                    smp[(size_t)(L::sNrm + k) * cap + o] = t.nN[k];
                }
                smp[(size_t)L::sPdf * cap + o] = t.dirPdf;
                smp[(size_t)L::sLi * cap + o] = t.Li;
                smp_onN[o] = t.onN ? 1 : 0;
            }
            ++o;
        }
    }

    // ---- training: forward + KL gradient + backward + weight gradients, all GEMMs on tcgen05 -------------------
    // (trainStepImpl, guided/integrator.cu:617-668: tcnn forward -> compute_dL_doutput_divergence -> tcnn backward)
    // One barrier-delimited GEMM: every thread has finished writing its operand rows / reading the accumulator;
    // thread 0 issues nk MMAs of K = 16 and commits; everybody waits for completion.
    WOST_DEV void tc_gemm(uint64_t *bar, uint32_t &phase, uint32_t tmem_d, uint32_t a_addr, uint32_t a_lbo, uint32_t a_sbo, uint32_t a_step,
                          uint32_t b_addr, uint32_t b_lbo, uint32_t b_sbo, uint32_t b_step, uint32_t idesc, int nk, bool accumulate)
    {
        tc::fence_proxy_async();
        tc::tc_fence_before();
        __syncthreads();
        if (threadIdx.x == 0)
        {
            tc::tc_fence_after();
            for (int k = 0; k < nk; ++k)
                tc::mma_f16(tmem_d, tc::make_desc(a_addr + k * a_step, a_lbo, a_sbo), tc::make_desc(b_addr + k * b_step, b_lbo, b_sbo), idesc,
                            (accumulate || k > 0) ? 1u : 0u);
            tc::mma_commit(bar);
        }
        tc::mbar_wait(bar, phase);
        phase ^= 1u;
        tc::tc_fence_after();
    }
    template <int K0>
    struct MlpTrainSmem
    {
        static constexpr int kWeightHalfs = kHidden * K0 + 2 * kHidden * kHidden + kOutPad * kHidden;
        static constexpr int kActBytes = kTileRows * kHidden * 2; // one 128 x 64 fp16 operand
        static constexpr int kA0Bytes = kTileRows * K0 * 2;
        static constexpr size_t kBytes = (size_t)kWeightHalfs * 2 + 3 * kActBytes + kA0Bytes + 4 * kActBytes + 64;
        __half *w, *a3, *a2, *a1, *a0, *dz4, *dz3, *dz2, *dz1;
        uint64_t *bar;
        uint32_t *tmem_slot;
        WOST_DEV void carve(char *base)
        {
            char *p = base;
            w = reinterpret_cast<__half *>(p), p += (size_t)kWeightHalfs * 2;
            a3 = reinterpret_cast<__half *>(p), p += kActBytes; // [a3 | a2] and [a1 | a0] are contiguous: stacked MN-major operands
            a2 = reinterpret_cast<__half *>(p), p += kActBytes;
            a1 = reinterpret_cast<__half *>(p), p += kActBytes;
            a0 = reinterpret_cast<__half *>(p), p += kA0Bytes;
            dz4 = reinterpret_cast<__half *>(p), p += kActBytes; // [dz4 | dz3], [dz2 | dz1]
            dz3 = reinterpret_cast<__half *>(p), p += kActBytes;
            dz2 = reinterpret_cast<__half *>(p), p += kActBytes;
            dz1 = reinterpret_cast<__half *>(p), p += kActBytes;
            bar = reinterpret_cast<uint64_t *>(p);
            tmem_slot = reinterpret_cast<uint32_t *>(bar + 1);
        }
    };
    struct TrainArgs
    {
        const float *smp;
        const uint8_t *smp_onN;
        int cap;
        TrainCtl *ctl;
        int iter;            // batch index within the frame
        int batch_size, min_batch, per_frame;
        int n_override;      // >= 0: probe call on exactly this many rows (wost_guided_train_batch)
        float loss_scale;
        float *grad;         // [n_params + 4]
        int n_params;
    };
    // read this thread's 64 fp32 accumulator columns, mask with the ReLU pattern of `act` (the forward activation of the
    // same layer) and write the fp16 result as the next operand `dz`
    WOST_DEV void tc_backprop_relu(uint32_t lane_base, const __half *act, __half *dz, int r)
    {
#pragma unroll
        for (int c16 = 0; c16 < kHidden / 16; ++c16)
        {
            float v[16];
            tc::tmem_ld16(lane_base + c16 * 16, v);
#pragma unroll
            for (int h = 0; h < 2; ++h)
            {
                const int chunk = c16 * 2 + h;
                const uint4 m = *reinterpret_cast<const uint4 *>(reinterpret_cast<const char *>(act) + (size_t)chunk * (kTileRows * 16) + (size_t)r * 16);
                const __half2 *mh = reinterpret_cast<const __half2 *>(&m);
#pragma unroll
                for (int j = 0; j < 4; ++j)
                {
                    if (!(__low2float(mh[j]) > 0.0f))
                        v[h * 8 + 2 * j] = 0.0f;
                    if (!(__high2float(mh[j]) > 0.0f))
                        v[h * 8 + 2 * j + 1] = 0.0f;
                }
                tc::store_chunk(dz, kTileRows, r, chunk, v + h * 8);
            }
        }
    }
    template <int D, int K0>
    __global__ void __launch_bounds__(kGBlock, 1) k_g_train(const __grid_constant__ GuideNetView net, const __grid_constant__ TrainArgs a)
    {
        using L = RecLayout<D>;
        extern __shared__ __align__(1024) char smem_raw[];
        MlpTrainSmem<K0> sm;
        sm.carve(smem_raw);
        // ---- which rows does this batch cover? (trainStepImpl :637-645)
        int row0, rows;
        {
            const int n = a.n_override >= 0 ? a.n_override : a.ctl->n_samples;
            const int nb = min(n / a.batch_size + 1, a.per_frame);
            row0 = a.iter * a.batch_size;
            rows = min(n - row0, a.batch_size);
            rows -= rows % kTileRows;
            const bool valid = a.iter < nb && rows > 0 && (a.n_override >= 0 || rows >= a.min_batch);
            if (blockIdx.x == 0 && threadIdx.x == 0)
            {
                a.ctl->batch_valid[a.iter] = valid ? 1 : 0;
                a.ctl->batch_rows[a.iter] = valid ? rows : 0;
                if (valid)
                    a.grad[a.n_params] = 1.0f;
            }
            if (!valid)
                return;
        }
        const int r = threadIdx.x;
        {
            uint64_t *wbar = sm.bar + 2;
            if (r == 0)
            {
                tc::mbar_init(sm.bar, 1);
                tc::mbar_init(wbar, 1);
                tc::fence_mbar_init();
                // forward weights: one TMA bulk copy (mlp_tc.cuh: bulk_g2s)
                constexpr uint32_t kWBytes = (uint32_t)MlpTrainSmem<K0>::kWeightHalfs * 2u;
                static_assert(kWBytes % 16u == 0u, "bulk copy size");
                tc::mbar_expect_tx(wbar, kWBytes);
                tc::bulk_g2s(sm.w, net.mlp, kWBytes, wbar);
            }
            if (r < 32)
                tc::tmem_alloc(sm.tmem_slot, 512);
            tc::tc_fence_before();
            __syncthreads();
            tc::tc_fence_after();
            tc::mbar_wait(wbar, 0);
        }
        const uint32_t tmem = *sm.tmem_slot;
        const uint32_t lane_base = tmem + ((uint32_t)(r & ~31) << 16);
        const uint32_t kColG43 = 64, kColG21 = 192;
        uint32_t phase = 0;
        const uint32_t AB = kTileRows * 16; // K-chunk stride of a 128-row operand
        const int off1 = kHidden * K0, off2 = off1 + kHidden * kHidden, off3 = off2 + kHidden * kHidden;
        const uint32_t w0 = tc::smem_u32(sm.w), w1 = tc::smem_u32(sm.w + off1), w2 = tc::smem_u32(sm.w + off2), w3 = tc::smem_u32(sm.w + off3);
        const float scale_over_n = dm_div(a.loss_scale, (float)rows);
        float loss_acc = 0.0f;
        bool did_any = false;
        for (int tile = blockIdx.x; tile < rows / kTileRows; tile += gridDim.x)
        {
            const size_t i = (size_t)row0 + (size_t)tile * kTileRows + r;
            Vec<D> pos;
            TrainSample<D> t;
#pragma unroll
            for (int k = 0; k < D; ++k)
            {
                pos[k] = a.smp[(size_t)(L::sPos + k) * a.cap + i];
                t.dir[k] = a.smp[(size_t)(L::sDir + k) * a.cap + i];
                t.nN[k] = a.smp[(size_t)(L::sNrm + k) * a.cap + i];
            }
            t.dirPdf = a.smp[(size_t)L::sPdf * a.cap + i];
            t.Li = a.smp[(size_t)L::sLi * a.cap + i];
            t.onN = a.smp_onN[i] != 0;
            // ---- forward, activations kept
            {
                float enc[K0];
                encode_point<D, K0>(net, pos, enc);
#pragma unroll
                for (int c = 0; c < K0 / 8; ++c)
                    tc::store_chunk(sm.a0, kTileRows, r, c, enc + c * 8);
            }
            const uint32_t id_f64 = tc::make_idesc(kTileRows, kHidden, 0, 0), id_f48 = tc::make_idesc(kTileRows, kOutPad, 0, 0);
            __half *acts[4] = {sm.a0, sm.a1, sm.a2, sm.a3};
#pragma unroll
            for (int l = 0; l < 3; ++l)
            {
                const int K = l == 0 ? K0 : kHidden;
                const uint32_t wl = l == 0 ? w0 : (l == 1 ? w1 : w2);
                tc_gemm(sm.bar, phase, tmem, tc::smem_u32(acts[l]), AB, 128, 2 * AB, wl, kHidden * 16, 128, 2 * kHidden * 16, id_f64, K / 16, false);
#pragma unroll
                for (int c16 = 0; c16 < kHidden / 16; ++c16)
                {
                    float v[16];
                    tc::tmem_ld16(lane_base + c16 * 16, v);
#pragma unroll
                    for (int j = 0; j < 16; ++j)
                        v[j] = fmaxf(v[j], 0.0f);
                    tc::store_chunk(acts[l + 1], kTileRows, r, c16 * 2, v);
                    tc::store_chunk(acts[l + 1], kTileRows, r, c16 * 2 + 1, v + 8);
                }
            }
            tc_gemm(sm.bar, phase, tmem, tc::smem_u32(sm.a3), AB, 128, 2 * AB, w3, kOutPad * 16, 128, 2 * kOutPad * 16, id_f48, kHidden / 16, false);
            // ---- loss gradient w.r.t. the raw outputs
            {
                float raw[kOutPad];
#pragma unroll
                for (int c16 = 0; c16 < kOutPad / 16; ++c16)
                    tc::tmem_ld16(lane_base + c16 * 16, raw + c16 * 16);
                float dl[kHidden];
                loss_acc += kl_dl_doutput<D>(raw, t, scale_over_n, dl);
#pragma unroll
                for (int k = kLobes * (D + 2) + 1; k < kHidden; ++k)
                    dl[k] = 0.0f;
#pragma unroll
                for (int k = 0; k < kLobes * (D + 2) + 1; ++k) // keep fp16 finite: one exploding sample must not poison the batch
                    dl[k] = dm_isnan(dl[k]) ? 0.0f : fminf(fmaxf(dl[k], -60000.0f), 60000.0f);
#pragma unroll
                for (int c = 0; c < kHidden / 8; ++c)
                    tc::store_chunk(sm.dz4, kTileRows, r, c, dl + c * 8);
            }
            // ---- backward through the layers: dA_l = dZ_{l+1} * W_l (B = the forward weight tile read MN-major)
            const uint32_t id_b64 = tc::make_idesc(kTileRows, kHidden, 0, 1), id_b0 = tc::make_idesc(kTileRows, K0, 0, 1);
            tc_gemm(sm.bar, phase, tmem, tc::smem_u32(sm.dz4), AB, 128, 2 * AB, w3, 128, kOutPad * 16, 256, id_b64, kOutPad / 16, false);
            tc_backprop_relu(lane_base, sm.a3, sm.dz3, r);
            tc_gemm(sm.bar, phase, tmem, tc::smem_u32(sm.dz3), AB, 128, 2 * AB, w2, 128, kHidden * 16, 256, id_b64, kHidden / 16, false);
            tc_backprop_relu(lane_base, sm.a2, sm.dz2, r);
            tc_gemm(sm.bar, phase, tmem, tc::smem_u32(sm.dz2), AB, 128, 2 * AB, w1, 128, kHidden * 16, 256, id_b64, kHidden / 16, false);
            tc_backprop_relu(lane_base, sm.a1, sm.dz1, r);
            tc_gemm(sm.bar, phase, tmem, tc::smem_u32(sm.dz1), AB, 128, 2 * AB, w0, 128, kHidden * 16, 256, id_b0, kHidden / 16, false);
            // ---- grid gradient (kernel_grid_backward, grid.h:395-540): scatter w * dL/d(encoded feature)
            {
                float de[K0];
#pragma unroll
                for (int c16 = 0; c16 < K0 / 16; ++c16)
                    tc::tmem_ld16(lane_base + c16 * 16, de + c16 * 16);
                float xn[D];
                normalize_coord<D>(net, pos, xn);
                float *gg = a.grad + net.n_mlp_params;
#pragma unroll
                for (int l = 0; l < K0 / kGridFeat; ++l)
                {
                    if (l >= net.n_levels)
                        break;
                    GridCell<D> cell;
                    grid_cell<D>(net, l, xn, cell);
                    // one 128-bit reduction per corner (red.global.add.v4.f32, sm_90+): the 4 features of a grid point are
                    // contiguous and 16-byte aligned -- 4x fewer atomics than tcnn's per-feature-pair adds (grid.h:437-446)
#pragma unroll
                    for (int corner = 0; corner < (1 << D); ++corner)
                    {
                        const float w = cell.w[corner];
                        atomicAdd(reinterpret_cast<float4 *>(gg + (size_t)cell.idx[corner] * kGridFeat),
                                  make_float4(w * de[l * kGridFeat], w * de[l * kGridFeat + 1], w * de[l * kGridFeat + 2], w * de[l * kGridFeat + 3]));
                    }
                }
            }
            // ---- weight gradients, accumulated in TMEM over all tiles of this CTA:
            //   [dZ4|dZ3]^T [a3|a2] -> dW3 (rows 0..47, cols 0..63) and dW2 (rows 64..127, cols 64..127)
            //   [dZ2|dZ1]^T [a1|a0] -> dW1 (rows 0..63, cols 0..63) and dW0 (rows 64..127, cols 64..64+K0-1)
            const uint32_t id_g43 = tc::make_idesc(kTileRows, 2 * kHidden, 1, 1), id_g21 = tc::make_idesc(kTileRows, kHidden + K0, 1, 1);
            tc_gemm(sm.bar, phase, tmem + kColG43, tc::smem_u32(sm.dz4), 128, AB, 256, tc::smem_u32(sm.a3), 128, AB, 256, id_g43, kTileRows / 16, did_any);
            tc_gemm(sm.bar, phase, tmem + kColG21, tc::smem_u32(sm.dz2), 128, AB, 256, tc::smem_u32(sm.a1), 128, AB, 256, id_g21, kTileRows / 16, did_any);
            did_any = true;
        }
        // ---- drain the weight gradients of this CTA into the global slab
        if (did_any)
        {
            const int o = r & 63;
            float *g43 = r < 64 ? a.grad + off3 + o * kHidden : a.grad + off2 + o * kHidden;
            const uint32_t c43 = kColG43 + (r < 64 ? 0 : kHidden);
            // tcgen05.ld is warp-collective (.sync.aligned): every lane loads, only the lanes that own a real row add
            const bool own43 = r >= 64 || o < kOutPad; // rows 48..63 of the padded dW3 block are padding
#pragma unroll
            for (int c16 = 0; c16 < kHidden / 16; ++c16)
            {
                float v[16];
                tc::tmem_ld16(lane_base + c43 + c16 * 16, v);
                if (own43)
                {
                    // 128-bit reductions (red.global.add.v4.f32): 4x fewer atomics than one per weight
#pragma unroll
                    for (int j = 0; j < 16; j += 4)
                        atomicAdd(reinterpret_cast<float4 *>(g43 + c16 * 16 + j), make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]));
                }
            }
            if (r < 64)
            {
#pragma unroll
                for (int c16 = 0; c16 < kHidden / 16; ++c16)
                {
                    float v[16];
                    tc::tmem_ld16(lane_base + kColG21 + c16 * 16, v);
#pragma unroll
                    for (int j = 0; j < 16; j += 4)
                        atomicAdd(reinterpret_cast<float4 *>(a.grad + off1 + o * kHidden + c16 * 16 + j), make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]));
                }
            }
            else
            {
#pragma unroll
                for (int c16 = 0; c16 < K0 / 16; ++c16)
                {
                    float v[16];
                    tc::tmem_ld16(lane_base + kColG21 + kHidden + c16 * 16, v);
#pragma unroll
                    for (int j = 0; j < 16; j += 4)
                        atomicAdd(reinterpret_cast<float4 *>(a.grad + o * K0 + c16 * 16 + j), make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]));
                }
            }
        }
#pragma unroll
        for (int s = 16; s > 0; s >>= 1)
            loss_acc += __shfl_xor_sync(kFullMask, loss_acc, s);
        if ((r & 31) == 0 && did_any)
            atomicAdd(&a.ctl->loss[a.iter], loss_acc);
        tc::tc_fence_before();
        __syncthreads();
        if (r < 32)
            tc::tmem_dealloc(tmem, 512);
    }

    // ---- optimiser: Adam nested in EMA, plus the fp16 working copies, in one pass over the parameter slab --------
    // (adam_step, optimizers/adam.h:48-118; ema_step_full_precision, optimizers/ema.h:44-60)
    WOST_HOSTDEV int canonical_index(int k0, int n_mlp, int i) // parameter i -> position in the fp16 working copy
    {
        if (i >= n_mlp)
            return i;
        int l = 0;
        if (i >= kHidden * k0)
            l = 1 + (i - kHidden * k0) / (kHidden * kHidden);
        const int off = l == 0 ? 0 : (kHidden * k0 + (l - 1) * kHidden * kHidden);
        const int K = l == 0 ? k0 : kHidden, N = l == 3 ? kOutPad : kHidden;
        const int o = (i - off) / K, k = (i - off) - o * K;
        return off + tile_off(N, o, k);
    }
    struct AdamArgs
    {
        int n_params, n_mlp, k0;
        AdamConfig cfg;
        float loss_scale;
        float *master, *m, *v, *ema;
        uint32_t *steps;
        const float *grad; // [n_params + 4], grad[n_params] = number of ranks that contributed
        __half *w16_train, *w16_infer;
        int *opt_steps;    // [0] optimiser steps so far (read), [1] written by thread 0 of block 0
    };
    __global__ void k_g_adam_ema(const AdamArgs a)
    {
        const float nvalid = a.grad[a.n_params];
        if (!(nvalid > 0.0f))
        {
            if (blockIdx.x == 0 && threadIdx.x == 0)
                a.opt_steps[1] = a.opt_steps[0];
            return;
        }
        const int step = a.opt_steps[0] + 1;
        if (blockIdx.x == 0 && threadIdx.x == 0)
            a.opt_steps[1] = step;
        const int i = blockIdx.x * blockDim.x + threadIdx.x;
        if (i >= a.n_params)
            return;
        float w = a.master[i];
        float g = a.grad[i] / (a.loss_scale * nvalid);
        const bool matrix = i < a.n_mlp;
        if (matrix || g != 0.0f) // grid parameters without gradient are left alone (adam.h:76-80)
        {
            if (matrix)
                g += a.cfg.l2_reg * w;
            const float m = a.m[i] = a.cfg.beta1 * a.m[i] + (1.0f - a.cfg.beta1) * g;
            const float v = a.v[i] = a.cfg.beta2 * a.v[i] + (1.0f - a.cfg.beta2) * g * g;
            const uint32_t ps = ++a.steps[i];
            const float lr = a.cfg.learning_rate * sqrtf(1.0f - powf(a.cfg.beta2, (float)ps)) / (1.0f - powf(a.cfg.beta1, (float)ps));
            const float eff = lr / (sqrtf(v) + a.cfg.epsilon);
            const float decayed = (1.0f - a.cfg.relative_decay * lr) * w - copysignf(a.cfg.absolute_decay * lr, w);
            w = decayed - eff * m;
            a.master[i] = w;
        }
        const int ci = canonical_index(a.k0, a.n_mlp, i);
        a.w16_train[ci] = __float2half_rn(w);
        const float debias_old = 1.0f - powf(a.cfg.ema_decay, (float)(step - 1));
        const float debias_new = 1.0f / (1.0f - powf(a.cfg.ema_decay, (float)step));
        const float e = (a.ema[i] * a.cfg.ema_decay * debias_old + w * (1.0f - a.cfg.ema_decay)) * debias_new;
        a.ema[i] = e;
        a.w16_infer[ci] = __float2half_rn(e);
    }
    // (re)build both fp16 working copies from the fp32 slabs (after initialisation / wost_guided_set_params)
    __global__ void k_g_sync_params(int n_params, int n_mlp, int k0, const float *master, const float *infer_src, __half *w16_train, __half *w16_infer)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x;
        if (i >= n_params)
            return;
        const int ci = canonical_index(k0, n_mlp, i);
        w16_train[ci] = __float2half_rn(master[i]);
        w16_infer[ci] = __float2half_rn(infer_src[i]);
    }
    __global__ void k_g_frame_ctl(TrainCtl *ctl, int *opt_steps, int last)
    {
        // start of a training frame: carry the optimiser step count over, clear the per-batch flags
        (void)opt_steps, (void)last;
        for (int i = 0; i < 8; ++i)
        {
            ctl->batch_valid[i] = ctl->batch_rows[i] = 0;
            ctl->loss[i] = 0.0f;
        }
    }
    __global__ void k_g_shift_step(int *opt_steps)
    {
        opt_steps[0] = opt_steps[1];
    }

    template <int D>
    __global__ void k_vmm_probe(const float *raw, const float *dir, const float *nrm, const uint8_t *onN, int n, uint32_t seed, float *pdf, float *grad,
                                float *sampled)
    {
        const int i = blockIdx.x * blockDim.x + threadIdx.x;
        if (i >= n)
            return;
        constexpr int kOut = kLobes * (D + 2) + 1;
        float r[kOut];
        for (int k = 0; k < kOut; ++k)
            r[k] = raw[(size_t)i * kOut + k];
        VMM<D> m;
        m.decode(r);
        Vec<D> w, nn;
        for (int k = 0; k < D; ++k)
            w[k] = dir[(size_t)i * D + k], nn[k] = nrm[(size_t)i * D + k];
        float g[kLobes * (D + 2)];
        pdf[i] = vmm_grad_prob<D>(m, w, onN[i] != 0, nn, g);
        for (int k = 0; k < kLobes * (D + 2); ++k)
            grad[(size_t)i * kLobes * (D + 2) + k] = g[k];
        const Philox4 q = wost_rng(seed, (uint32_t)i, 0u, 0u, WOST_RNG_SELECT, 0);
        const int lobe = m.pick(u32_to_unit_float(q.y));
        const Vec<D> s = lobe_sample(m.kappa[lobe], m.mu[lobe], seed, (uint32_t)i, 0u, 0u);
        for (int k = 0; k < D; ++k)
            sampled[(size_t)i * D + k] = s[k];
    }

    // =============================================================================================== host side
    static float u01(uint32_t a, uint32_t b, uint32_t seed, uint32_t stream)
    {
        return u32_to_unit_float(philox4x32_10(a, b, 0u, 0u, seed, stream).x);
    }
    static void net_fill_view(GuideNet &n, const wost_problem *p)
    {
        GuideNetView v;
        memset(&v, 0, sizeof(v));
        v.dim = n.dim;
        v.n_levels = n.n_levels;
        v.k0 = n.k0;
        const float log2s = std::log2(n.per_level_scale);
        uint32_t off = 0;
        for (int l = 0; l < n.n_levels; ++l)
        {
            // grid_scale / grid_resolution (grid.h:193-202) and the per-level sizes of GridEncodingTemplated (:982-1012)
            const float scale = std::exp2((float)l * log2s) * (float)n.base_res - 1.0f;
            const uint32_t res = (uint32_t)std::ceil(scale) + 1u;
            uint64_t pts = 1;
            for (int d = 0; d < n.dim; ++d)
                pts *= res;
            pts = (pts + 7u) / 8u * 8u;
            v.scale[l] = scale;
            v.res[l] = res;
            v.off[l] = off;
            off += (uint32_t)pts;
        }
        v.off[n.n_levels] = off;
        n.n_grid = (int)off * kGridFeat;
        n.n_mlp = kHidden * n.k0 + 2 * kHidden * kHidden + kOutPad * kHidden;
        n.n_params = n.n_mlp + n.n_grid;
        v.n_mlp_params = n.n_mlp;
        v.n_grid_params = n.n_grid;
        // normalizeSpatialCoord: AABB inflated by 0.5 % of its diagonal length (train.h:148-155)
        float d2 = 0.0f;
        for (int k = 0; k < n.dim; ++k)
        {
            const float e = p->desc.aabb_max[k] - p->desc.aabb_min[k];
            d2 += e * e;
        }
        const float infl = std::sqrt(d2) * 0.005f;
        for (int k = 0; k < 3; ++k)
        {
            v.lo[k] = p->desc.aabb_min[k] - infl;
            v.hi[k] = p->desc.aabb_max[k] + infl;
        }
        n.view_train = v;
        n.view_infer = v;
    }
    static wost_status net_alloc_and_init(wost_integrator *it)
    {
        GuideNet &n = *it->net;
        const size_t P = (size_t)n.n_params;
        if (n.master.alloc(P) != cudaSuccess || n.adam_m.alloc(P) != cudaSuccess || n.adam_v.alloc(P) != cudaSuccess || n.ema.alloc(P) != cudaSuccess ||
            n.grad.alloc(P + 4) != cudaSuccess || n.steps.alloc(P) != cudaSuccess || n.w16_train.alloc(P) != cudaSuccess ||
            n.w16_infer.alloc(P) != cudaSuccess || n.ctl.alloc(1) != cudaSuccess)
        {
            set_last_error("guiding network: device allocation failed");
            return WOST_ERR_CUDA;
        }
        n.view_train.mlp = n.w16_train.ptr;
        n.view_train.grid = n.w16_train.ptr + n.n_mlp;
        n.view_infer.mlp = n.w16_infer.ptr;
        n.view_infer.grid = n.w16_infer.ptr + n.n_mlp;
        // initialize_params: Xavier-uniform matrices (gpu_matrix.h:291-306), grid uniform in [-1e-4, 1e-4] (grid.h:1361)
        std::vector<float> h(P);
        size_t o = 0;
        for (int l = 0; l < 4; ++l)
        {
            const int K = l == 0 ? n.k0 : kHidden, N = l == 3 ? kOutPad : kHidden;
            const float s = std::sqrt(6.0f / (float)(K + N));
            for (int i = 0; i < K * N; ++i, ++o)
                h[o] = (2.0f * u01((uint32_t)i, (uint32_t)l, n.seed, 1u) - 1.0f) * s;
        }
        for (int i = 0; i < n.n_grid; ++i, ++o)
            h[o] = (2.0f * u01((uint32_t)i, 7u, n.seed, 2u) - 1.0f) * 1e-4f;
        return wost_guided_set_params(it, h.data());
    }
} // namespace wost

using namespace wost;

namespace wost
{
    wost_status guided_alloc(wost_integrator *it)
    {
        const int D = it->dim;
        const size_t cap = it->queue_capacity;
        for (int i = 0; i < 2; ++i)
        {
            for (int j = 0; j < 3; ++j)
                if (it->gq[i][j].alloc(cap) != cudaSuccess)
                    return WOST_ERR_CUDA;
            if (it->gq3[i].alloc(cap) != cudaSuccess)
                return WOST_ERR_CUDA;
        }
        GuidedBuffers &g = it->gb;
        const size_t npix = (size_t)std::max(1, it->n_active); // walk records are indexed by active rank: sized by the shard, not the frame
        const int planes = 3 * D + 5;
        g.cap = (int)std::min<size_t>((size_t)kMaxTrainDepth * std::max(1, it->n_active), (size_t)1 << 30);
        if (g.rec.alloc((size_t)planes * kMaxTrainDepth * npix) != cudaSuccess || g.rec_onN.alloc((size_t)kMaxTrainDepth * npix) != cudaSuccess ||
            g.rec_count.alloc(npix) != cudaSuccess || g.total.alloc(3 * npix) != cudaSuccess || g.sample_off.alloc((size_t)it->n_active + 1) != cudaSuccess ||
            g.smp.alloc((size_t)(3 * D + 2) * g.cap) != cudaSuccess || g.smp_onN.alloc((size_t)g.cap) != cudaSuccess)
            return WOST_ERR_CUDA;
        g.scan_tmp_bytes = 0;
        cub::DeviceScan::ExclusiveSum(nullptr, g.scan_tmp_bytes, g.sample_off.ptr, g.sample_off.ptr, it->n_active + 1);
        if (g.scan_tmp.alloc(g.scan_tmp_bytes + 16) != cudaSuccess)
            return WOST_ERR_CUDA;
        if (it->s.mlp_impl == 1 && g.infer_out.alloc(cap * kOutPad) != cudaSuccess)
            return WOST_ERR_CUDA;
        cudaMemset(g.rec_count.ptr, 0, sizeof(int) * npix);
        cudaMemset(g.total.ptr, 0, sizeof(float) * 3 * npix);
        return WOST_OK;
    }

    template <int D, int K0>
    static void launch_bounce(wost_integrator *it, const WalkCtx<D> &ctx, const GArgs &a, int grid)
    {
        const size_t smem = stack_smem_bytes(kGBlock, kBounceStackLevels) + MlpFwdSmem<K0>::kBytes;
        static bool attr_done = false; // per instantiation
        if (!attr_done)
        {
            cudaFuncSetAttribute(k_g_bounce<D, K0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            attr_done = true;
        }
        k_g_bounce<D, K0><<<grid, kGBlock, smem, it->stream>>>(ctx, a, it->net->view_infer);
    }
    template <int D, int K0>
    static int bounce_grid()
    {
        const size_t smem = stack_smem_bytes(kGBlock, kBounceStackLevels) + MlpFwdSmem<K0>::kBytes;
        cudaFuncSetAttribute(k_g_bounce<D, K0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        int occ = 1;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_g_bounce<D, K0>, kGBlock, smem) != cudaSuccess)
            occ = 1;
        return sm_count() * std::max(2, std::min(occ, 512 / kTmemColsFwd)); // >= 2: CTAs beyond residency just find the cursor exhausted
    }
    template <int D, int K0>
    static wost_status launch_train(wost_integrator *it, const TrainArgs &ta)
    {
        const size_t smem = MlpTrainSmem<K0>::kBytes;
        WOST_CUDA_TRY(cudaFuncSetAttribute(k_g_train<D, K0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k_g_train<D, K0><<<sm_count(), kGBlock, smem, it->stream>>>(it->net->view_train, ta);
        return WOST_OK;
    }
    static wost_status launch_adam(wost_integrator *it, float loss_scale, int *opt_steps)
    {
        GuideNet &n = *it->net;
        AdamArgs aa;
        aa.n_params = n.n_params, aa.n_mlp = n.n_mlp, aa.k0 = n.k0;
        aa.cfg = n.opt;
        aa.loss_scale = loss_scale;
        aa.master = n.master.ptr, aa.m = n.adam_m.ptr, aa.v = n.adam_v.ptr, aa.ema = n.ema.ptr;
        aa.steps = n.steps.ptr;
        aa.grad = n.grad.ptr;
        aa.w16_train = n.w16_train.ptr, aa.w16_infer = n.w16_infer.ptr;
        aa.opt_steps = opt_steps;
        k_g_adam_ema<<<div_up(n.n_params, 256), 256, 0, it->stream>>>(aa);
        k_g_shift_step<<<1, 1, 0, it->stream>>>(opt_steps);
        WOST_CUDA_TRY(cudaGetLastError());
        return WOST_OK;
    }
#define WOST_DISPATCH_K0(D_, k0, ...)                          \
    switch (k0)                                                \
    {                                                          \
    case 16: { constexpr int K0_ = 16; __VA_ARGS__; break; }   \
    case 32: { constexpr int K0_ = 32; __VA_ARGS__; break; }   \
    default: { constexpr int K0_ = 64; __VA_ARGS__; break; }   \
    }

    // one training frame: records -> samples -> <= batches_per_frame optimiser steps (trainStepImpl, :617-668)
    template <int D>
    static wost_status train_frame(wost_integrator *it, const WalkCtx<D> &ctx)
    {
        GuideNet &n = *it->net;
        GuidedBuffers &g = it->gb;
        cudaStream_t st = it->stream;
        int bs = it->s.train_batch_size > 0 ? it->s.train_batch_size : 524288;
        int minb = it->s.train_min_batch > 0 ? it->s.train_min_batch : 65536;
        const int per_frame = std::min(8, it->s.train_batches_per_frame > 0 ? it->s.train_batches_per_frame : 5);
        // data-parallel training (gradient all-reduce installed): one optimiser step consumes the UNION of the ranks' batches,
        // so every rank takes 1/num_shards of the global batch -- the same number of optimiser steps per frame and the same
        // global batch size as a single-GPU run. The step count must be identical on all ranks (collectives): it is sized from
        // the frame-wide active-pixel count, which every rank knows.
        const int shards = (it->allreduce && it->s.num_shards > 1) ? it->s.num_shards : 1;
        const long long cap_frame = (long long)kMaxTrainDepth * std::max(1, it->allreduce ? it->n_active_frame : it->n_active);
        const int max_frame_batches = (int)std::min<long long>(per_frame, cap_frame / bs + 1);
        if (shards > 1)
        {
            bs = std::max(128, bs / shards / 128 * 128);
            minb = std::max(128, minb / shards / 128 * 128);
        }
        int *opt_steps = &n.ctl.ptr->opt_steps[0];
        k_g_frame_ctl<<<1, 1, 0, st>>>(n.ctl.ptr, opt_steps, 0);
        k_g_count_samples<D><<<div_up(it->n_active + 1, 256), 256, 0, st>>>(ctx, it->n_active, g.rec.ptr, g.rec_onN.ptr,
                                                                               g.rec_count.ptr, g.total.ptr, g.sample_off.ptr);
        size_t tb = g.scan_tmp_bytes;
        WOST_CUDA_TRY(cub::DeviceScan::ExclusiveSum(g.scan_tmp.ptr, tb, g.sample_off.ptr, g.sample_off.ptr, it->n_active + 1, st));
        k_g_scatter_samples<D><<<div_up(std::max(1, it->n_active), 256), 256, 0, st>>>(ctx, it->n_active, g.rec.ptr, g.rec_onN.ptr,
                                                                             g.rec_count.ptr, g.total.ptr, g.sample_off.ptr, g.cap, g.smp.ptr, g.smp_onN.ptr, n.ctl.ptr);
        it->launches += 4;
        // with a gradient all-reduce every rank must issue the same number of collectives, whatever its own sample count
        const int max_batches = it->allreduce ? max_frame_batches : std::min(per_frame, g.cap / bs + 1);
        for (int iter = 0; iter < max_batches; ++iter)
        {
            WOST_CUDA_TRY(cudaMemsetAsync(n.grad.ptr, 0, sizeof(float) * ((size_t)n.n_params + 4), st));
            TrainArgs ta;
            ta.smp = g.smp.ptr, ta.smp_onN = g.smp_onN.ptr, ta.cap = g.cap;
            ta.ctl = n.ctl.ptr;
            ta.iter = iter;
            ta.batch_size = bs, ta.min_batch = minb, ta.per_frame = per_frame;
            ta.n_override = -1;
            ta.loss_scale = 128.0f; // TRAIN_LOSS_SCALE (parameters.h:15)
            ta.grad = n.grad.ptr;
            ta.n_params = n.n_params;
            wost_status rc = WOST_OK;
            WOST_DISPATCH_K0(D, n.k0, (rc = launch_train<D, K0_>(it, ta)));
            if (rc != WOST_OK)
                return rc;
            if (it->allreduce && it->allreduce(it->allreduce_user, n.grad.ptr, (uint64_t)n.n_params + 4, (void *)st) != 0)
            {
                set_last_error("gradient all-reduce callback failed");
                return WOST_ERR_INVALID;
            }
            rc = launch_adam(it, ta.loss_scale, opt_steps);
            if (rc != WOST_OK)
                return rc;
            it->launches += 3;
        }
        return WOST_OK;
    }

    template <int D, int F>
    static wost_status solve_guided_f(wost_integrator *it, uint64_t *walk_steps);
    // arrive / finish kernels compiled for the features the problem has (walk.cuh: kFeatNeumann / kFeatSource)
    template <int D>
    static wost_status solve_guided_impl(wost_integrator *it, uint64_t *walk_steps)
    {
        if (it->problem->neumann != nullptr && !it->problem->neumann_flux_enabled && it->problem->d_source.ptr == nullptr)
            return solve_guided_f<D, kFeatNeumann>(it, walk_steps); // data/ladybug, data/fille: reflecting boundary, no source
        if (it->problem->neumann != nullptr)
            return solve_guided_f<D, kFeatAll>(it, walk_steps);
        if (it->problem->d_source.ptr != nullptr)
            return solve_guided_f<D, kFeatSource>(it, walk_steps);
        return solve_guided_f<D, 0>(it, walk_steps);
    }
    template <int D, int F>
    static wost_status solve_guided_f(wost_integrator *it, uint64_t *walk_steps)
    {
        if (!it->net)
        {
            set_last_error("guided integrator: wost_reset_network has not been called");
            return WOST_ERR_CONFIG;
        }
        WalkCtx<D> ctx;
        it->problem->fill_ctx(ctx);
        ctx.frame_w = it->s.frame_w, ctx.frame_h = it->s.frame_h, ctx.max_depth = it->s.max_walking_depth, ctx.eps = it->s.epsilon_shell, ctx.seed = it->s.seed;
        GuideNet &n = *it->net;
        GuidedBuffers &g = it->gb;
        cudaStream_t st = it->stream;
        const int spp = it->s.samples_per_pixel, maxd = it->s.max_walking_depth;
        const int S = it->spp_batch;
        const size_t stack_smem = stack_smem_bytes(kGBlock);
        int occ = 1;
        WOST_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_g_arrive<D, false, F>, kGBlock, stack_smem));
        const int grid_full = sm_count() * std::max(1, occ);
        int grid_bounce = sm_count();
        // Launches over few items (training batches: one sample per pixel) leave ~2 items per resident lane: the persistent lanes
        // run out of refills early and the launch ends in a long tail of sparsely filled warps. Fewer, longer-lived lanes
        // (>= items_per_lane items each) trade latency hiding for a shorter tail.
        static const int items_per_lane = []
        {
            const char *e = getenv("WOST_G_ITEMS_PER_LANE");
            return e ? std::max(1, atoi(e)) : 1;
        }();
        auto grid_for = [&](long long max_items)
        { return (int)std::max<long long>(sm_count(), std::min<long long>(grid_full, (max_items + (long long)kGBlock * items_per_lane - 1) / ((long long)kGBlock * items_per_lane))); };
        WOST_DISPATCH_K0(D, n.k0, (grid_bounce = bounce_grid<D, K0_>()));
        const int max_train_depth = std::min(kMaxTrainDepth, it->s.max_train_depth > 0 ? it->s.max_train_depth : 3);

        WOST_CUDA_TRY(cudaMemsetAsync(it->d_acc.ptr, 0, sizeof(float4) * (size_t)S * std::max(1, it->n_active), st));
        WOST_CUDA_TRY(cudaMemsetAsync(it->d_steps.ptr, 0, sizeof(unsigned long long), st));
        it->launches = 0, it->step_launches = 0, it->step_kernel_ms = 0.0f;
        it->train_samples = 0, it->train_steps = 0;
        it->slot_n.assign((size_t)S, 0);
        // optional per-kernel-class device timing (WOST_PROFILE=1): CUDA events around every launch of the solve
        const bool prof = getenv("WOST_PROFILE") != nullptr;
        size_t prof_used = 0;
        std::vector<int> prof_cls;
        auto prof_mark = [&](int cls)
        {
            if (!prof)
                return;
            if (prof_used == it->prof_ev.size())
            {
                cudaEvent_t e;
                cudaEventCreate(&e);
                it->prof_ev.push_back(e);
            }
            cudaEventRecord(it->prof_ev[prof_used++], st);
            prof_cls.push_back(cls);
        };
        std::vector<std::pair<int, int>> batches; // (first sample, count)
        for (int s0 = 0; s0 < spp;)
        {
            const bool training = s0 < it->s.train_spp_count;
            int Sb = training ? 1 : it->frame_batch(s0, std::min(S, spp - s0)); // cut after every dumped sample index (progressive metrics)
            batches.emplace_back(s0, Sb);
            s0 += Sb;
        }
        while (it->batch_ev.size() < 2 * batches.size())
        {
            cudaEvent_t e;
            WOST_CUDA_TRY(cudaEventCreate(&e));
            it->batch_ev.push_back(e);
        }
        if (it->n_active > 0 && !batches.empty())
        {
            if (it->d_first.count < (size_t)it->n_active && it->d_first.alloc((size_t)it->n_active) != cudaSuccess)
            {
                set_last_error("wost_solve: first-hit cache allocation failed");
                return WOST_ERR_CUDA;
            }
            const int grid = std::max(1, std::min(grid_full, div_up(it->n_active, kGBlock)));
            k_first_hit<D, false><<<grid, kGBlock, stack_smem, st>>>(ctx, it->d_active.ptr, it->n_active, it->d_first.ptr, it->d_overflow.ptr);
            ++it->launches;
        }
        // A rank without active pixels (all its row tiles masked, or fewer tiles than ranks) launches no walk kernels but still
        // takes part in every gradient all-reduce of the training phase: it contributes a zero slab with valid = 0.
        const bool walks = it->n_active > 0;
        for (size_t bi = 0; (walks || it->allreduce) && bi < batches.size(); ++bi)
        {
            const int s0 = batches[bi].first, Sb = batches[bi].second;
            const bool training = s0 < it->s.train_spp_count;
            if (!walks)
            {
                if (training)
                {
                    WOST_CUDA_TRY(cudaMemsetAsync(g.rec_count.ptr, 0, sizeof(int), st));
                    const wost_status rc = train_frame<D>(it, ctx);
                    if (rc != WOST_OK)
                        return rc;
                }
                continue;
            }
            // solveImpl :991-996: after trainSppCount samples switch to the guiding-phase settings
            const int maxGuided = training ? it->s.max_guided_depth_training : it->s.max_guided_depth_guiding;
            const float ufrac = training ? it->s.uniform_fraction_training : it->s.uniform_fraction_guiding;
            const int G = std::max(0, std::min(maxGuided, maxd));
            WOST_CUDA_TRY(cudaMemsetAsync(it->d_counts.ptr, 0, sizeof(int) * (size_t)(3 * (maxd + 2)), st));
            WOST_CUDA_TRY(cudaMemsetAsync(it->d_cursor64.ptr, 0, sizeof(unsigned long long), st));
            if (training)
                WOST_CUDA_TRY(cudaMemsetAsync(g.rec_count.ptr, 0, sizeof(int) * (size_t)std::max(1, it->n_active), st));
            WOST_CUDA_TRY(cudaEventRecord(it->batch_ev[2 * bi], st));
            GArgs a;
            memset(&a, 0, sizeof(a));
            a.steps = it->d_steps.ptr;
            a.active = it->d_active.ptr, a.n_active = it->n_active;
            a.first = it->d_first.ptr;
            a.sample0 = (uint32_t)s0, a.batch_samples = Sb;
            a.acc = it->d_acc.ptr, a.max_depth = maxd, a.overflow = it->d_overflow.ptr;
            a.slot0 = training ? s0 % S : 0; // a training batch is one sample: rotate it through the S accumulator slots
            for (int j = 0; j < Sb; ++j)
                ++it->slot_n[(size_t)(a.slot0 + j)];
            a.training = training ? 1 : 0;
            a.refill = default_refill_threshold(D, F);
            a.max_train_depth = max_train_depth;
            a.uniform_fraction = ufrac;
            a.rec = g.rec.ptr, a.rec_onN = g.rec_onN.ptr, a.rec_count = g.rec_count.ptr, a.total = g.total.ptr;
            auto queue = [&](int i)
            { return GQueue{it->gq[i][0].ptr, it->gq[i][1].ptr, it->gq[i][2].ptr, it->gq3[i].ptr}; };
            int *counts = it->d_counts.ptr, *cur_b = counts + (maxd + 2), *cur_a = counts + 2 * (maxd + 2);
            // depth 0: evaluation points -> out-shell items of depth 0 in queue 0
            a.in = queue(1), a.out = queue(0);
            a.in_count = nullptr, a.out_count = counts + 0, a.cursor = cur_a + 0, a.cursor64 = it->d_cursor64.ptr, a.depth = 0;
            {
                const long long n0 = (long long)it->n_active * Sb;
                const int grid = (int)std::min<long long>(grid_full, (n0 + kGBlock - 1) / kGBlock);
                prof_mark(0);
                k_g_arrive<D, true, F><<<std::max(1, grid), kGBlock, stack_smem, st>>>(ctx, a);
                ++it->launches, ++it->step_launches;
            }
            // out-shell items of depth d live in queue d & 1
            for (int d = 0; d < G; ++d)
            {
                a.in = queue(d & 1), a.out = queue(d & 1); // the bounce moves its items in place
                a.in_count = counts + d, a.out_count = nullptr, a.cursor = cur_b + d, a.depth = (uint32_t)d;
                a.precomputed = nullptr;
                if (it->s.mlp_impl == 1)
                {
                    k_g_mlp_simt<D><<<sm_count() * 8, 128, 0, st>>>(n.view_infer, a.in.q0, nullptr, a.in_count, 0, g.infer_out.ptr);
                    a.precomputed = g.infer_out.ptr;
                    ++it->launches;
                }
                prof_mark(1);
                WOST_DISPATCH_K0(D, n.k0, (launch_bounce<D, K0_>(it, ctx, a, grid_bounce)));
                ++it->launches, ++it->step_launches;
                if (d + 1 < maxd)
                {
                    a.in = queue(d & 1), a.out = queue((d + 1) & 1);
                    a.in_count = counts + d, a.out_count = counts + d + 1, a.cursor = cur_a + d + 1, a.depth = (uint32_t)(d + 1);
                    prof_mark(0);
                    k_g_arrive<D, false, F><<<grid_for((long long)it->n_active * Sb), kGBlock, stack_smem, st>>>(ctx, a);
                    ++it->launches, ++it->step_launches;
                }
            }
            if (G < maxd)
            {
                a.in = queue(G & 1), a.out = queue((G + 1) & 1);
                a.in_count = counts + G, a.out_count = nullptr, a.cursor = cur_b + G, a.depth = (uint32_t)G;
                prof_mark(2);
                k_g_finish<D, F><<<grid_for((long long)it->n_active * Sb), kGBlock, stack_smem, st>>>(ctx, a);
                ++it->launches, ++it->step_launches;
            }
            prof_mark(3);
            WOST_CUDA_TRY(cudaEventRecord(it->batch_ev[2 * bi + 1], st));
            WOST_CUDA_TRY(cudaGetLastError());
            if (training)
            {
                prof_mark(4);
                const wost_status rc = train_frame<D>(it, ctx);
                if (rc != WOST_OK)
                    return rc;
            }
            prof_mark(3);
            if (it->frame_cb)
            {
                const wost_status frc = emit_frames(it, s0 + Sb - 1);
                if (frc != WOST_OK)
                    return frc;
            }
            // slot s of every batch adds to acc[s]; k_resolve sums all S slots at the end
        }
        launch_resolve(it, S, (float)spp);
        ++it->launches;
        WOST_CUDA_TRY(cudaGetLastError());
        unsigned long long h_steps = 0;
        int h_over = 0;
        TrainCtl h_ctl;
        memset(&h_ctl, 0, sizeof(h_ctl));
        WOST_CUDA_TRY(cudaMemcpyAsync(&h_steps, it->d_steps.ptr, sizeof(h_steps), cudaMemcpyDeviceToHost, st));
        WOST_CUDA_TRY(cudaMemcpyAsync(&h_over, it->d_overflow.ptr, sizeof(int), cudaMemcpyDeviceToHost, st));
        WOST_CUDA_TRY(cudaMemcpyAsync(&h_ctl, n.ctl.ptr, sizeof(h_ctl), cudaMemcpyDeviceToHost, st));
        WOST_CUDA_TRY(cudaEventRecord(it->ev1, st));
        WOST_CUDA_TRY(cudaStreamSynchronize(st));
        if (walk_steps)
            *walk_steps = h_steps;
        for (size_t bi = 0; walks && bi < batches.size(); ++bi)
        {
            float f = 0.0f;
            WOST_CUDA_TRY(cudaEventElapsedTime(&f, it->batch_ev[2 * bi], it->batch_ev[2 * bi + 1]));
            it->step_kernel_ms += f;
        }
        for (int k = 0; k < 8; ++k)
            it->prof_ms[k] = 0.0f;
        for (size_t k = 0; k + 1 < prof_used; ++k)
        {
            float f = 0.0f;
            cudaEventElapsedTime(&f, it->prof_ev[k], it->prof_ev[k + 1]);
            it->prof_ms[prof_cls[k]] += f; // the time from a mark to the next one belongs to that mark's class
        }
        it->train_steps = (uint64_t)h_ctl.opt_steps[0];
        it->train_samples = (uint64_t)h_ctl.n_samples;
        for (int i = 7; i >= 0; --i)
            if (h_ctl.batch_valid[i])
            {
                it->last_loss = h_ctl.loss[i] / (float)std::max(1, h_ctl.batch_rows[i]);
                break;
            }
        if (h_over)
        {
            cudaMemsetAsync(it->d_overflow.ptr, 0, sizeof(int), st);
            set_last_error("LBVH traversal stack overflow during solve");
            return WOST_ERR_OVERFLOW;
        }
        return WOST_OK;
    }
    wost_status solve_guided(wost_integrator *it, uint64_t *walk_steps)
    {
        return it->dim == 2 ? solve_guided_impl<2>(it, walk_steps) : solve_guided_impl<3>(it, walk_steps);
    }
} // namespace wost

// ======================================================================================================= C ABI
extern "C"
{
    wost_status wost_reset_network(wost_integrator *it, const char *tcnn_json)
    {
        if (!it || !tcnn_json || it->kind != WOST_GUIDED)
        {
            set_last_error("wost_reset_network: needs a guided integrator and a JSON string");
            return WOST_ERR_INVALID;
        }
        std::unique_ptr<GuideNet> n(new GuideNet());
        try
        {
            const wjson::Value cfg = wjson::parse(tcnn_json);
            const wjson::Value &enc = cfg.at("encoding");
            const wjson::Value &mlp = cfg.at("network");
            const std::string otype = enc.string_or("otype", "Grid"), gtype = enc.string_or("type", otype == "DenseGrid" ? "Dense" : "Hash");
            if (gtype != "Dense" || enc.string_or("interpolation", "Linear") != "Linear" || (int)enc.number_or("n_features_per_level", 2) != kGridFeat)
            {
                set_last_error("wost_reset_network: only the DenseGrid encoding with Linear interpolation and 4 features per level is built");
                return WOST_ERR_UNSUPPORTED;
            }
            if ((int)mlp.number_or("n_neurons", 128) != kHidden || (int)mlp.number_or("n_hidden_layers", 5) != 3 ||
                mlp.string_or("activation", "ReLU") != "ReLU" || mlp.string_or("output_activation", "None") != "None")
            {
                set_last_error("wost_reset_network: only the 64-neuron, 3-hidden-layer ReLU MLP is built");
                return WOST_ERR_UNSUPPORTED;
            }
            n->dim = it->dim;
            n->n_levels = (int)enc.number_or("n_levels", 16);
            n->base_res = (int)enc.number_or("base_resolution", 16);
            n->per_level_scale = (float)enc.number_or("per_level_scale", 2.0);
            if (n->n_levels < 1 || n->n_levels > kMaxLevels)
            {
                set_last_error("wost_reset_network: n_levels must be in [1, 16]");
                return WOST_ERR_UNSUPPORTED;
            }
            const int feat = n->n_levels * kGridFeat;
            n->k0 = feat <= 16 ? 16 : (feat <= 32 ? 32 : 64);
            if (const wjson::Value *o = cfg.find("optimizer"))
            {
                n->opt.ema_decay = (float)o->number_or("decay", n->opt.ema_decay);
                const wjson::Value *ad = o->find("nested");
                if (!ad)
                    ad = o;
                n->opt.learning_rate = (float)ad->number_or("learning_rate", n->opt.learning_rate);
                n->opt.beta1 = (float)ad->number_or("beta1", n->opt.beta1);
                n->opt.beta2 = (float)ad->number_or("beta2", n->opt.beta2);
                n->opt.epsilon = (float)ad->number_or("epsilon", n->opt.epsilon);
                n->opt.l2_reg = (float)ad->number_or("l2_reg", n->opt.l2_reg);
                n->opt.relative_decay = (float)ad->number_or("relative_decay", 0.0);
                n->opt.absolute_decay = (float)ad->number_or("absolute_decay", 0.0);
            }
            n->seed = (uint32_t)cfg.number_or("seed", 1337.0);
        }
        catch (const wjson::Error &e)
        {
            set_last_error(std::string("wost_reset_network: ") + e.what());
            return WOST_ERR_CONFIG;
        }
        net_fill_view(*n, it->problem);
        it->net = std::move(n);
        return net_alloc_and_init(it);
    }
    wost_status wost_reset_training(wost_integrator *it)
    {
        if (!it || !it->net)
            return WOST_ERR_INVALID;
        return net_alloc_and_init(it);
    }
    wost_status wost_guided_network_info(wost_integrator *it, int *n_params, int *n_mlp_params, int *encoded_width)
    {
        if (!it || !it->net)
            return WOST_ERR_INVALID;
        if (n_params)
            *n_params = it->net->n_params;
        if (n_mlp_params)
            *n_mlp_params = it->net->n_mlp;
        if (encoded_width)
            *encoded_width = it->net->k0;
        return WOST_OK;
    }
    wost_status wost_guided_get_params(wost_integrator *it, int which, float *params)
    {
        if (!it || !it->net || !params)
            return WOST_ERR_INVALID;
        GuideNet &n = *it->net;
        WOST_CUDA_TRY(cudaMemcpyAsync(params, which == 1 ? n.ema.ptr : n.master.ptr, sizeof(float) * (size_t)n.n_params, cudaMemcpyDeviceToHost, it->stream));
        WOST_CUDA_TRY(cudaStreamSynchronize(it->stream));
        return WOST_OK;
    }
    wost_status wost_guided_set_params(wost_integrator *it, const float *params)
    {
        if (!it || !it->net || !params)
            return WOST_ERR_INVALID;
        GuideNet &n = *it->net;
        const size_t P = (size_t)n.n_params;
        cudaStream_t st = it->stream;
        WOST_CUDA_TRY(cudaMemcpyAsync(n.master.ptr, params, sizeof(float) * P, cudaMemcpyHostToDevice, st));
        // optimiser state restarts; the inference ("EMA") weights start as the parameters themselves (trainer.h:83-97)
        WOST_CUDA_TRY(cudaMemcpyAsync(n.ema.ptr, n.master.ptr, sizeof(float) * P, cudaMemcpyDeviceToDevice, st));
        WOST_CUDA_TRY(cudaMemsetAsync(n.adam_m.ptr, 0, sizeof(float) * P, st));
        WOST_CUDA_TRY(cudaMemsetAsync(n.adam_v.ptr, 0, sizeof(float) * P, st));
        WOST_CUDA_TRY(cudaMemsetAsync(n.steps.ptr, 0, sizeof(uint32_t) * P, st));
        WOST_CUDA_TRY(cudaMemsetAsync(n.grad.ptr, 0, sizeof(float) * (P + 4), st));
        WOST_CUDA_TRY(cudaMemsetAsync(n.ctl.ptr, 0, sizeof(TrainCtl), st));
        k_g_sync_params<<<div_up(n.n_params, 256), 256, 0, st>>>(n.n_params, n.n_mlp, n.k0, n.master.ptr, n.master.ptr, n.w16_train.ptr, n.w16_infer.ptr);
        WOST_CUDA_TRY(cudaGetLastError());
        WOST_CUDA_TRY(cudaStreamSynchronize(st));
        return WOST_OK;
    }
    wost_status wost_guided_infer(wost_integrator *it, const float *pts, int n, int impl, int which_weights, float *out)
    {
        if (!it || !it->net || !pts || !out || n < 0)
            return WOST_ERR_INVALID;
        if (n == 0)
            return WOST_OK;
        GuideNet &net = *it->net;
        const int D = it->dim;
        DevBuf<float> d_pts, d_out;
        cudaStream_t st = it->stream;
        WOST_CUDA_TRY(d_pts.upload(pts, (size_t)n * D, st));
        WOST_CUDA_TRY(d_out.alloc((size_t)n * kOutPad));
        const GuideNetView &v = which_weights == 1 ? net.view_infer : net.view_train;
        if (impl == 1)
        {
            if (D == 2)
                k_g_mlp_simt<2><<<div_up(n, 128), 128, 0, st>>>(v, nullptr, d_pts.ptr, nullptr, n, d_out.ptr);
            else
                k_g_mlp_simt<3><<<div_up(n, 128), 128, 0, st>>>(v, nullptr, d_pts.ptr, nullptr, n, d_out.ptr);
        }
        else
        {
            const int grid = std::min(div_up(n, kTileRows), sm_count() * 2);
#define WOST_INFER_TC(D_)                                                                                                      \
    WOST_DISPATCH_K0(D_, net.k0, {                                                                                            \
        cudaFuncSetAttribute(k_g_infer_tc<D_, K0_>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)MlpFwdSmem<K0_>::kBytes); \
        k_g_infer_tc<D_, K0_><<<grid, kGBlock, MlpFwdSmem<K0_>::kBytes, st>>>(v, d_pts.ptr, n, d_out.ptr);                     \
    })
            if (D == 2)
            {
                WOST_INFER_TC(2);
            }
            else
            {
                WOST_INFER_TC(3);
            }
        }
        WOST_CUDA_TRY(cudaGetLastError());
        WOST_CUDA_TRY(d_out.download(out, (size_t)n * kOutPad, st));
        WOST_CUDA_TRY(cudaStreamSynchronize(st));
        return WOST_OK;
    }
    wost_status wost_query_network(wost_integrator *it, const float *p, float *out)
    {
        if (!it || !it->net || !p || !out)
            return WOST_ERR_INVALID;
        float raw[kOutPad];
        const wost_status rc = wost_guided_infer(it, p, 1, 1, 1, raw);
        if (rc != WOST_OK)
            return rc;
        memcpy(out, raw, sizeof(float) * (size_t)(kLobes * (it->dim + 2) + 1));
        return WOST_OK;
    }
    wost_status wost_guided_train_batch(wost_integrator *it, const float *pos, const float *dir, const float *nrm, const float *dir_pdf, const float *li,
                                        const uint8_t *on_neumann, int n, float *loss, float *grad_out)
    {
        if (!it || !it->net || !pos || !dir || !nrm || !dir_pdf || !li || !on_neumann || n < kTileRows)
        {
            set_last_error("wost_guided_train_batch: invalid argument (need >= 128 samples)");
            return WOST_ERR_INVALID;
        }
        GuideNet &net = *it->net;
        const int D = it->dim;
        cudaStream_t st = it->stream;
        const int cap = n;
        std::vector<float> h((size_t)(3 * D + 2) * cap);
        for (int i = 0; i < n; ++i)
        {
            for (int k = 0; k < D; ++k)
            {
                h[(size_t)(0 + k) * cap + i] = pos[(size_t)i * D + k];
                h[(size_t)(D + k) * cap + i] = dir[(size_t)i * D + k];
                h[(size_t)(2 * D + k) * cap + i] = nrm[(size_t)i * D + k];
            }
            h[(size_t)(3 * D) * cap + i] = dir_pdf[i];
            h[(size_t)(3 * D + 1) * cap + i] = li[i];
        }
        DevBuf<float> d_smp;
        DevBuf<uint8_t> d_on;
        WOST_CUDA_TRY(d_smp.upload(h.data(), h.size(), st));
        WOST_CUDA_TRY(d_on.upload(on_neumann, (size_t)n, st));
        int *opt_steps = &net.ctl.ptr->opt_steps[0];
        k_g_frame_ctl<<<1, 1, 0, st>>>(net.ctl.ptr, opt_steps, 0);
        WOST_CUDA_TRY(cudaMemsetAsync(net.grad.ptr, 0, sizeof(float) * ((size_t)net.n_params + 4), st));
        TrainArgs ta;
        ta.smp = d_smp.ptr, ta.smp_onN = d_on.ptr, ta.cap = cap;
        ta.ctl = net.ctl.ptr;
        ta.iter = 0, ta.batch_size = n, ta.min_batch = kTileRows, ta.per_frame = 1, ta.n_override = n;
        ta.loss_scale = 128.0f;
        ta.grad = net.grad.ptr, ta.n_params = net.n_params;
        wost_status rc = WOST_OK;
        if (D == 2)
        {
            WOST_DISPATCH_K0(2, net.k0, (rc = launch_train<2, K0_>(it, ta)));
        }
        else
        {
            WOST_DISPATCH_K0(3, net.k0, (rc = launch_train<3, K0_>(it, ta)));
        }
        if (rc != WOST_OK)
            return rc;
        WOST_CUDA_TRY(cudaGetLastError());
        if (grad_out)
            WOST_CUDA_TRY(cudaMemcpyAsync(grad_out, net.grad.ptr, sizeof(float) * (size_t)net.n_params, cudaMemcpyDeviceToHost, st));
        if (it->allreduce && it->allreduce(it->allreduce_user, net.grad.ptr, (uint64_t)net.n_params + 4, (void *)st) != 0)
            return WOST_ERR_INVALID;
        rc = launch_adam(it, ta.loss_scale, opt_steps);
        if (rc != WOST_OK)
            return rc;
        TrainCtl h_ctl;
        WOST_CUDA_TRY(cudaMemcpyAsync(&h_ctl, net.ctl.ptr, sizeof(h_ctl), cudaMemcpyDeviceToHost, st));
        WOST_CUDA_TRY(cudaStreamSynchronize(st));
        if (loss)
            *loss = h_ctl.loss[0] / (float)std::max(1, h_ctl.batch_rows[0]);
        return WOST_OK;
    }
    wost_status wost_vmm_eval(int dim, const float *raw, const float *dir, const float *nrm, const uint8_t *on_neumann, int n, uint32_t seed, float *pdf,
                              float *grad, float *sampled_dir)
    {
        if ((dim != 2 && dim != 3) || !raw || !dir || !nrm || !on_neumann || !pdf || !grad || !sampled_dir || n <= 0)
            return WOST_ERR_INVALID;
        const int kOut = kLobes * (dim + 2) + 1, kG = kLobes * (dim + 2);
        DevBuf<float> d_raw, d_dir, d_nrm, d_pdf, d_grad, d_s;
        DevBuf<uint8_t> d_on;
        WOST_CUDA_TRY(d_raw.upload(raw, (size_t)n * kOut));
        WOST_CUDA_TRY(d_dir.upload(dir, (size_t)n * dim));
        WOST_CUDA_TRY(d_nrm.upload(nrm, (size_t)n * dim));
        WOST_CUDA_TRY(d_on.upload(on_neumann, (size_t)n));
        WOST_CUDA_TRY(d_pdf.alloc(n));
        WOST_CUDA_TRY(d_grad.alloc((size_t)n * kG));
        WOST_CUDA_TRY(d_s.alloc((size_t)n * dim));
        if (dim == 2)
            k_vmm_probe<2><<<div_up(n, 128), 128>>>(d_raw.ptr, d_dir.ptr, d_nrm.ptr, d_on.ptr, n, seed, d_pdf.ptr, d_grad.ptr, d_s.ptr);
        else
            k_vmm_probe<3><<<div_up(n, 128), 128>>>(d_raw.ptr, d_dir.ptr, d_nrm.ptr, d_on.ptr, n, seed, d_pdf.ptr, d_grad.ptr, d_s.ptr);
        WOST_CUDA_TRY(cudaGetLastError());
        WOST_CUDA_TRY(d_pdf.download(pdf, n));
        WOST_CUDA_TRY(d_grad.download(grad, (size_t)n * kG));
        WOST_CUDA_TRY(d_s.download(sampled_dir, (size_t)n * dim));
        WOST_CUDA_TRY(cudaDeviceSynchronize());
        return WOST_OK;
    }
    wost_status wost_set_allreduce(wost_integrator *it, wost_allreduce_fn fn, void *user)
    {
        if (!it)
            return WOST_ERR_INVALID;
        it->allreduce = fn;
        it->allreduce_user = user;
        return WOST_OK;
    }
    wost_status wost_guided_last_profile(wost_integrator *it, float *ms8)
    {
        if (!it || !ms8)
            return WOST_ERR_INVALID;
        for (int k = 0; k < 8; ++k)
            ms8[k] = it->prof_ms[k];
        return WOST_OK;
    }
    // parity probe: the compacted training set of the LAST training frame (generate_training_data, guided/train.h:422-471), one row
    // per sample: pos[dim], dir[dim], nN[dim], dirPdf, Li, onN (3 * dim + 3 floats)
    wost_status wost_guided_last_samples(wost_integrator *it, float *rows, uint64_t capacity_rows, uint64_t *n_rows)
    {
        if (!it || it->kind != WOST_GUIDED || !n_rows)
            return WOST_ERR_INVALID;
        const GuidedBuffers &g = it->gb;
        const uint64_t n = std::min<uint64_t>(it->train_samples, (uint64_t)std::max(0, g.cap));
        *n_rows = n;
        if (!rows || n == 0)
            return WOST_OK;
        if (capacity_rows < n)
        {
            set_last_error("wost_guided_last_samples: buffer too small");
            return WOST_ERR_INVALID;
        }
        const int D = it->dim, planes = 3 * D + 2, cap = g.cap;
        std::vector<float> h((size_t)planes * cap);
        std::vector<uint8_t> on((size_t)cap);
        WOST_CUDA_TRY(cudaMemcpyAsync(h.data(), g.smp.ptr, sizeof(float) * h.size(), cudaMemcpyDeviceToHost, it->stream));
        WOST_CUDA_TRY(cudaMemcpyAsync(on.data(), g.smp_onN.ptr, on.size(), cudaMemcpyDeviceToHost, it->stream));
        WOST_CUDA_TRY(cudaStreamSynchronize(it->stream));
        const int w = planes + 1;
        for (uint64_t i = 0; i < n; ++i)
        {
            for (int k = 0; k < planes; ++k)
                rows[i * w + k] = h[(size_t)k * cap + i];
            rows[i * w + planes] = on[i] ? 1.0f : 0.0f;
        }
        return WOST_OK;
    }
    wost_status wost_guided_last_stats(wost_integrator *it, uint64_t *train_samples, uint64_t *train_steps, float *last_loss)
    {
        if (!it)
            return WOST_ERR_INVALID;
        if (train_samples)
            *train_samples = it->train_samples;
        if (train_steps)
            *train_steps = it->train_steps;
        if (last_loss)
            *last_loss = it->last_loss;
        return WOST_OK;
    }
}
