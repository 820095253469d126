// Problem + integrator objects and the wavefront walk kernels (uniform integrator).
//
// Wavefront scheme (replaces solveImpl, integrator/uniform/integrator.cu:528-623):
//   reference:  for spp: for depth: 1 single-thread reset kernel + 3..5 full-frame stage kernels, W*H threads each,
//               one global atomic per pushed item, queues in managed memory.
//   here:       for each batch of S samples-per-pixel: for depth: ONE fused kernel (walk.cuh) over the compacted queue
//               of live walks. Persistent grid (multiple of the SM count); each warp claims 32 items at a time with one
//               atomic (dynamic load balance), keeps the walk state in registers for the whole step, and appends
//               survivors to the next queue with one warp-aggregated atomic. Depth 0 synthesises its items from
//               (pixel, sample) -- no generate kernel, no queue read. Per-(slot, pixel) accumulators make the result
//               independent of scheduling: a pixel's samples of slot s are summed in sample order (== the oracle's
//               order when S == 1), slots are summed in slot order by the resolve kernel.
#include "integrator.cuh"
#include <mutex>
#include "walk.cuh"

#include <algorithm>
#include <chrono>
#include <cstdlib>
#include <cstring>

namespace wost
{
    constexpr int kWalkBlock = 128;

    WOST_DEV void load_item(const QueueView &q, size_t i, const uint32_t *active, WalkState<2> &w)
    {
        const float4 a = q.a[i], b = q.b[i];
        w.x[0] = a.x, w.x[1] = a.y;
        w.ai = __float_as_uint(a.z);
        w.pixel = __ldg(active + w.ai);
        const uint32_t sf = __float_as_uint(a.w);
        w.sample = sf & 0x7fffffffu;
        w.onN = (sf >> 31) != 0;
        w.nN[0] = b.x, w.nN[1] = b.y;
        w.thp = b.z;
    }
    WOST_DEV void store_item(const QueueView &q, size_t i, const WalkState<2> &w)
    {
        q.a[i] = make_float4(w.x[0], w.x[1], __uint_as_float(w.ai), __uint_as_float(w.sample | (w.onN ? 0x80000000u : 0u)));
        q.b[i] = make_float4(w.nN[0], w.nN[1], w.thp, 0.0f);
    }
    WOST_DEV void load_item(const QueueView &q, size_t i, const uint32_t *active, WalkState<3> &w)
    {
        const float4 a = q.a[i], b = q.b[i];
        const uint32_t sf = q.c[i];
        w.x[0] = a.x, w.x[1] = a.y, w.x[2] = a.z;
        w.thp = a.w;
        w.nN[0] = b.x, w.nN[1] = b.y, w.nN[2] = b.z;
        w.ai = __float_as_uint(b.w);
        w.pixel = __ldg(active + w.ai);
        w.sample = sf & 0x7fffffffu;
        w.onN = (sf >> 31) != 0;
    }
    WOST_DEV void store_item(const QueueView &q, size_t i, const WalkState<3> &w)
    {
        q.a[i] = make_float4(w.x[0], w.x[1], w.x[2], w.thp);
        q.b[i] = make_float4(w.nN[0], w.nN[1], w.nN[2], __uint_as_float(w.ai));
        q.c[i] = w.sample | (w.onN ? 0x80000000u : 0u);
    }

    struct StepArgs
    {
        QueueView in, out;
        const int *in_count; // nullptr for depth 0
        int *out_count;
        int *cursor;              // claim cursor of this bounce
        unsigned long long *steps; // walk-step counter
        const uint32_t *active;
        int n_active;
        uint32_t sample0; // first sample of the batch
        int batch_samples;
        uint32_t depth;
        float4 *acc; // [slot][active rank]
        int last_depth; // 1: survivors are dropped (depth == maxWalkingDepth-1)
        int *overflow;
    };

    template <int D, bool First, bool RefOrder>
    __global__ void __launch_bounds__(kWalkBlock) k_step_uniform(const __grid_constant__ WalkCtx<D> c, const __grid_constant__ StepArgs a)
    {
        extern __shared__ uint2 smem_stack[];
        SmemStack st;
        uint2 spill_mem[kSpillLevels];
        st.init(smem_stack, a.overflow, spill_mem);
        const int lane = threadIdx.x & 31;
        const long long n = First ? (long long)a.n_active * a.batch_samples : (long long)*a.in_count;
        if (blockIdx.x == 0 && threadIdx.x == 0 && n > 0)
            atomicAdd(a.steps, (unsigned long long)n);
        for (;;)
        {
            long long base = 0;
            if (lane == 0)
                base = (long long)atomicAdd(a.cursor, 32);
            base = __shfl_sync(0xffffffffu, base, 0);
            if (base >= n)
                break;
            const long long i = base + lane;
            const bool valid = i < n;
            bool alive = false;
            WalkState<D> w;
            if (valid)
            {
                if (First)
                {
                    const uint32_t sj = (uint32_t)(i / a.n_active);
                    w.ai = (uint32_t)(i - (long long)sj * a.n_active);
                    w.pixel = __ldg(a.active + w.ai);
                    w.sample = a.sample0 + sj;
                    w.x = eval_point(c, w.pixel);
                    w.thp = 1.0f;
                    w.onN = false;
#pragma unroll
                    for (int k = 0; k < D; ++k)
                        w.nN[k] = 0.0f;
                }
                else
                    load_item(a.in, (size_t)i, a.active, w);
            }
            else
            {
                w.pixel = 0, w.ai = 0, w.sample = a.sample0, w.thp = 0.0f, w.onN = false;
#pragma unroll
                for (int k = 0; k < D; ++k)
                    w.x[k] = w.nN[k] = 0.0f;
            }
            {
                // every lane of the warp enters the step (the closest-point traversal is warp-synchronous)
                PixelAcc acc;
                acc.init(reinterpret_cast<float *>(a.acc + (size_t)(w.sample - a.sample0) * a.n_active + w.ai));
                alive = walk_step_uniform<D, RefOrder>(c, w, a.depth, st, acc, valid);
                acc.flush();
            }
            if (!a.last_depth)
            {
                // warp-aggregated append: one atomic per warp
                const uint32_t mask = __ballot_sync(0xffffffffu, alive);
                if (mask)
                {
                    int slot0 = 0;
                    if (lane == 0)
                        slot0 = atomicAdd(a.out_count, __popc(mask));
                    slot0 = __shfl_sync(0xffffffffu, slot0, 0);
                    if (alive)
                        store_item(a.out, (size_t)slot0 + __popc(mask & ((1u << lane) - 1u)), w);
                }
            }
        }
    }

    // ---------------------------------------------------------------------------------------------------------
    // Persistent walk kernel (default mode): ONE launch per batch of S samples-per-pixel. Each lane owns a whole walk
    // -- state in registers from the evaluation point to the epsilon-shell -- and the warp runs a two-phase loop:
    //   traversal trips   every lane with a closest-point query in flight opens one node / evaluates one leaf;
    //   tail + refill     as soon as refill_threshold lanes (walk.cuh: default_refill_threshold) have finished their query (or nobody is traversing any
    //                     more) those lanes run the rest of their walk step (walk_step_tail: epsilon-shell test,
    //                     Dirichlet read-out, silhouette query, source / Neumann samples, direction sample + ray),
    //                     lanes whose walk ended claim new (pixel, sample) jobs with one atomic per warp, and all of
    //                     them start their next query.
    // So the closest-point traversal -- >90% of the instructions -- runs with nearly all lanes busy although the
    // number of visits per query varies 10x between neighbouring lanes, and a bounce costs neither a kernel launch nor a
    // queue round trip through HBM. Same Philox stream (pixel, sample, depth) and same per-(slot, pixel) accumulators
    // as the wavefront kernel, hence the same film bit for bit.

    struct PersistArgs
    {
        const uint32_t *active;
        const uint2 *first; // first-step cache (k_first_hit), indexed like `active`
        int n_active;
        uint32_t sample0;
        int batch_samples;
        float4 *acc; // [slot][active rank]
        int max_depth;
        int refill_threshold;
        int stack_levels; // shared-memory levels of the traversal stack (deeper levels spill to local memory)
        int *cursor;
        unsigned long long *steps;
        int *overflow;
    };

    // resident CTAs per SM the lean 3D variants are compiled for (64 registers: a few spilled bytes buy an eighth CTA; c4 +8%);
    // 0 = no request, the compiler's own budget (an explicit request cost the 2D kernels 2-7%)
    constexpr int persist_min_ctas(int D, bool ref_order, int F)
    {
        return ((D == 3 && !ref_order && !(F & kFeatNeumann)) || (D == 2 && !ref_order && F == kFeatNeumann)) ? 8 : 0;
    }
    template <int D, bool RefOrder, int F>
    __global__ void __launch_bounds__(kWalkBlock, persist_min_ctas(D, RefOrder, F)) k_walk_persistent(const __grid_constant__ WalkCtx<D> c, const __grid_constant__ PersistArgs a)
    {
        extern __shared__ uint2 smem_stack[];
        SmemStack st;
        uint2 spill_mem[kSpillLevels];
        st.init(smem_stack, a.overflow, spill_mem, a.stack_levels);
        const int lane = threadIdx.x & 31;
        const long long njobs = (long long)a.n_active * a.batch_samples;
        WalkState<D> w;
        NearestTrav<D> t;
        uint32_t depth = 0;
        bool has_walk = false, exhausted = false, trav = false, fresh = false;
        unsigned int my_steps = 0;
        float bound = dm_inf(); // upper bound of the next closest-point distance: R_D of this step + the distance moved
        w.pixel = 0, w.ai = 0, w.sample = a.sample0, w.thp = 0.0f, w.onN = false;
#pragma unroll
        for (int k = 0; k < D; ++k)
            w.x[k] = w.nN[k] = 0.0f;
        t.win = kInvalid, t.best = 0.0f;
#ifdef WOST_TRAV_DEBUG
        t.dbg_dec = 0;
        long long dbg_t0 = clock64();
        unsigned long long *dbg_cyc = reinterpret_cast<unsigned long long *>(a.overflow + 32); // [0..7] trip-iteration cycles, [8..15] tail-iteration cycles per decade
#endif
        for (;;)
        {
            const unsigned trav_mask = __ballot_sync(kFullMask, trav);
            const unsigned pend_mask = __ballot_sync(kFullMask, !trav && (has_walk || !exhausted));
#ifdef WOST_TRAV_DEBUG
            {
                // attribute the previous iteration of this warp to the farthest query decade among its lanes
                int dec = (trav || has_walk) ? (t.dbg_dec & 7) : 0;
#pragma unroll
                for (int o = 16; o > 0; o >>= 1)
                    dec = max(dec, __shfl_xor_sync(kFullMask, dec, o));
                const long long now = clock64();
                const bool tail_iter = pend_mask != 0u && (__popc(pend_mask) >= a.refill_threshold || trav_mask == 0u);
                if (lane == 0)
                    atomicAdd(dbg_cyc + (tail_iter ? 8 : 0) + dec, (unsigned long long)(now - dbg_t0));
                dbg_t0 = now;
            }
#endif
            if (pend_mask != 0u && (__popc(pend_mask) >= a.refill_threshold || trav_mask == 0u))
            {
                if (!trav && has_walk)
                {
                    PixelAcc acc;
                    acc.init(reinterpret_cast<float *>(a.acc + (size_t)(w.sample - a.sample0) * a.n_active + w.ai));
                    w.pixel = __ldg(a.active + w.ai); // only the tail needs the pixel id (Philox key): not live across the traversal trips
                    const Vec<D> x0 = w.x;
                    const float R_D = nearest_distance(t);
                    const bool alive = walk_step_tail<D, F>(c, w, depth, st, acc, R_D, t.win);
                    acc.flush();
                    ++depth;
                    if (!alive || depth >= (uint32_t)a.max_depth)
                        has_walk = false;
                    // upper bound of the next closest-point distance: the distance from the new point to the primitive that
                    // was closest a step ago (<= R_D + |step| by the triangle inequality, usually much tighter), rounding slack on top
                    if (alive && t.win != kInvalid)
                        bound = prim_distance(load_prim(c.dir.leaf_prim + t.win), w.x) * 1.00001f;
                    else
                        bound = (R_D + vlen(vsub(w.x, x0))) * 1.00001f;
                }
                const bool want = !has_walk && !exhausted;
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
                        const long long i = (long long)base + __popc(want_mask & ((1u << lane) - 1u));
                        if (i < njobs)
                        {
                            const uint32_t sj = (uint32_t)(i / a.n_active);
                            const long long ai = i - (long long)sj * a.n_active;
                            w.ai = (uint32_t)ai;
                            w.sample = a.sample0 + sj;
                            w.x = eval_point(c, __ldg(a.active + ai));
                            w.thp = 1.0f;
                            w.onN = false;
#pragma unroll
                            for (int k = 0; k < D; ++k)
                                w.nN[k] = 0.0f;
                            depth = 0;
                            has_walk = true;
                            bound = dm_inf();
                            // every sample of a pixel starts at the same point: its closest-point query was answered once
                            // per solve by k_first_hit (same traversal, same bits)
                            const uint2 f = __ldg(a.first + ai);
                            t.win = f.x;
                            t.best = __uint_as_float(f.y);
                            fresh = true;
                        }
                        else
                            exhausted = true;
                    }
                }
                if (!trav && has_walk)
                {
                    ++my_steps;
                    if (fresh)
                        fresh = false; // result already in t: the lane stays pending and takes the next tail phase
                    else
                        trav = nearest_begin<D, RefOrder>(c.dir, w.x, st, t, bound, c.eps, (F & kFeatNeumann) != 0);
                }
                // walks that are ~1e8 units away from the scene: the warp scans the leaves together (lbvh.cuh, phase 4)
                // (only in the variants with a Neumann mesh: walks leave a scene through the corners of closed Neumann polygons)
                if (!RefOrder && D == 2 && (F & kFeatNeumann) && (nearest_brute_warp<D>(c.dir, w.x, trav && t.phase == 4, t) >> lane & 1u))
                    trav = false;
            }
            else if (trav_mask == 0u)
                break;
            if (trav)
                trav = nearest_trip<D, RefOrder>(c.dir, w.x, st, t);
        }
        unsigned int ws = my_steps;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1)
            ws += __shfl_xor_sync(kFullMask, ws, o);
        if (lane == 0 && ws)
            atomicAdd(a.steps, (unsigned long long)ws);
    }

    // solution/spp -> Film (integrator.cu:614-620): {r,g,b,1, weight 1}. Accumulators are indexed by the active rank of
    // the pixel (shard-sized); slots are summed in slot order and scattered to the pixel's film entry. The film entries of
    // pixels this shard does not own keep their initial {0,0,0,1,1} (k_film_init).
    // se != nullptr: also the standard error of the pixel mean from the slot means (slot s holds slot_n[s] samples of the
    // pixel): se^2 = sum_s n_s (m_s - m)^2 / ((S_used - 1) * spp) -- the exact sample variance / spp when every slot holds one sample.
    __global__ void k_resolve(int n_active, int nslots, const float4 *acc, const uint32_t *active, float spp, float *film, const int *slot_n, float *se)
    {
        const int ai = blockIdx.x * blockDim.x + threadIdx.x;
        if (ai >= n_active)
            return;
        float r = 0.0f, g = 0.0f, b = 0.0f;
        for (int s = 0; s < nslots; ++s)
        {
            const float4 v = acc[(size_t)s * n_active + ai];
            r = r + v.x;
            g = g + v.y;
            b = b + v.z;
        }
        const uint32_t p = active[ai];
        float *f = film + (size_t)p * 5;
        f[0] = dm_div(r, spp);
        f[1] = dm_div(g, spp);
        f[2] = dm_div(b, spp);
        f[3] = 1.0f;
        f[4] = 1.0f;
        if (se)
        {
            const double m[3] = {(double)r / spp, (double)g / spp, (double)b / spp};
            double q[3] = {0.0, 0.0, 0.0};
            int used = 0;
            for (int s = 0; s < nslots; ++s)
            {
                const int n = slot_n[s];
                if (n <= 0)
                    continue;
                ++used;
                const float4 v = acc[(size_t)s * n_active + ai];
                const double d0 = (double)v.x / n - m[0], d1 = (double)v.y / n - m[1], d2 = (double)v.z / n - m[2];
                q[0] += n * d0 * d0, q[1] += n * d1 * d1, q[2] += n * d2 * d2;
            }
            for (int ch = 0; ch < 3; ++ch)
                se[(size_t)p * 3 + ch] = used > 1 ? (float)sqrt(q[ch] / ((double)(used - 1) * spp)) : 0.0f;
        }
    }
    __global__ void k_film_init(int npix, float *film)
    {
        const int p = blockIdx.x * blockDim.x + threadIdx.x;
        if (p >= npix)
            return;
        float *f = film + (size_t)p * 5;
        f[0] = f[1] = f[2] = 0.0f;
        f[3] = f[4] = 1.0f;
    }

    void launch_resolve(wost_integrator *it, int nslots, float spp)
    {
        if (it->n_active > 0)
            k_resolve<<<div_up(it->n_active, 256), 256, 0, it->stream>>>(it->n_active, nslots, it->d_acc.ptr, it->d_active.ptr, spp, it->film[WOST_CH_SOLUTION].ptr, nullptr, nullptr);
    }

    wost_status emit_frames(wost_integrator *it, int sample_id)
    {
        const int kinds = it->frame_kinds(sample_id);
        if (!kinds)
            return WOST_OK;
        cudaStream_t st = it->stream;
        launch_resolve(it, it->spp_batch, (float)(sample_id + 1));
        ++it->launches;
        it->frame_host.resize((size_t)it->npix * 5);
        WOST_CUDA_TRY(cudaMemcpyAsync(it->frame_host.data(), it->film[WOST_CH_SOLUTION].ptr, sizeof(float) * 5 * (size_t)it->npix, cudaMemcpyDeviceToHost, st));
        WOST_CUDA_TRY(cudaStreamSynchronize(st));
        const uint64_t ms = (uint64_t)std::chrono::duration_cast<std::chrono::milliseconds>(std::chrono::high_resolution_clock::now() - it->solve_t0).count();
        for (int kind = 0; kind < 2; ++kind)
            if (kinds & (1 << kind))
                it->frame_cb(it->frame_user, kind, sample_id, ms, it->frame_host.data(), it->s.frame_w, it->s.frame_h);
        return WOST_OK;
    }

    // renderDirichletSDF / renderSilhouetteSDF / renderSource (integrator/common.h:51-163)
    template <int D>
    __global__ void __launch_bounds__(kWalkBlock) k_render_channel(const __grid_constant__ WalkCtx<D> c, int channel, int npix, float *film, int *overflow)
    {
        extern __shared__ uint2 smem_stack[];
        SmemStack st;
        uint2 spill_mem[kSpillLevels];
        st.init(smem_stack, overflow, spill_mem);
        for (int base = blockIdx.x * blockDim.x; base < npix; base += gridDim.x * blockDim.x)
        {
            const int p = base + threadIdx.x;
            const bool valid = p < npix;
            const Vec<D> x = eval_point(c, (uint32_t)(valid ? p : 0));
            float v[3] = {0.0f, 0.0f, 0.0f};
            if (channel == WOST_CH_DIRICHLET_SDF) // uniform branch: query_nearest is warp-synchronous
            {
                float d = dm_inf();
                if (c.dir.nprims > 0)
                {
                    uint32_t pos;
                    query_nearest(c.dir, x, st, pos, d, valid);
                }
                v[0] = v[1] = v[2] = d;
            }
            if (!valid)
                continue;
            if (channel == WOST_CH_NEUMANN_SDF)
            {
                float d = dm_inf();
                if (c.neu.nprims > 0)
                    d = query_silhouette(c.neu, x, false, st);
                v[0] = v[1] = v[2] = d;
            }
            else if (channel == WOST_CH_SOURCE)
            {
                if (c.source)
                {
                    float w3[3];
                    w3[0] = x[0], w3[1] = x[1], w3[2] = (D == 3) ? x[D - 1] : 0.0f;
                    source_lookup(c, w3, v);
                    v[0] *= c.source_intensity, v[1] *= c.source_intensity, v[2] *= c.source_intensity;
                }
            }
            float *f = film + (size_t)p * 5;
            f[0] = v[0], f[1] = v[1], f[2] = v[2], f[3] = 1.0f, f[4] = 1.0f;
        }
    }

    template <int D, bool RefOrder>
    static wost_status solve_uniform(wost_integrator *it, uint64_t *walk_steps)
    {
        WalkCtx<D> ctx;
        it->problem->fill_ctx(ctx);
        ctx.frame_w = it->s.frame_w;
        ctx.frame_h = it->s.frame_h;
        ctx.max_depth = it->s.max_walking_depth;
        ctx.eps = it->s.epsilon_shell;
        ctx.seed = it->s.seed;
        const int S = it->spp_batch;
        const int spp = it->s.samples_per_pixel;
        const int maxd = it->s.max_walking_depth;
        cudaStream_t st = it->stream;
        const size_t smem = stack_smem_bytes(kWalkBlock);
        int occ = 1;
        WOST_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_step_uniform<D, false, RefOrder>, kWalkBlock, smem));
        const int grid_full = sm_count() * std::max(1, occ);

        WOST_CUDA_TRY(cudaMemsetAsync(it->d_acc.ptr, 0, sizeof(float4) * (size_t)S * std::max(1, it->n_active), st));
        WOST_CUDA_TRY(cudaMemsetAsync(it->d_steps.ptr, 0, sizeof(unsigned long long), st));
        it->launches = 0;
        it->step_launches = 0;
        it->step_kernel_ms = 0.0f;
        it->slot_n.assign((size_t)S, 0);
        for (int s0 = 0; s0 < spp; s0 += S)
            for (int j = 0; j < std::min(S, spp - s0); ++j)
                ++it->slot_n[(size_t)j];
        const int nbatches = it->n_active > 0 ? (spp + S - 1) / S : 0;
        while ((int)it->batch_ev.size() < 2 * nbatches)
        {
            cudaEvent_t e;
            WOST_CUDA_TRY(cudaEventCreate(&e));
            it->batch_ev.push_back(e);
        }
        if (it->n_active > 0)
        {
            for (int s0 = 0, bi = 0; s0 < spp; s0 += S, ++bi)
            {
                const int Sb = std::min(S, spp - s0);
                WOST_CUDA_TRY(cudaMemsetAsync(it->d_counts.ptr, 0, sizeof(int) * (size_t)(2 * (maxd + 1)), st));
                WOST_CUDA_TRY(cudaEventRecord(it->batch_ev[2 * bi], st));
                for (int depth = 0; depth < maxd; ++depth)
                {
                    StepArgs a;
                    a.in = QueueView{it->qa[depth & 1].ptr, it->qb[depth & 1].ptr, it->qc[depth & 1].ptr};
                    a.out = QueueView{it->qa[(depth + 1) & 1].ptr, it->qb[(depth + 1) & 1].ptr, it->qc[(depth + 1) & 1].ptr};
                    a.in_count = depth == 0 ? nullptr : it->d_counts.ptr + depth;
                    a.out_count = it->d_counts.ptr + depth + 1;
                    a.cursor = it->d_counts.ptr + (maxd + 1) + depth;
                    a.steps = it->d_steps.ptr;
                    a.active = it->d_active.ptr;
                    a.n_active = it->n_active;
                    a.sample0 = (uint32_t)s0;
                    a.batch_samples = Sb;
                    a.depth = (uint32_t)depth;
                    a.acc = it->d_acc.ptr;
                    a.last_depth = (depth == maxd - 1) ? 1 : 0;
                    a.overflow = it->d_overflow.ptr;
                    if (depth == 0)
                    {
                        const long long n0 = (long long)it->n_active * Sb;
                        const int grid = (int)std::min<long long>(grid_full, (n0 + kWalkBlock - 1) / kWalkBlock);
                        k_step_uniform<D, true, RefOrder><<<std::max(1, grid), kWalkBlock, smem, st>>>(ctx, a);
                    }
                    else
                        k_step_uniform<D, false, RefOrder><<<grid_full, kWalkBlock, smem, st>>>(ctx, a);
                    ++it->launches;
                    ++it->step_launches;
                }
                WOST_CUDA_TRY(cudaEventRecord(it->batch_ev[2 * bi + 1], st));
            }
        }
        launch_resolve(it, S, (float)spp);
        ++it->launches;
        WOST_CUDA_TRY(cudaGetLastError());
        unsigned long long h_steps = 0;
        int h_over = 0;
        WOST_CUDA_TRY(cudaMemcpyAsync(&h_steps, it->d_steps.ptr, sizeof(h_steps), cudaMemcpyDeviceToHost, st));
        WOST_CUDA_TRY(cudaMemcpyAsync(&h_over, it->d_overflow.ptr, sizeof(int), cudaMemcpyDeviceToHost, st));
        WOST_CUDA_TRY(cudaEventRecord(it->ev1, st));
        WOST_CUDA_TRY(cudaStreamSynchronize(st));
        if (walk_steps)
            *walk_steps = h_steps;
        for (int bi = 0; bi < nbatches; ++bi)
        {
            float f = 0.0f;
            WOST_CUDA_TRY(cudaEventElapsedTime(&f, it->batch_ev[2 * bi], it->batch_ev[2 * bi + 1]));
            it->step_kernel_ms += f;
        }
        if (h_over)
        {
            cudaMemsetAsync(it->d_overflow.ptr, 0, sizeof(int), st);
            set_last_error("LBVH traversal stack overflow during solve");
            return WOST_ERR_OVERFLOW;
        }
        return WOST_OK;
    }
    // Blackwell L2 persistence for the top BVH levels: when the Dirichlet tree (fused records + leaf records) does not fit
    // the L2, the contiguous slab of its top levels (scene.cu: breadth-first record order, <= 32 MB) gets a persisting
    // access-policy window on the integrator's stream (hitRatio 1: every line of the window is marked persisting; missProp would
    // only apply to window lines left out by a hitRatio < 1, so it has no effect here). Accesses outside the window keep the
    // normal policy and compete for the rest of the L2. WOST_L2_PERSIST=0 disables it.
    // The carve-out (cudaLimitPersistingL2CacheSize) is device-wide: it is raised to the largest window any live integrator of the
    // device holds and never lowered under one; persisting lines are reset only when the LAST holder on the device goes away
    // (cudaCtxResetPersistingL2Cache would otherwise drop the lines of every other live integrator, e.g. the c4-slices groups).
    namespace
    {
        constexpr int kL2MaxDevices = 64;
        std::mutex g_l2_mu;
        size_t g_l2_limit[kL2MaxDevices] = {};
        int g_l2_holders[kL2MaxDevices] = {};
    } // namespace
    static bool acquire_l2_carveout(int device, size_t bytes)
    {
        if (device < 0 || device >= kL2MaxDevices)
            return false;
        std::lock_guard<std::mutex> lock(g_l2_mu);
        if (bytes > g_l2_limit[device])
        {
            if (cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, bytes) != cudaSuccess)
            {
                cudaGetLastError();
                return false;
            }
            g_l2_limit[device] = bytes;
        }
        ++g_l2_holders[device];
        return true;
    }
    void release_l2_window(int device)
    {
        if (device < 0 || device >= kL2MaxDevices)
            return;
        std::lock_guard<std::mutex> lock(g_l2_mu);
        if (g_l2_holders[device] > 0 && --g_l2_holders[device] == 0)
        {
            int cur = 0;
            cudaGetDevice(&cur);
            if (cur != device)
                cudaSetDevice(device);
            cudaCtxResetPersistingL2Cache();
            cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, 0);
            cudaGetLastError();
            if (cur != device)
                cudaSetDevice(cur);
            g_l2_limit[device] = 0;
        }
    }
    static void apply_l2_window(wost_integrator *it)
    {
        it->l2_window_bytes = 0;
        const wost_scene *sc = it->problem->dirichlet;
        const char *e = getenv("WOST_L2_PERSIST");
        if (!sc || sc->nprims < 2 || (e && atoi(e) == 0))
            return;
        int l2 = 0, max_persist = 0, max_window = 0;
        cudaDeviceGetAttribute(&l2, cudaDevAttrL2CacheSize, it->device);
        cudaDeviceGetAttribute(&max_persist, cudaDevAttrMaxPersistingL2CacheSize, it->device);
        cudaDeviceGetAttribute(&max_window, cudaDevAttrMaxAccessPolicyWindowSize, it->device);
        const size_t leaf_bytes = (size_t)sc->nprims * (sc->dim == 2 ? sizeof(LeafPrim<2>) : sizeof(LeafPrim<3>));
        const bool force = e && atoi(e) >= 2; // 2: also for trees that fit (A/B measurements)
        if ((!force && sc->tnode_bytes() + leaf_bytes <= (size_t)l2) || max_persist <= 0 || max_window <= 0)
            return;
        const size_t bytes = std::min(std::min(sc->top_bytes(), (size_t)max_window), (size_t)max_persist);
        if (bytes == 0 || !acquire_l2_carveout(it->device, bytes))
            return;
        cudaStreamAttrValue v;
        memset(&v, 0, sizeof(v));
        v.accessPolicyWindow.base_ptr = sc->d_tnodes.ptr;
        v.accessPolicyWindow.num_bytes = bytes;
        v.accessPolicyWindow.hitRatio = 1.0f;
        v.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
        v.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
        if (cudaStreamSetAttribute(it->stream, cudaStreamAttributeAccessPolicyWindow, &v) == cudaSuccess)
            it->l2_window_bytes = bytes;
        else
        {
            cudaGetLastError();
            release_l2_window(it->device);
        }
    }
    template <int D, bool RefOrder, int F>
    static wost_status solve_persistent_f(wost_integrator *it, uint64_t *walk_steps);
    // picks the kernel variant compiled for the features this problem has (walk.cuh: kFeatNeumann / kFeatSource); the
    // reference-order stand-in mode always runs the full-feature variant
    template <int D, bool RefOrder>
    static wost_status solve_persistent(wost_integrator *it, uint64_t *walk_steps)
    {
        if (RefOrder)
            return solve_persistent_f<D, RefOrder, kFeatAll>(it, walk_steps);
        const bool neu = it->problem->neumann != nullptr, src = it->problem->d_source.ptr != nullptr, flux = neu && it->problem->neumann_flux_enabled;
        if (neu && !flux && !src) // data/ladybug, data/fille: reflecting boundary, no source
            return solve_persistent_f<D, false, kFeatNeumann>(it, walk_steps);
        if (neu)
            return solve_persistent_f<D, false, kFeatAll>(it, walk_steps);
        if (src)
            return solve_persistent_f<D, false, kFeatSource>(it, walk_steps);
        return solve_persistent_f<D, false, 0>(it, walk_steps);
    }
    template <int D, bool RefOrder, int F>
    static wost_status solve_persistent_f(wost_integrator *it, uint64_t *walk_steps)
    {
        WalkCtx<D> ctx;
        it->problem->fill_ctx(ctx);
        ctx.frame_w = it->s.frame_w;
        ctx.frame_h = it->s.frame_h;
        ctx.max_depth = it->s.max_walking_depth;
        ctx.eps = it->s.epsilon_shell;
        ctx.seed = it->s.seed;
        const int S = it->spp_batch;
        const int spp = it->s.samples_per_pixel;
        cudaStream_t st = it->stream;
        const size_t smem_first = stack_smem_bytes(kWalkBlock);
        const int stack_levels = walk_stack_levels(D, F);
        const size_t smem = stack_smem_bytes(kWalkBlock, stack_levels);
        int occ = 1;
        WOST_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_walk_persistent<D, RefOrder, F>, kWalkBlock, smem));
        const int grid_full = sm_count() * std::max(1, occ);
        const int refill_threshold = default_refill_threshold(D, F);
        WOST_CUDA_TRY(cudaMemsetAsync(it->d_acc.ptr, 0, sizeof(float4) * (size_t)S * std::max(1, it->n_active), st));
        WOST_CUDA_TRY(cudaMemsetAsync(it->d_steps.ptr, 0, sizeof(unsigned long long), st));
        it->launches = 0;
        it->step_launches = 0;
        it->step_kernel_ms = 0.0f;
        // sample batches: S at a time, cut right after every sample index the progressive metrics dump
        std::vector<std::pair<int, int>> batches;
        for (int s0 = 0; it->n_active > 0 && s0 < spp;)
        {
            const int Sb = it->frame_batch(s0, std::min(S, spp - s0));
            batches.emplace_back(s0, Sb);
            s0 += Sb;
        }
        it->slot_n.assign((size_t)S, 0);
        for (const auto &b : batches)
            for (int j = 0; j < b.second; ++j)
                ++it->slot_n[(size_t)j];
        const int nbatches = (int)batches.size();
        while ((int)it->batch_ev.size() < 2 * nbatches)
        {
            cudaEvent_t e;
            WOST_CUDA_TRY(cudaEventCreate(&e));
            it->batch_ev.push_back(e);
        }
        if (nbatches > 0)
        {
            if (it->d_first.count < (size_t)it->n_active && it->d_first.alloc((size_t)it->n_active) != cudaSuccess)
            {
                set_last_error("wost_solve: first-hit cache allocation failed");
                return WOST_ERR_CUDA;
            }
            const int grid = std::max(1, std::min(grid_full, div_up(it->n_active, kWalkBlock)));
            k_first_hit<D, RefOrder><<<grid, kWalkBlock, smem_first, st>>>(ctx, it->d_active.ptr, it->n_active, it->d_first.ptr, it->d_overflow.ptr);
            ++it->launches;
        }
        for (int bi = 0; bi < nbatches; ++bi)
        {
            const int s0 = batches[bi].first, Sb = batches[bi].second;
            WOST_CUDA_TRY(cudaMemsetAsync(it->d_counts.ptr, 0, sizeof(int), st));
            WOST_CUDA_TRY(cudaEventRecord(it->batch_ev[2 * bi], st));
            PersistArgs a;
            a.active = it->d_active.ptr;
            a.first = it->d_first.ptr;
            a.n_active = it->n_active;
            a.sample0 = (uint32_t)s0;
            a.batch_samples = Sb;
            a.acc = it->d_acc.ptr;
            a.max_depth = it->s.max_walking_depth;
            a.refill_threshold = refill_threshold;
            a.stack_levels = stack_levels;
            a.cursor = it->d_counts.ptr;
            a.steps = it->d_steps.ptr;
            a.overflow = it->d_overflow.ptr;
            const long long n0 = (long long)it->n_active * Sb;
            const int grid = (int)std::min<long long>(grid_full, (n0 + kWalkBlock - 1) / kWalkBlock);
            k_walk_persistent<D, RefOrder, F><<<std::max(1, grid), kWalkBlock, smem, st>>>(ctx, a);
            ++it->launches;
            ++it->step_launches;
            WOST_CUDA_TRY(cudaEventRecord(it->batch_ev[2 * bi + 1], st));
            if (it->frame_cb)
            {
                const wost_status frc = emit_frames(it, s0 + Sb - 1);
                if (frc != WOST_OK)
                    return frc;
            }
        }
        launch_resolve(it, S, (float)spp);
        ++it->launches;
        WOST_CUDA_TRY(cudaGetLastError());
        unsigned long long h_steps = 0;
        int h_over = 0;
        WOST_CUDA_TRY(cudaMemcpyAsync(&h_steps, it->d_steps.ptr, sizeof(h_steps), cudaMemcpyDeviceToHost, st));
        WOST_CUDA_TRY(cudaMemcpyAsync(&h_over, it->d_overflow.ptr, sizeof(int), cudaMemcpyDeviceToHost, st));
        WOST_CUDA_TRY(cudaEventRecord(it->ev1, st));
        WOST_CUDA_TRY(cudaStreamSynchronize(st));
        if (walk_steps)
            *walk_steps = h_steps;
        for (int bi = 0; bi < nbatches; ++bi)
        {
            float f = 0.0f;
            WOST_CUDA_TRY(cudaEventElapsedTime(&f, it->batch_ev[2 * bi], it->batch_ev[2 * bi + 1]));
            it->step_kernel_ms += f;
        }
        if (h_over)
        {
            cudaMemsetAsync(it->d_overflow.ptr, 0, sizeof(int), st);
            set_last_error("LBVH traversal stack overflow during solve");
            return WOST_ERR_OVERFLOW;
        }
        return WOST_OK;
    }
} // namespace wost

using namespace wost;

extern "C"
{
    wost_status wost_problem_create(const wost_problem_desc *d, wost_problem **out)
    {
        if (!d || !out || (d->dim != 2 && d->dim != 3))
        {
            set_last_error("wost_problem_create: invalid argument");
            return WOST_ERR_INVALID;
        }
        if ((d->dirichlet && d->dirichlet->dim != d->dim) || (d->neumann && d->neumann->dim != d->dim))
        {
            set_last_error("wost_problem_create: scene dimensionality mismatch");
            return WOST_ERR_INVALID;
        }
        wost_problem *p = new wost_problem();
        p->dim = d->dim;
        cudaGetDevice(&p->device);
        p->dirichlet = d->dirichlet;
        p->neumann = d->neumann;
        p->desc = *d;
        auto fail = [&](wost_status rc, const char *msg)
        {
            set_last_error(msg);
            delete p;
            return rc;
        };
        if (d->dirichlet && d->color_dirichlet)
        {
            if (d->n_color_dirichlet < d->dirichlet->nverts)
                return fail(WOST_ERR_INVALID, "wost_problem_create: fewer Dirichlet colour pairs than vertices");
            if (p->d_color_dir.upload(d->color_dirichlet, (size_t)d->n_color_dirichlet * 6) != cudaSuccess)
                return fail(WOST_ERR_CUDA, "wost_problem_create: colour upload failed");
        }
        if (d->neumann && d->color_neumann)
        {
            if (d->n_color_neumann < d->neumann->nverts)
                return fail(WOST_ERR_INVALID, "wost_problem_create: fewer Neumann colour pairs than vertices");
            if (p->d_color_neu.upload(d->color_neumann, (size_t)d->n_color_neumann * 6) != cudaSuccess)
                return fail(WOST_ERR_CUDA, "wost_problem_create: colour upload failed");
            for (size_t i = 0; i < (size_t)d->n_color_neumann * 6; ++i)
                if (d->color_neumann[i] != 0.0f)
                {
                    p->neumann_flux_enabled = true;
                    break;
                }
            if (d->neumann_intensity == 0.0f)
                p->neumann_flux_enabled = false;
        }
        if (d->source)
        {
            const size_t n = (size_t)d->source_dims[0] * d->source_dims[1] * d->source_dims[2] * 3;
            if (n == 0 || !(d->source_voxel > 0.0f))
                return fail(WOST_ERR_INVALID, "wost_problem_create: bad source grid");
            if (p->d_source.upload(d->source, n) != cudaSuccess)
                return fail(WOST_ERR_CUDA, "wost_problem_create: source upload failed");
        }
        if (d->mask)
        {
            p->mask_w = d->mask_w;
            p->mask_h = d->mask_h;
            if (p->d_mask.upload(d->mask, (size_t)d->mask_w * d->mask_h) != cudaSuccess)
                return fail(WOST_ERR_CUDA, "wost_problem_create: mask upload failed");
        }
        p->desc.color_dirichlet = nullptr;
        p->desc.color_neumann = nullptr;
        p->desc.source = nullptr;
        if (cudaDeviceSynchronize() != cudaSuccess)
            return fail(WOST_ERR_CUDA, "wost_problem_create: sync failed");
        // keep the mask on the host for the active-pixel list
        if (d->mask)
        {
            uint8_t *copy = new uint8_t[(size_t)d->mask_w * d->mask_h];
            memcpy(copy, d->mask, (size_t)d->mask_w * d->mask_h);
            p->desc.mask = copy;
        }
        *out = p;
        return WOST_OK;
    }
    void wost_problem_destroy(wost_problem *p)
    {
        if (!p)
            return;
        delete[] p->desc.mask;
        delete p;
    }

    wost_status wost_integrator_create(wost_problem *problem, const wost_settings *settings, int kind, wost_integrator **out)
    {
        if (!problem || !settings || !out)
        {
            set_last_error("wost_integrator_create: invalid argument");
            return WOST_ERR_INVALID;
        }
        if (kind != WOST_UNIFORM && kind != WOST_GUIDED)
        {
            set_last_error("wost_integrator_create: unknown integrator kind");
            return WOST_ERR_UNSUPPORTED;
        }
        const wost_settings &s = *settings;
        if (s.frame_w <= 0 || s.frame_h <= 0 || s.max_walking_depth <= 0 || s.samples_per_pixel <= 0)
        {
            set_last_error("wost_integrator_create: bad settings");
            return WOST_ERR_INVALID;
        }
        if (problem->desc.mask && (problem->mask_w != s.frame_w || problem->mask_h != s.frame_h))
        {
            // the reference reads a 1024^2 default mask out of bounds for larger frames (core/problem.cu:245-247)
            set_last_error("wost_integrator_create: mask size differs from frame size");
            return WOST_ERR_INVALID;
        }
        wost_integrator *it = new wost_integrator();
        it->problem = problem;
        it->s = s;
        it->kind = kind;
        it->dim = problem->dim;
        cudaGetDevice(&it->device);
        it->npix = s.frame_w * s.frame_h;
        // shard: contiguous pixel range and/or interleaved row tiles
        int pb = s.pixel_begin, pe = s.pixel_end > 0 ? s.pixel_end : it->npix;
        pb = std::max(0, std::min(pb, it->npix));
        pe = std::max(pb, std::min(pe, it->npix));
        std::vector<uint32_t> active;
        active.reserve(pe - pb);
        const uint8_t *mask = problem->desc.mask;
        // frame-wide count of evaluated pixels: identical on every rank whatever its pixel range / shard (it sizes the
        // data-parallel training batches, so that all ranks issue the same number of gradient all-reduces)
        if (mask)
        {
            int cnt = 0;
            for (int p = 0; p < it->npix; ++p) // branch-free: vectorised by the compiler
                cnt += mask[p] != 0;
            it->n_active_frame = cnt;
        }
        else
            it->n_active_frame = it->npix;
        const bool sharded = s.pixel_stride_rows > 0 && s.num_shards > 1;
        for (int row = pb / s.frame_w; row <= (pe - 1) / s.frame_w && pe > pb; ++row)
        {
            if (sharded && (row / s.pixel_stride_rows) % s.num_shards != s.shard)
                continue; // rows of other shards are skipped whole
            const int p0 = std::max(pb, row * s.frame_w), p1 = std::min(pe, (row + 1) * s.frame_w);
            for (int p = p0; p < p1; ++p)
                if (!mask || mask[p])
                    active.push_back((uint32_t)p);
        }
        it->n_active = (int)active.size();
        int S = s.spp_batch;
        if (S <= 0)
        {
            // auto: walks in flight per launch. Persistent uniform kernel: ~64M (16 B of accumulator per walk; fewer launches = fewer
            // drain phases: C3 2 454 M at 44 spp per launch, 2 505 M at 256). Wavefront modes and the guided integrator also hold
            // 64-100 B of queue per walk: ~16M.
            const bool persistent_uniform = kind == WOST_UNIFORM && s.walk_mode != WOST_MODE_REFERENCE_ORDER && s.walk_mode != WOST_MODE_WAVEFRONT;
            const long long target = persistent_uniform ? (64ll << 20) : (16ll << 20);
            S = (int)std::max<long long>(1, target / std::max(1, it->n_active));
        }
        S = std::max(1, std::min(S, s.samples_per_pixel));
        // the walk kernels claim jobs through 32-bit cursors: keep n_active * S below 2^31 (more samples simply take more batches)
        S = (int)std::min<long long>(S, std::max<long long>(1, ((1ll << 31) - 4096) / std::max(1, it->n_active)));
        it->spp_batch = S;
        it->queue_capacity = (size_t)std::max(1, it->n_active) * S;
        auto fail = [&](const char *msg)
        {
            set_last_error(std::string(msg) + ": " + cudaGetErrorString(cudaGetLastError()));
            delete it;
            return WOST_ERR_CUDA;
        };
        if (it->d_active.upload(active.data(), active.size()) != cudaSuccess)
            return fail("active pixel upload");
        // ping-pong queues: wavefront modes only (the persistent kernel keeps its walks in registers)
        for (int i = 0; i < 2 && kind == WOST_UNIFORM && (s.walk_mode == WOST_MODE_REFERENCE_ORDER || s.walk_mode == WOST_MODE_WAVEFRONT); ++i)
        {
            if (it->qa[i].alloc(it->queue_capacity) != cudaSuccess || it->qb[i].alloc(it->queue_capacity) != cudaSuccess)
                return fail("queue allocation");
            if (it->dim == 3 && it->qc[i].alloc(it->queue_capacity) != cudaSuccess)
                return fail("queue allocation");
        }
        if (kind == WOST_GUIDED && (it->d_cursor64.alloc(1) != cudaSuccess || guided_alloc(it) != WOST_OK))
            return fail("guided buffer allocation");
        if (it->d_counts.alloc((size_t)3 * (s.max_walking_depth + 2)) != cudaSuccess || it->d_steps.alloc(1) != cudaSuccess ||
            it->d_acc.alloc((size_t)S * std::max(1, it->n_active)) != cudaSuccess || it->d_overflow.alloc(96) != cudaSuccess)
            return fail("buffer allocation");
        for (int c = 0; c < WOST_CH_COUNT; ++c)
        {
            if (it->film[c].alloc((size_t)it->npix * 5) != cudaSuccess)
                return fail("film allocation");
            if (c == WOST_CH_SOLUTION)
                k_film_init<<<div_up(it->npix, 256), 256>>>(it->npix, it->film[c].ptr); // pixels outside the mask / shard: solution 0, weight 1
            else
                cudaMemset(it->film[c].ptr, 0, sizeof(float) * (size_t)it->npix * 5);
        }
        cudaMemset(it->d_overflow.ptr, 0, 96 * sizeof(int));
        if (cudaStreamCreateWithFlags(&it->stream, cudaStreamNonBlocking) != cudaSuccess || cudaEventCreate(&it->ev0) != cudaSuccess ||
            cudaEventCreate(&it->ev1) != cudaSuccess)
            return fail("stream/event creation");
        apply_l2_window(it);
        if (cudaDeviceSynchronize() != cudaSuccess)
            return fail("sync");
        *out = it;
        return WOST_OK;
    }
    void wost_integrator_destroy(wost_integrator *it) { delete it; }

    wost_status wost_solve(wost_integrator *it, uint64_t *ms, uint64_t *walk_steps, float *device_ms)
    {
        if (!it)
            return WOST_ERR_INVALID;
        const auto t0 = std::chrono::high_resolution_clock::now();
        it->solve_t0 = t0;
        WOST_CUDA_TRY(cudaEventRecord(it->ev0, it->stream));
        wost_status rc;
        if (it->kind == WOST_GUIDED)
            rc = solve_guided(it, walk_steps);
        else
        switch (it->s.walk_mode)
        {
        case WOST_MODE_REFERENCE_ORDER:
            rc = (it->dim == 2) ? solve_uniform<2, true>(it, walk_steps) : solve_uniform<3, true>(it, walk_steps);
            break;
        case WOST_MODE_WAVEFRONT:
            rc = (it->dim == 2) ? solve_uniform<2, false>(it, walk_steps) : solve_uniform<3, false>(it, walk_steps);
            break;
        case WOST_MODE_PERSISTENT_REFERENCE_ORDER:
            rc = (it->dim == 2) ? solve_persistent<2, true>(it, walk_steps) : solve_persistent<3, true>(it, walk_steps);
            break;
        default:
            rc = (it->dim == 2) ? solve_persistent<2, false>(it, walk_steps) : solve_persistent<3, false>(it, walk_steps);
            break;
        }
        if (rc != WOST_OK)
            return rc;
        const auto t1 = std::chrono::high_resolution_clock::now();
        if (ms)
            *ms = (uint64_t)std::chrono::duration_cast<std::chrono::milliseconds>(t1 - t0).count();
        if (device_ms)
        {
            float f = 0.0f;
            WOST_CUDA_TRY(cudaEventElapsedTime(&f, it->ev0, it->ev1));
            *device_ms = f;
        }
        return WOST_OK;
    }

    wost_status wost_set_frame_callback(wost_integrator *it, wost_frame_fn fn, void *user)
    {
        if (!it)
            return WOST_ERR_INVALID;
        it->frame_cb = fn;
        it->frame_user = user;
        return WOST_OK;
    }

    wost_status wost_render_channel(wost_integrator *it, int channel)
    {
        if (!it || channel <= WOST_CH_SOLUTION || channel >= WOST_CH_COUNT)
        {
            set_last_error("wost_render_channel: invalid channel");
            return WOST_ERR_INVALID;
        }
        const size_t smem = stack_smem_bytes(kWalkBlock);
        const int grid = std::max(1, std::min(div_up(it->npix, kWalkBlock), sm_count() * 8));
        if (it->dim == 2)
        {
            WalkCtx<2> ctx;
            it->problem->fill_ctx(ctx);
            ctx.frame_w = it->s.frame_w, ctx.frame_h = it->s.frame_h, ctx.max_depth = it->s.max_walking_depth, ctx.eps = it->s.epsilon_shell, ctx.seed = it->s.seed;
            k_render_channel<2><<<grid, kWalkBlock, smem, it->stream>>>(ctx, channel, it->npix, it->film[channel].ptr, it->d_overflow.ptr);
        }
        else
        {
            WalkCtx<3> ctx;
            it->problem->fill_ctx(ctx);
            ctx.frame_w = it->s.frame_w, ctx.frame_h = it->s.frame_h, ctx.max_depth = it->s.max_walking_depth, ctx.eps = it->s.epsilon_shell, ctx.seed = it->s.seed;
            k_render_channel<3><<<grid, kWalkBlock, smem, it->stream>>>(ctx, channel, it->npix, it->film[channel].ptr, it->d_overflow.ptr);
        }
        WOST_CUDA_TRY(cudaGetLastError());
        WOST_CUDA_TRY(cudaStreamSynchronize(it->stream));
        return WOST_OK;
    }

    wost_status wost_film_download(wost_integrator *it, int channel, float *out)
    {
        if (!it || !out || channel < 0 || channel >= WOST_CH_COUNT)
            return WOST_ERR_INVALID;
        WOST_CUDA_TRY(cudaMemcpyAsync(out, it->film[channel].ptr, sizeof(float) * (size_t)it->npix * 5, cudaMemcpyDeviceToHost, it->stream));
        WOST_CUDA_TRY(cudaStreamSynchronize(it->stream));
        return WOST_OK;
    }
    wost_status wost_film_stderr(wost_integrator *it, float *out_se_rgb)
    {
        if (!it || !out_se_rgb)
            return WOST_ERR_INVALID;
        if (it->slot_n.empty())
        {
            set_last_error("wost_film_stderr: no solve has run on this integrator");
            return WOST_ERR_INVALID;
        }
        const size_t n = (size_t)it->npix * 3;
        if (it->d_se.count < n && it->d_se.alloc(n) != cudaSuccess)
        {
            set_last_error("wost_film_stderr: allocation failed");
            return WOST_ERR_CUDA;
        }
        if (it->d_slot_n.upload(it->slot_n.data(), it->slot_n.size()) != cudaSuccess)
        {
            set_last_error("wost_film_stderr: upload failed");
            return WOST_ERR_CUDA;
        }
        WOST_CUDA_TRY(cudaMemsetAsync(it->d_se.ptr, 0, sizeof(float) * n, it->stream));
        if (it->n_active > 0)
            k_resolve<<<div_up(it->n_active, 256), 256, 0, it->stream>>>(it->n_active, (int)it->slot_n.size(), it->d_acc.ptr, it->d_active.ptr,
                                                                         (float)it->s.samples_per_pixel, it->film[WOST_CH_SOLUTION].ptr, it->d_slot_n.ptr, it->d_se.ptr);
        WOST_CUDA_TRY(cudaGetLastError());
        WOST_CUDA_TRY(cudaMemcpyAsync(out_se_rgb, it->d_se.ptr, sizeof(float) * n, cudaMemcpyDeviceToHost, it->stream));
        WOST_CUDA_TRY(cudaStreamSynchronize(it->stream));
        return WOST_OK;
    }
    wost_status wost_film_device_ptr(wost_integrator *it, int channel, void **d_ptr)
    {
        if (!it || !d_ptr || channel < 0 || channel >= WOST_CH_COUNT)
            return WOST_ERR_INVALID;
        *d_ptr = it->film[channel].ptr;
        return WOST_OK;
    }
    wost_status wost_traversal_stats(wost_integrator *it, uint64_t *full_reference_order, uint64_t *restricted_second_pass, int reset)
    {
        if (!it)
            return WOST_ERR_INVALID;
        int h[4] = {0, 0, 0, 0};
        WOST_CUDA_TRY(cudaMemcpyAsync(h, it->d_overflow.ptr, sizeof(h), cudaMemcpyDeviceToHost, it->stream));
        WOST_CUDA_TRY(cudaStreamSynchronize(it->stream));
        if (full_reference_order)
            *full_reference_order = (uint64_t)(unsigned)h[1];
        if (restricted_second_pass)
            *restricted_second_pass = (uint64_t)(unsigned)h[2];
        if (reset)
            WOST_CUDA_TRY(cudaMemsetAsync(it->d_overflow.ptr + 1, 0, 2 * sizeof(int), it->stream));
        return WOST_OK;
    }
    // diagnostics (library built with -DWOST_TRAV_DEBUG); the caller's buffer holds 48 words, see include/wost.h
    wost_status wost_debug_counters(wost_integrator *it, unsigned int *out48)
    {
        if (!it || !out48)
            return WOST_ERR_INVALID;
        // out: [0..15] trip / query counts, [16..47] 16 x 64-bit warp-cycle sums (trip iterations, tail iterations) per decade
        WOST_CUDA_TRY(cudaMemcpy(out48, it->d_overflow.ptr + 8, 16 * sizeof(int), cudaMemcpyDeviceToHost));
        WOST_CUDA_TRY(cudaMemcpy(out48 + 16, it->d_overflow.ptr + 32, 32 * sizeof(int), cudaMemcpyDeviceToHost));
        WOST_CUDA_TRY(cudaMemset(it->d_overflow.ptr + 8, 0, 88 * sizeof(int)));
        return WOST_OK;
    }
    wost_status wost_l2_window_bytes(wost_integrator *it, uint64_t *bytes)
    {
        if (!it || !bytes)
            return WOST_ERR_INVALID;
        *bytes = (uint64_t)it->l2_window_bytes;
        return WOST_OK;
    }
    wost_status wost_last_launch_count(wost_integrator *it, uint64_t *kernel_launches)
    {
        if (!it || !kernel_launches)
            return WOST_ERR_INVALID;
        *kernel_launches = it->launches;
        return WOST_OK;
    }
    wost_status wost_last_step_kernel_time(wost_integrator *it, uint64_t *step_launches, float *step_kernel_ms)
    {
        if (!it || !step_launches || !step_kernel_ms)
            return WOST_ERR_INVALID;
        *step_launches = it->step_launches;
        *step_kernel_ms = it->step_kernel_ms;
        return WOST_OK;
    }
}
