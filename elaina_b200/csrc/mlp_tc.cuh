// tcgen05 / TMEM building blocks of the guiding MLP (sm_100a only) and the fused forward pass of one 128-row tile.
//
// Shape of the problem: rows = walk items (M = 128 per CTA tile), K = 16..64, N = 64 / 48 -- four tiny GEMMs back to
// back with a ReLU between them. One CTA = 128 threads = 4 warps; thread r owns row r end to end:
//   * operands live in shared memory in the UMMA "K-major, no swizzle" canonical layout, stored chunk-major:
//       byte offset(row r, column k) = (k / 8) * (R * 16) + r * 16 + (k % 8) * 2        (R rows, 16-byte chunks)
//     i.e. core matrices (8 rows x 16 bytes) are contiguous, 8-row groups are 128 B apart (SBO) and K-chunks are
//     R * 16 B apart (LBO). A thread writes its row as one 16-byte store per chunk: consecutive lanes hit consecutive
//     16-byte slots, so stores are conflict-free, and the same buffer doubles as an MN-major operand (K = rows) for
//     the weight-gradient GEMMs of training;
//   * one elected thread issues tcgen05.mma (kind::f16, fp16 x fp16 -> fp32) K/16 times per layer into a TMEM
//     accumulator of N columns; completion is signalled with tcgen05.commit on an mbarrier;
//   * warp w reads TMEM lanes 32w..32w+31 with tcgen05.ld.32x32b (lane == row == thread), applies ReLU, converts to
//     fp16 and writes the next layer's A operand.
// Weights are staged once per CTA (<= 26 KB fp16, already in canonical layout in global memory, see guiding.cuh).
#pragma once
#include "guiding.cuh"

namespace wost
{
    namespace tc
    {
        WOST_DEV uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

        WOST_DEV void mbar_init(uint64_t *bar, uint32_t count)
        {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
        }
        WOST_DEV void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
        WOST_DEV void mbar_wait(uint64_t *bar, uint32_t parity)
        {
            const uint32_t a = smem_u32(bar);
            uint32_t done, spins = 0;
            do
            {
                if (++spins > (1u << 22)) // a lost tcgen05.commit would otherwise hang the GPU: fail loudly instead
                    __trap();
                asm volatile(
                    "{\n\t"
                    ".reg .pred p;\n\t"
                    "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                    "selp.u32 %0, 1, 0, p;\n\t"
                    "}"
                    : "=r"(done)
                    : "r"(a), "r"(parity)
                    : "memory");
            } while (!done);
        }
        // TMA bulk copy (1-D, cp.async.bulk): one thread moves `bytes` (multiple of 16, both addresses 16-byte aligned) from
        // global to shared memory through the async proxy and the mbarrier counts the bytes -- the same proxy tcgen05.mma reads
        // operands through, so no generic-proxy staging loop and no proxy fence for the weights
        WOST_DEV void mbar_expect_tx(uint64_t *bar, uint32_t bytes)
        {
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
        }
        WOST_DEV void bulk_g2s(void *smem_dst, const void *gsrc, uint32_t bytes, uint64_t *bar)
        {
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(smem_dst)),
                         "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
                         : "memory");
        }
        WOST_DEV void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
        WOST_DEV void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
        WOST_DEV void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

        // one full warp allocates `cols` (power of two >= 32) TMEM columns; the base address lands in *slot (smem)
        WOST_DEV void tmem_alloc(uint32_t *slot, uint32_t cols)
        {
            asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot)), "r"(cols) : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
        }
        WOST_DEV void tmem_dealloc(uint32_t taddr, uint32_t cols)
        {
            asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
        }

        // Shared-memory matrix descriptor (cute::UMMA::SmemDescriptor): start >> 4 [0,14), leading byte offset >> 4
        // [16,30), stride byte offset >> 4 [32,46), version 1 [46,48), layout type 0 = no swizzle [61,64).
        // K-major operand of R rows stored chunk-major: LBO = R * 16 (K-chunk stride), SBO = 128 (8-row group stride).
        // MN-major operand (K = rows of the same buffer): LBO = 128 (8-row k-group stride), SBO = R * 16 (mn-chunk stride).
        WOST_DEV uint64_t make_desc(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes)
        {
            uint64_t d = 0;
            d |= (uint64_t)((smem_addr >> 4) & 0x3FFFu);
            d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16;
            d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32;
            d |= (uint64_t)1 << 46;
            return d;
        }
        // Instruction descriptor (cute::UMMA::InstrDescriptor) for kind::f16: C = F32 [4,6) = 1, A/B = F16 (0),
        // a_major [15], b_major [16] (0 = K-major, 1 = MN-major), N >> 3 at [17,23), M >> 4 at [24,29).
        WOST_HOSTDEV uint32_t make_idesc(int M, int N, int a_mn_major, int b_mn_major)
        {
            return (1u << 4) | ((uint32_t)a_mn_major << 15) | ((uint32_t)b_mn_major << 16) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
        }
        // D[tmem] (+)= A[smem] * B[smem]^T ; issued by ONE thread
        WOST_DEV void mma_f16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate)
        {
            asm volatile(
                "{\n\t"
                ".reg .pred p;\n\t"
                "setp.ne.b32 p, %4, 0;\n\t"
                "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
                "}" ::"r"(tmem_d),
                "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
                : "memory");
        }
        // arrive on the mbarrier when every previously issued tcgen05.mma of this thread has completed
        WOST_DEV void mma_commit(uint64_t *bar)
        {
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
        }
        // 16 consecutive fp32 columns of this thread's TMEM lane (warp w may only touch lanes 32w..32w+31)
        WOST_DEV void tmem_ld16(uint32_t taddr, float *v)
        {
            uint32_t r[16];
            asm volatile(
                "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
                  "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                : "r"(taddr)
                : "memory");
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
            for (int i = 0; i < 16; ++i)
                v[i] = __uint_as_float(r[i]);
        }
        WOST_DEV uint32_t pack_half2(float a, float b)
        {
            const __half2 h = __floats2half2_rn(a, b);
            return *reinterpret_cast<const uint32_t *>(&h);
        }
        // write 8 consecutive columns [k0, k0+8) of row r into a chunk-major operand buffer of R rows
        WOST_DEV void store_chunk(__half *buf, int R, int r, int chunk, const float *v8)
        {
            uint4 u;
            u.x = pack_half2(v8[0], v8[1]);
            u.y = pack_half2(v8[2], v8[3]);
            u.z = pack_half2(v8[4], v8[5]);
            u.w = pack_half2(v8[6], v8[7]);
            *reinterpret_cast<uint4 *>(reinterpret_cast<char *>(buf) + (size_t)chunk * (R * 16) + (size_t)r * 16) = u;
        }
    } // namespace tc

    // Shared-memory carve-up of the forward pass (inference): weights + one activation buffer + sync words.
    constexpr int kTmemColsFwd = 64;
    template <int K0>
    struct MlpFwdSmem
    {
        static constexpr int kWeightHalfs = kHidden * K0 + 2 * kHidden * kHidden + kOutPad * kHidden;
        static constexpr size_t kBytes = (size_t)kWeightHalfs * 2 + (size_t)kTileRows * kHidden * 2 + 32;
        __half *w;   // canonical weights W0|W1|W2|W3
        __half *a;   // 128 x 64 activation operand
        uint64_t *bar;
        uint32_t *tmem_slot;
        WOST_DEV void carve(char *base)
        {
            w = reinterpret_cast<__half *>(base);
            a = reinterpret_cast<__half *>(base + (size_t)kWeightHalfs * 2);
            bar = reinterpret_cast<uint64_t *>(base + (size_t)kWeightHalfs * 2 + (size_t)kTileRows * kHidden * 2);
            tmem_slot = reinterpret_cast<uint32_t *>(bar + 1);
        }
    };

    // CTA-wide one-time setup: stage the weights, init the mbarrier, allocate TMEM. Returns the TMEM base address.
    template <int K0>
    WOST_DEV uint32_t mlp_fwd_setup(MlpFwdSmem<K0> &sm, const __half *weights_canonical)
    {
        uint64_t *wbar = sm.bar + 2; // second mbarrier of the sync words: weight arrival
        if (threadIdx.x == 0)
        {
            tc::mbar_init(sm.bar, 1);
            tc::mbar_init(wbar, 1);
            tc::fence_mbar_init();
            // weights: ONE TMA bulk copy (26 KB, already in the canonical UMMA layout in global memory)
            constexpr uint32_t kWBytes = (uint32_t)MlpFwdSmem<K0>::kWeightHalfs * 2u;
            static_assert(kWBytes % 16u == 0u, "bulk copy size");
            tc::mbar_expect_tx(wbar, kWBytes);
            tc::bulk_g2s(sm.w, weights_canonical, kWBytes, wbar);
        }
        if (threadIdx.x < 32)
            tc::tmem_alloc(sm.tmem_slot, kTmemColsFwd);
        tc::tc_fence_before();
        __syncthreads();
        tc::tc_fence_after();
        tc::mbar_wait(wbar, 0);
        return *sm.tmem_slot;
    }
    template <int K0>
    WOST_DEV void mlp_fwd_teardown(MlpFwdSmem<K0> &, uint32_t tmem_base)
    {
        tc::tc_fence_before();
        __syncthreads();
        if (threadIdx.x < 32)
            tc::tmem_dealloc(tmem_base, kTmemColsFwd);
    }

    // Forward pass of one tile. Every thread of the CTA (128) calls it with its row's encoded features enc[K0];
    // on return out[48] holds the row's raw network outputs (fp32). `phase` is the CTA-uniform mbarrier parity.
    template <int K0>
    WOST_DEV void mlp_fwd_tile(MlpFwdSmem<K0> &sm, uint32_t tmem_base, uint32_t &phase, const float *enc, float *out)
    {
        const int r = threadIdx.x;
        const uint32_t lane_base = tmem_base + ((uint32_t)(r & ~31) << 16);
#pragma unroll
        for (int c = 0; c < K0 / 8; ++c)
            tc::store_chunk(sm.a, kTileRows, r, c, enc + c * 8);
#pragma unroll
        for (int l = 0; l < 4; ++l)
        {
            const int K = l == 0 ? K0 : kHidden;
            const int N = l == 3 ? kOutPad : kHidden;
            const int woff = l == 0 ? 0 : (kHidden * K0 + (l - 1) * kHidden * kHidden);
            tc::fence_proxy_async();
            tc::tc_fence_before();
            __syncthreads(); // operand complete; previous accumulator fully read
            if (r == 0)
            {
                tc::tc_fence_after();
                const uint32_t idesc = tc::make_idesc(kTileRows, N, 0, 0);
                const uint32_t a_addr = tc::smem_u32(sm.a), b_addr = tc::smem_u32(sm.w + woff);
#pragma unroll
                for (int k = 0; k < K / 16; ++k)
                {
                    const uint64_t da = tc::make_desc(a_addr + k * 2 * (kTileRows * 16), kTileRows * 16, 128);
                    const uint64_t db = tc::make_desc(b_addr + k * 2 * (N * 16), N * 16, 128);
                    tc::mma_f16(tmem_base, da, db, idesc, k > 0 ? 1u : 0u);
                }
                tc::mma_commit(sm.bar);
            }
            tc::mbar_wait(sm.bar, phase);
            phase ^= 1u;
            tc::tc_fence_after();
            if (l < 3)
            {
#pragma unroll
                for (int c16 = 0; c16 < kHidden / 16; ++c16)
                {
                    float v[16];
                    tc::tmem_ld16(lane_base + c16 * 16, v);
#pragma unroll
                    for (int i = 0; i < 16; ++i)
                        v[i] = fmaxf(v[i], 0.0f);
                    tc::store_chunk(sm.a, kTileRows, r, c16 * 2, v);
                    tc::store_chunk(sm.a, kTileRows, r, c16 * 2 + 1, v + 8);
                }
            }
            else
            {
#pragma unroll
                for (int c16 = 0; c16 < kOutPad / 16; ++c16)
                    tc::tmem_ld16(lane_base + c16 * 16, out + c16 * 16);
            }
        }
        // the next tile's first operand store must not overtake other threads' reads of this tile: the stores of the
        // next call are followed by __syncthreads before any MMA, and every thread has finished its tcgen05.ld here.
    }
} // namespace wost
