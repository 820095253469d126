// Counter-based RNG for the walker: Philox-4x32-10 (Salmon et al., SC'11), host + device.
//
// Replaces the reference's stateful per-pixel PCG32 kept in global memory
// (core/sampler.h:10-103: every get1D() is a 16 B load + 8 B store through a tagged pointer).
// Here a draw is a pure function of (seed, pixel, sample, depth, purpose, iteration): no RNG state is
// ever stored, the result is independent of scheduling and of the number of GPUs, and the scalar
// CPU oracle reproduces the device stream bit for bit.
//
// u32 -> float uses the reference's mapping (core/sampler.h:87-98): (u >> 9 | 0x3f800000) - 1 in [0,1).
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define WOST_HD __host__ __device__ __forceinline__
#else
#define WOST_HD inline
#endif

struct Philox4
{
    uint32_t x, y, z, w;
};

WOST_HD uint32_t philox_mulhi(uint32_t a, uint32_t b)
{
#if defined(__CUDA_ARCH__)
    return __umulhi(a, b);
#else
    return (uint32_t)(((uint64_t)a * (uint64_t)b) >> 32);
#endif
}

WOST_HD Philox4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1)
{
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int r = 0; r < 10; ++r)
    {
        const uint32_t hi0 = philox_mulhi(M0, c0), lo0 = M0 * c0;
        const uint32_t hi1 = philox_mulhi(M1, c2), lo1 = M1 * c2;
        const uint32_t n0 = hi1 ^ c1 ^ k0;
        const uint32_t n2 = hi0 ^ c3 ^ k1;
        c0 = n0;
        c1 = lo1;
        c2 = n2;
        c3 = lo0;
        k0 += W0;
        k1 += W1;
    }
    Philox4 o;
    o.x = c0;
    o.y = c1;
    o.z = c2;
    o.w = c3;
    return o;
}

WOST_HD float u32_to_unit_float(uint32_t u)
{
    union
    {
        uint32_t u;
        float f;
    } v;
    v.u = (u >> 9) | 0x3f800000u;
    return v.f - 1.0f;
}

// Purposes (the 3rd counter word): which consumer inside one walk step draws the numbers.
enum WostRngPurpose
{
    WOST_RNG_WALK = 0,      // direction for the next step (oneStepWalk / uniform sampling)
    WOST_RNG_NEUMANN = 1,   // boundary-sample descent + point on primitive
    WOST_RNG_SOURCE_DIR = 2, // direction of the source sample
    WOST_RNG_SOURCE_RAD = 3, // radial rejection loop (4th counter word = iteration)
    WOST_RNG_SELECT = 4,    // guided: selection draw + lobe choice
    WOST_RNG_GUIDED = 5,    // guided: lobe sampling (4th counter word = rejection iteration)
    WOST_RNG_TRAIN = 6      // training-set shuffling / pixel offset
};

struct WostRngKey
{
    uint32_t k0, k1;
};

// key = (seed, pixel); counter = (sample, depth, purpose, iteration)
WOST_HD Philox4 wost_rng(uint32_t seed, uint32_t pixel, uint32_t sample, uint32_t depth, uint32_t purpose, uint32_t iter)
{
    return philox4x32_10(sample, depth, purpose, iter, seed, pixel);
}
