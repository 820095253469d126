// Deterministic single-precision elementary functions, identical on host (g++ -ffp-contract=off)
// and device (nvcc -fmad=false): every operation below is an IEEE-754 fp32 add/mul/div/sqrt or an
// integer/bit operation, so x86-64 and sm_100a return the same bits.
//
// Why: the reference's geometry code calls libm / CUDA-libdevice transcendentals (acos/asin/atan2 in
// the cone test ext/lbvh/.../core/cone.cuh:78-121, cos/sin in the cone merge :266-318, log in
// util/green.h:25-28, cos/sin in util/sampling.h:28-33). glibc and libdevice differ by ulps, which
// would make "same walk on CPU oracle and GPU" impossible. With one arithmetic the oracle reproduces the
// device walk trajectories bit for bit; accuracy is ~1-2 ulp (Cephes-style minimax polynomials), which is
// tested against libm in tests/test_detmath.py.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define DM_HD __host__ __device__ __forceinline__
#else
#define DM_HD inline
#endif

#define DM_PI 3.14159265358979323846f
#define DM_PI_2 1.57079632679489661923f
#define DM_PI_4 0.78539816339744830962f
#define DM_2PI 6.28318530717958647692f
#define DM_4PI 12.5663706143591729539f

DM_HD uint32_t dm_f2u(float f)
{
    union
    {
        float f;
        uint32_t u;
    } v;
    v.f = f;
    return v.u;
}
DM_HD float dm_u2f(uint32_t u)
{
    union
    {
        float f;
        uint32_t u;
    } v;
    v.u = u;
    return v.f;
}
DM_HD float dm_abs(float x) { return dm_u2f(dm_f2u(x) & 0x7fffffffu); }
DM_HD float dm_copysign(float mag, float sgn) { return dm_u2f((dm_f2u(mag) & 0x7fffffffu) | (dm_f2u(sgn) & 0x80000000u)); }
DM_HD bool dm_isnan(float x) { return (dm_f2u(x) & 0x7fffffffu) > 0x7f800000u; }
DM_HD bool dm_isinf(float x) { return (dm_f2u(x) & 0x7fffffffu) == 0x7f800000u; }
DM_HD float dm_inf() { return dm_u2f(0x7f800000u); }
DM_HD float dm_nan() { return dm_u2f(0x7fc00000u); }

// IEEE sqrt: correctly rounded on both sides (sqrtf with -prec-sqrt=true / SSE sqrtss).
#if defined(__CUDA_ARCH__)
DM_HD float dm_sqrt(float x) { return __fsqrt_rn(x); }
DM_HD float dm_div(float a, float b) { return __fdiv_rn(a, b); }
#else
#include <math.h>
DM_HD float dm_sqrt(float x) { return sqrtf(x); }
DM_HD float dm_div(float a, float b) { return a / b; }
#endif

// std::min / std::max semantics of the reference call sites (not NaN-suppressing): (b < a) ? b : a
DM_HD float dm_min(float a, float b) { return (b < a) ? b : a; }
DM_HD float dm_max(float a, float b) { return (a < b) ? b : a; }
// ::fmin / ::fmax semantics (NaN-suppressing, IEEE minNum/maxNum), written out so that host and device agree on
// the sign of zero as well (used where the result is stored and compared bit for bit: box merges of the build)
DM_HD float dm_fmin_strict(float a, float b)
{
    if (dm_isnan(a))
        return b;
    if (dm_isnan(b))
        return a;
    return (b < a) ? b : a;
}
DM_HD float dm_fmax_strict(float a, float b)
{
    if (dm_isnan(a))
        return b;
    if (dm_isnan(b))
        return a;
    return (a < b) ? b : a;
}
// Hot-path flavour: one FMNMX on the device (same NaN suppression; may differ from the host only in the sign of a
// zero result, which no caller can observe: the values are squared, compared or clamped to positive numbers).
#if defined(__CUDA_ARCH__)
DM_HD float dm_fmin(float a, float b) { return fminf(a, b); }
DM_HD float dm_fmax(float a, float b) { return fmaxf(a, b); }
#else
DM_HD float dm_fmin(float a, float b) { return dm_fmin_strict(a, b); }
DM_HD float dm_fmax(float a, float b) { return dm_fmax_strict(a, b); }
#endif

// ---------------------------------------------------------------- sin / cos
// kernel on |r| <= pi/4
DM_HD float dm_sin_kernel(float r)
{
    const float z = r * r;
    float p = -1.9515295891e-4f;
    p = p * z + 8.3321608736e-3f;
    p = p * z + -1.6666654611e-1f;
    return (p * z) * r + r;
}
DM_HD float dm_cos_kernel(float r)
{
    const float z = r * r;
    float p = 2.443315711809948e-5f;
    p = p * z + -1.388731625493765e-3f;
    p = p * z + 4.166664568298827e-2f;
    return (p * z) * z + (1.0f - 0.5f * z);
}

// sin and cos of x (radians), |x| up to ~1e5 with 3-term Cody-Waite reduction
DM_HD void dm_sincos(float x, float *s, float *c)
{
    const float ax = dm_abs(x);
    // j = nearest integer to ax / (pi/2)
    int j = (int)(ax * 0.636619772367581343f + 0.5f);
    const float fj = (float)j;
    // pi/2 split in three parts (Cephes DP1..DP3 doubled)
    float r = ax - fj * 1.5703125f;
    r = r - fj * 4.837512969970703125e-4f;
    r = r - fj * 7.54978995489188e-8f;
    const float sk = dm_sin_kernel(r);
    const float ck = dm_cos_kernel(r);
    float ss, cc;
    switch (j & 3)
    {
    case 0:
        ss = sk;
        cc = ck;
        break;
    case 1:
        ss = ck;
        cc = -sk;
        break;
    case 2:
        ss = -sk;
        cc = -ck;
        break;
    default:
        ss = -ck;
        cc = sk;
        break;
    }
    *s = (x < 0.0f) ? -ss : ss;
    *c = cc;
}
DM_HD float dm_sin(float x)
{
    float s, c;
    dm_sincos(x, &s, &c);
    return s;
}
DM_HD float dm_cos(float x)
{
    float s, c;
    dm_sincos(x, &s, &c);
    return c;
}

// ---------------------------------------------------------------- log / exp
DM_HD float dm_log(float x)
{
    if (dm_isnan(x))
        return x;
    if (x < 0.0f)
        return dm_nan();
    if (x == 0.0f)
        return -dm_inf();
    if (dm_isinf(x))
        return x;
    uint32_t ix = dm_f2u(x);
    int e = 0;
    if (ix < 0x00800000u)
    { // subnormal: scale by 2^23
        x = x * 8388608.0f;
        ix = dm_f2u(x);
        e = -23;
    }
    e += (int)(ix >> 23) - 126;                       // x = m * 2^e, m in [0.5, 1)
    float m = dm_u2f((ix & 0x007fffffu) | 0x3f000000u);
    if (m < 0.707106781186547524f)
    {
        e -= 1;
        m = m + m - 1.0f;
    }
    else
    {
        m = m - 1.0f;
    }
    const float z = m * m;
    float y = 7.0376836292e-2f;
    y = y * m + -1.1514610310e-1f;
    y = y * m + 1.1676998740e-1f;
    y = y * m + -1.2420140846e-1f;
    y = y * m + 1.4249322787e-1f;
    y = y * m + -1.6668057665e-1f;
    y = y * m + 2.0000714765e-1f;
    y = y * m + -2.4999993993e-1f;
    y = y * m + 3.3333331174e-1f;
    y = y * m * z;
    const float fe = (float)e;
    y = y + -2.12194440e-4f * fe;
    y = y + -0.5f * z;
    float r = m + y;
    r = r + 0.693359375f * fe;
    return r;
}

DM_HD float dm_exp(float x)
{
    if (dm_isnan(x))
        return x;
    if (x > 88.72283905206835f)
        return dm_inf();
    if (x < -103.278929903431851103f)
        return 0.0f;
    // n = round(x / ln2)
    float fn = x * 1.44269504088896341f;
    fn = (fn < 0.0f) ? fn - 0.5f : fn + 0.5f;
    const int n = (int)fn;
    fn = (float)n;
    float r = x - fn * 0.693359375f;
    r = r - fn * -2.12194440e-4f;
    const float z = r * r;
    float p = 1.9875691500e-4f;
    p = p * r + 1.3981999507e-3f;
    p = p * r + 8.3334519073e-3f;
    p = p * r + 4.1665795894e-2f;
    p = p * r + 1.6666665459e-1f;
    p = p * r + 5.0000001201e-1f;
    float y = p * z + r + 1.0f;
    // scale by 2^n in two steps so that subnormal results round once more but never overflow early
    int n1 = n / 2, n2 = n - n1;
    y = y * dm_u2f((uint32_t)(n1 + 127) << 23);
    y = y * dm_u2f((uint32_t)(n2 + 127) << 23);
    return y;
}

// ---------------------------------------------------------------- atan / atan2 / asin / acos
DM_HD float dm_atan(float xx)
{
    if (dm_isnan(xx))
        return xx;
    float x = dm_abs(xx);
    float y;
    if (x > 2.414213562373095f)
    { // tan(3pi/8)
        y = DM_PI_2;
        x = -dm_div(1.0f, x);
    }
    else if (x > 0.4142135623730950f)
    { // tan(pi/8)
        y = DM_PI_4;
        x = dm_div(x - 1.0f, x + 1.0f);
    }
    else
    {
        y = 0.0f;
    }
    const float z = x * x;
    float p = 8.05374449538e-2f;
    p = p * z + -1.38776856032e-1f;
    p = p * z + 1.99777106478e-1f;
    p = p * z + -3.33329491539e-1f;
    y = y + (p * z * x + x);
    return (xx < 0.0f) ? -y : y;
}

DM_HD float dm_atan2(float y, float x)
{
    if (dm_isnan(x) || dm_isnan(y))
        return dm_nan();
    const bool xneg = (dm_f2u(x) >> 31) != 0;
    if (y == 0.0f)
        return xneg ? dm_copysign(DM_PI, y) : dm_copysign(0.0f, y);
    if (x == 0.0f)
        return dm_copysign(DM_PI_2, y);
    if (dm_isinf(x))
    {
        if (dm_isinf(y))
            return dm_copysign(xneg ? 3.0f * DM_PI_4 : DM_PI_4, y);
        return xneg ? dm_copysign(DM_PI, y) : dm_copysign(0.0f, y);
    }
    if (dm_isinf(y))
        return dm_copysign(DM_PI_2, y);
    const float a = dm_atan(dm_div(y, x));
    if (!xneg)
        return a;
    return (y < 0.0f) ? a - DM_PI : a + DM_PI;
}

DM_HD float dm_asin(float xx)
{
    if (dm_isnan(xx))
        return xx;
    const float a = dm_abs(xx);
    if (a > 1.0f)
        return dm_nan();
    float x, z;
    bool big = false;
    if (a > 0.5f)
    {
        z = 0.5f * (1.0f - a);
        x = dm_sqrt(z);
        big = true;
    }
    else
    {
        x = a;
        z = x * x;
    }
    float p = 4.2163199048e-2f;
    p = p * z + 2.4181311049e-2f;
    p = p * z + 4.5470025998e-2f;
    p = p * z + 7.4953002686e-2f;
    p = p * z + 1.6666752422e-1f;
    float r = p * z * x + x;
    if (big)
    {
        r = r + r;
        r = DM_PI_2 - r;
    }
    return (xx < 0.0f) ? -r : r;
}

DM_HD float dm_acos(float x)
{
    if (dm_isnan(x))
        return x;
    if (x < -1.0f || x > 1.0f)
        return dm_nan();
    if (x < -0.5f)
        return DM_PI - 2.0f * dm_asin(dm_sqrt(0.5f * (1.0f + x)));
    if (x > 0.5f)
        return 2.0f * dm_asin(dm_sqrt(0.5f * (1.0f - x)));
    return DM_PI_2 - dm_asin(x);
}

// cube root for x >= 0 (util/green.h:104 uses cbrt(u1*u1)): exp(log(x)/3) polished by two Newton steps
DM_HD float dm_cbrt(float x)
{
    if (!(x > 0.0f))
        return x; // 0, -0, NaN pass through; negatives are not used by the walker
    float y = dm_exp(dm_log(x) * 0.333333333333333333f);
    y = y - dm_div(y * y * y - x, 3.0f * y * y);
    y = y - dm_div(y * y * y - x, 3.0f * y * y);
    return y;
}

// logistic (core/math/include/krrmath/functors.h:182)
DM_HD float dm_logistic(float x) { return dm_div(1.0f, 1.0f + dm_exp(-x)); }
