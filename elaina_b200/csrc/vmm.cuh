// Guiding distribution: mixture of 8 von Mises lobes (2D) / von Mises-Fisher lobes (3D), decoded from the raw
// network output, with pdf, sampling and the gradient of the mixture pdf w.r.t. the activated lobe parameters.
//
// Follows integrator/guided/distribution.h:136-444 (VMM<2|3,8>), util/vonmises.h:64-209, util/vmf.h:21-84 and the
// output activations of integrator/guided/train.h:49-105. Differences, all stated here:
//   * fp32 transcendentals: on the device the hardware approximations (ex2 / lg2 / rcp based __expf, __logf, __fdividef,
//     rsqrtf) -- what the reference's own --use_fast_math build executes (CMakeLists.txt:64-70); host compiles of this header
//     use detmath.h. The guided layer is compared statistically and at stated tolerances, never bit for bit;
//   * the von Mises density is evaluated as exp(kappa (cos - 1) + [kappa - log I0(kappa)] - log 2 pi) with the bracket formed
//     without cancellation (0.5 log kappa - log(poly) for kappa >= 3.75): the reference's kappa cos - log 2 pi - log I0(kappa)
//     loses |kappa| ulps to cancellation (kappa reaches e^15); same value up to that rounding noise, and the per-lobe
//     normalisers are computed once per decode instead of once per density evaluation;
//   * the Best-Fisher rejection sampler of the von Mises lobe keeps the reference's double arithmetic for the proposal
//     parameter r = 1 + O(1/kappa), the candidate f and the acceptance variable c (util/vonmises.h:95-118: not representable in
//     fp32 for the kappa <= e^15 the activation allows); cos(pi u), the logarithm of the acceptance test and the final arc
//     cosine (written 2 asin(sqrt((1 -+ f) / 2)), exact near f = +-1) run in fp32 -- they only perturb the proposal draw / the
//     acceptance boundary at the 1e-7 level. Numbers come from the counter-based Philox stream with the rejection iteration as a
//     counter word; the loop gives up after 64 iterations (the reference loops forever);
//   * d log I0 / d kappa is written as the Horner derivative of the same polynomials instead of the reference's
//     expanded expression (util/vonmises.h:135-170) -- algebraically identical.
#pragma once
#include "geom.cuh"
#include "philox.h"

namespace wost
{
    constexpr int kLobes = 8;                       // parameters.h:22,33
    constexpr float kMEps = 1e-5f;                  // M_EPSILON, krrmath/constants.h:19
    constexpr float kInv2Pi = 0.15915494309189535f; // 1/(2 pi)
    constexpr float kInv4Pi = 0.07957747154594767f;
    template <int D>
    struct GuideDims
    {
        static constexpr int kLobeDims = D + 2;                   // lambda, kappa, mu (N_DIM_VMF: 4 / 5)
        static constexpr int kOut = kLobes * kLobeDims + 1;       // + selection probability (33 / 41)
        static constexpr int kOutPadded = 48;                     // parameters.h:25,36
    };

    // ---- transcendental / division flavour of the guided layer (see the header comment)
    WOST_HOSTDEV float g_exp(float x)
    {
#if defined(__CUDA_ARCH__)
        return __expf(x);
#else
        return dm_exp(x);
#endif
    }
    WOST_HOSTDEV float g_log(float x)
    {
#if defined(__CUDA_ARCH__)
        return __logf(x);
#else
        return dm_log(x);
#endif
    }
    WOST_HOSTDEV float g_div(float a, float b)
    {
#if defined(__CUDA_ARCH__)
        return __fdividef(a, b);
#else
        return dm_div(a, b);
#endif
    }
    WOST_HOSTDEV float g_logistic(float x) { return g_div(1.0f, 1.0f + g_exp(-x)); }
    WOST_HOSTDEV float g_rsqrt(float x)
    {
#if defined(__CUDA_ARCH__)
        return rsqrtf(x);
#else
        return dm_div(1.0f, dm_sqrt(x));
#endif
    }

    // ---- Abramowitz-Stegun polynomials of log I0 (util/vonmises.h:18-92)
    WOST_HOSTDEV float i0_small_poly(float y, float *dy)
    {
        const float c[7] = {1.0f, 3.5156229f, 3.0899424f, 1.2067492f, 0.2659732f, 0.360768e-1f, 0.45813e-2f};
        float p = c[6], d = 0.0f;
#pragma unroll
        for (int i = 5; i >= 0; --i)
        {
            d = p + y * d;
            p = c[i] + y * p;
        }
        if (dy)
            *dy = d;
        return p;
    }
    WOST_HOSTDEV float i0_large_poly(float y, float *dy)
    {
        const float c[9] = {0.39894228f, 0.1328592e-1f, 0.225319e-2f, -0.157565e-2f, 0.916281e-2f, -0.2057706e-1f, 0.2635537e-1f, -0.1647633e-1f, 0.392377e-2f};
        float p = c[8], d = 0.0f;
#pragma unroll
        for (int i = 7; i >= 0; --i)
        {
            d = p + y * d;
            p = c[i] + y * p;
        }
        if (dy)
            *dy = d;
        return p;
    }
    // logModifiedBesselFn(x, 0), util/vonmises.h:75-92
    WOST_HOSTDEV float log_bessel_i0(float x)
    {
        if (x < 3.75f)
        {
            float y = g_div(x, 3.75f);
            y *= y;
            return g_log(i0_small_poly(y, nullptr));
        }
        const float y = g_div(3.75f, x);
        return x - 0.5f * g_log(x) + g_log(i0_large_poly(y, nullptr));
    }
    // kappa - log I0(kappa) - log(2 pi): the log-normaliser of the von Mises density written around cos - 1, no cancellation
    WOST_HOSTDEV float vm_lognorm(float x)
    {
        if (x < 3.75f)
        {
            float y = x * (1.0f / 3.75f);
            y *= y;
            return x - g_log(i0_small_poly(y, nullptr)) - 1.8378770664093453f;
        }
        const float y = g_div(3.75f, x);
        return 0.5f * g_log(x) - g_log(i0_large_poly(y, nullptr)) - 1.8378770664093453f;
    }
    // d/dkappa log I0(kappa)
    WOST_HOSTDEV float dlog_bessel_i0(float x)
    {
        if (x < 3.75f)
        {
            const float t = g_div(x, 3.75f);
            float d;
            const float p = i0_small_poly(t * t, &d);
            return g_div(d, p) * g_div(2.0f * x, 3.75f * 3.75f);
        }
        const float y = g_div(3.75f, x);
        float d;
        const float p = i0_large_poly(y, &d);
        return 1.0f - g_div(0.5f, x) - g_div(d, p) * g_div(y, x);
    }
    // VonMises::eval / d_eval_d_kappa (util/vonmises.h:128-183)
    WOST_HOSTDEV float vm_eval_ln(float kappa, float lognorm, float cosTheta)
    {
        return (kappa < 1e-3f) ? kInv2Pi : g_exp(kappa * (cosTheta - 1.0f) + lognorm);
    }
    WOST_HOSTDEV float vm_eval(float kappa, float cosTheta)
    {
        if (kappa < 1e-3f)
            return kInv2Pi;
        return vm_eval_ln(kappa, vm_lognorm(kappa), cosTheta);
    }
    WOST_HOSTDEV float vm_deval_dkappa(float kappa, float cosTheta)
    {
        return vm_eval(kappa, cosTheta) * (cosTheta - dlog_bessel_i0(kappa));
    }
    // VMF::eval (util/vmf.h:27-32): exp(kappa min(0, cos - 1)) * kappa / (2 pi (1 - exp(-2 kappa)))
    WOST_HOSTDEV float vmf_norm(float kappa) { return g_div(kappa, DM_2PI * (1.0f - g_exp(-2.0f * kappa))); }
    WOST_HOSTDEV float vmf_eval_n(float kappa, float norm, float cosTheta)
    {
        return (kappa < kMEps) ? kInv4Pi : g_exp(kappa * fmin_nn(0.0f, cosTheta - 1.0f)) * norm;
    }
    WOST_HOSTDEV float vmf_eval(float kappa, float cosTheta)
    {
        if (kappa < kMEps)
            return kInv4Pi;
        return vmf_eval_n(kappa, vmf_norm(kappa), cosTheta);
    }

    // ---- activations (train.h:49-105)
    WOST_HOSTDEV float act_exp(float v) { return g_exp(fmin_nn(fmax_nn(v, -10.0f), 15.0f)); }

    template <int D>
    struct VMM
    {
        float lambda[kLobes], kappa[kLobes], weight[kLobes];
        float norm[kLobes]; // 2D: log-normaliser kappa - log I0(kappa) - log 2 pi; 3D: kappa / (2 pi (1 - exp(-2 kappa)))
        Vec<D> mu[kLobes], mu_orig[kLobes];
        float total;

        // VMM(T *data), distribution.h:146-168 / 289-313
        WOST_HOSTDEV void decode(const float *raw)
        {
            total = 0.0f;
#pragma unroll
            for (int i = 0; i < kLobes; ++i)
            {
                const float *p = raw + i * (D + 2);
                lambda[i] = act_exp(p[0]);
                kappa[i] = act_exp(p[1]);
                norm[i] = (D == 2) ? vm_lognorm(kappa[i]) : vmf_norm(kappa[i]);
                float n2 = 0.0f;
#pragma unroll
                for (int k = 0; k < D; ++k)
                {
                    mu_orig[i][k] = p[2 + k];
                    n2 += p[2 + k] * p[2 + k];
                }
                const float inv = (n2 > 0.0f) ? g_rsqrt(n2) : 1.0f; // Eigen normalized(): x / norm when norm > 0
#pragma unroll
                for (int k = 0; k < D; ++k)
                    mu[i][k] = mu_orig[i][k] * inv;
                total += lambda[i];
            }
            const float inv_total = g_div(1.0f, total);
#pragma unroll
            for (int i = 0; i < kLobes; ++i)
                weight[i] = lambda[i] * inv_total;
        }
        WOST_HOSTDEV float lobe_pdf(int i, const Vec<D> &w) const
        {
            const float c = vdot(w, mu[i]);
            return D == 2 ? vm_eval_ln(kappa[i], norm[i], c) : vmf_eval_n(kappa[i], norm[i], c);
        }
        WOST_HOSTDEV float pdf(const Vec<D> &w) const
        {
            float v = 0.0f;
#pragma unroll
            for (int i = 0; i < kLobes; ++i)
                v += weight[i] * lobe_pdf(i, w);
            return v;
        }
        // lobe choice of VMM::sample (distribution.h:185-197)
        WOST_HOSTDEV int pick(float u) const
        {
#pragma unroll
            for (int i = 0; i < kLobes; ++i)
            {
                if (u < weight[i])
                    return i;
                u -= weight[i];
            }
            return 0;
        }
    };

    // reflect (util/transformation.h:69-72)
    template <int D>
    WOST_HOSTDEV Vec<D> vreflect(const Vec<D> &d, const Vec<D> &n)
    {
        const float s = 2.0f * vdot(d, n);
        Vec<D> r;
#pragma unroll
        for (int k = 0; k < D; ++k)
            r[k] = d[k] - s * n[k];
        return r;
    }

    // ---- lobe sampling
    WOST_HOSTDEV double u64_to_unit_double(uint32_t hi, uint32_t lo)
    {
        const unsigned long long b = ((unsigned long long)hi << 32) | lo;
        return (double)(b >> 11) * (1.0 / 9007199254740992.0);
    }
    // rejectionSample + VonMises::sample(sampler, mu) (util/vonmises.h:95-118,185-194)
    WOST_HOSTDEV Vec<2> vm_sample(float kappa, const Vec<2> &mu, uint32_t seed, uint32_t pixel, uint32_t sample, uint32_t depth)
    {
        float theta;
        if (kappa < 1e-3f)
        {
            const Philox4 q = wost_rng(seed, pixel, sample, depth, WOST_RNG_GUIDED, 0);
            theta = DM_2PI * u32_to_unit_float(q.x);
        }
        else
        {
            const double k = (double)kappa;
            const double tau = 1.0 + sqrt(1.0 + 4.0 * k * k);
            const double rho = (tau - sqrt(2.0 * tau)) / (2.0 * k);
            const double pr = (kappa < 1e-5f) ? (1.0 / k + k) : (1.0 + rho * rho) / (2.0 * rho);
            double f = 1.0, sgn = 1.0;
            for (uint32_t it = 0; it < 64u; ++it)
            {
                const Philox4 q = wost_rng(seed, pixel, sample, depth, WOST_RNG_GUIDED, it);
                const double u1 = u64_to_unit_double(q.x, q.y), u2 = u64_to_unit_double(q.z, q.w);
                sgn = (q.w & 1u) ? 1.0 : -1.0;
#if defined(__CUDA_ARCH__)
                const double z = (double)cospif((float)u1);
#else
                const double z = cos(3.14159265358979323846 * u1);
#endif
                f = (1.0 + pr * z) / (pr + z);
                const double c = k * (pr - f);
                if ((c * (2.0 - c) - u2) > 0.0)
                    break;
#if defined(__CUDA_ARCH__)
                if (__logf((float)c) - __logf((float)u2) + 1.0f - (float)c >= 0.0f)
                    break;
#else
                if (log(c / u2) + 1.0 - c >= 0.0)
                    break;
#endif
            }
            f = f > 1.0 ? 1.0 : (f < -1.0 ? -1.0 : f);
#if defined(__CUDA_ARCH__)
            // acos(f) = 2 asin(sqrt((1 - f) / 2)) for f >= 0, pi - 2 asin(sqrt((1 + f) / 2)) below: 1 -+ f is formed in double
            const float half = (float)((f >= 0.0 ? 1.0 - f : 1.0 + f) * 0.5);
            const float a = 2.0f * asinf(sqrtf(half));
            theta = (float)sgn * (f >= 0.0 ? a : 3.14159265358979323846f - a);
#else
            theta = (float)(sgn * acos(f));
#endif
        }
        float s, c;
        dm_sincos(theta, &s, &c);
        // frameFromTangent(mu).toWorld((cos, sin)): T = mu, N = perp(mu) (util/transformation.h:36-50)
        Vec<2> r;
        r[0] = mu[0] * c + -mu[1] * s;
        r[1] = mu[1] * c + mu[0] * s;
        return r;
    }
    WOST_HOSTDEV Vec<3> frame_to_world_3d(const Vec<3> &n, const Vec<3> &v) // frameFromNormal(n).toWorld(v), transformation.h:62-67 + math_utils.h:141-151
    {
        const float ax = dm_abs(n[0]), ay = dm_abs(n[1]), az = dm_abs(n[2]);
        const uint32_t uyx = (ax - ay) < 0 ? 1 : 0, uzx = (ax - az) < 0 ? 1 : 0, uzy = (ay - az) < 0 ? 1 : 0;
        const uint32_t xm = uyx & uzx, ym = (1 ^ xm) & uzy, zm = 1 ^ (xm | ym);
        Vec<3> a;
        a[0] = (float)xm, a[1] = (float)ym, a[2] = (float)zm;
        Vec<3> t = vcross(n, a);
        float l = vlen(t);
        if (l > 0.0f)
            t[0] = g_div(t[0], l), t[1] = g_div(t[1], l), t[2] = g_div(t[2], l);
        Vec<3> b = vcross(n, t);
        l = vlen(b);
        if (l > 0.0f)
            b[0] = g_div(b[0], l), b[1] = g_div(b[1], l), b[2] = g_div(b[2], l);
        Vec<3> r;
#pragma unroll
        for (int k = 0; k < 3; ++k)
            r[k] = t[k] * v[0] + b[k] * v[1] + n[k] * v[2];
        return r;
    }
    // VMF::sample (util/vmf.h:45-56,70-72)
    WOST_HOSTDEV Vec<3> vmf_sample(float kappa, const Vec<3> &mu, uint32_t seed, uint32_t pixel, uint32_t sample, uint32_t depth)
    {
        const Philox4 q = wost_rng(seed, pixel, sample, depth, WOST_RNG_GUIDED, 0);
        const float u0 = u32_to_unit_float(q.x), u1 = u32_to_unit_float(q.y);
        Vec<3> l;
        if (kappa < kMEps)
        {
            const float z = 1 - 2 * u0;
            const float rr = dm_sqrt(fmax_nn(0.0f, 1 - z * z));
            float s, c;
            dm_sincos(DM_2PI * u1, &s, &c);
            l[0] = rr * c, l[1] = rr * s, l[2] = z;
            return l; // uniformSampleSphere: not rotated (vmf.h:46-47)
        }
        // cosTheta = 1 + log1p(-u + exp(-2k) u) / k ; log1p(x) evaluated as log(1 + x) on the detmath log
        const float cosT = 1.0f + g_div(g_log(1.0f + (-u0 + g_exp(-2.0f * kappa) * u0)), kappa);
        const float sinT = dm_sqrt(fmax_nn(0.0f, 1.0f - cosT * cosT));
        float s, c;
        dm_sincos(DM_2PI * u1, &s, &c);
        l[0] = c * sinT, l[1] = s * sinT, l[2] = cosT;
        return frame_to_world_3d(mu, l);
    }
    WOST_HOSTDEV Vec<2> lobe_sample(float kappa, const Vec<2> &mu, uint32_t seed, uint32_t pixel, uint32_t sample, uint32_t depth)
    {
        return vm_sample(kappa, mu, seed, pixel, sample, depth);
    }
    WOST_HOSTDEV Vec<3> lobe_sample(float kappa, const Vec<3> &mu, uint32_t seed, uint32_t pixel, uint32_t sample, uint32_t depth)
    {
        return vmf_sample(kappa, mu, seed, pixel, sample, depth);
    }

    // ---- gradient of the mixture pdf w.r.t. the ACTIVATED lobe parameters (VMM::gradients_probability,
    // distribution.h:200-264 / 347-431). grad: kLobes * (D+2) floats (lambda, kappa, mu...). Returns the pdf
    // (including the reflected direction when the sample sits on a Neumann boundary).
    template <int D>
    WOST_HOSTDEV float vmm_grad_prob(const VMM<D> &m, const Vec<D> &wi, bool onN, const Vec<D> &nN, float *grad)
    {
        Vec<D> wr = wi;
        if (onN)
            wr = vreflect(wi, nN);
        float p[kLobes], pr[kLobes];
        float prob = 0.0f;
#pragma unroll
        for (int i = 0; i < kLobes; ++i)
        {
            p[i] = m.lobe_pdf(i, wi);
            pr[i] = onN ? m.lobe_pdf(i, wr) : 0.0f;
            prob += m.weight[i] * p[i];
            if (onN)
                prob += m.weight[i] * pr[i];
        }
#pragma unroll
        for (int sg = 0; sg < kLobes; ++sg)
        {
            float *g = grad + sg * (D + 2);
            const float lambda = m.lambda[sg], kappa = m.kappa[sg], w = m.weight[sg];
            float dl = g_div((p[sg] + pr[sg]) * (m.total - lambda), m.total * m.total);
#pragma unroll
            for (int k = 0; k < kLobes; ++k)
                if (k != sg)
                {
                    dl -= g_div(m.weight[k], m.total) * p[k];
                    if (onN)
                        dl -= g_div(m.weight[k], m.total) * pr[k];
                }
            g[0] = dl;
            float n2 = 0.0f;
#pragma unroll
            for (int k = 0; k < D; ++k)
                n2 += m.mu_orig[sg][k] * m.mu_orig[sg][k];
            float denom = n2 * dm_sqrt(n2); // pow(n2, 1.5)
            if (denom < kMEps)
                denom = kMEps;
            if (D == 2)
            {
                float dk = w * vm_deval_dkappa(kappa, vdot(wi, m.mu[sg]));
                if (onN)
                    dk += w * vm_deval_dkappa(kappa, vdot(wr, m.mu[sg]));
                g[1] = dk;
                const float mx = m.mu_orig[sg][0], my = m.mu_orig[sg][1];
                float dx = g_div(w * p[sg] * kappa * my * (-mx * wi[1] + my * wi[0]), denom);
                float dy = g_div(w * p[sg] * kappa * mx * (mx * wi[1] - my * wi[0]), denom);
                if (onN)
                {
                    dx += g_div(w * pr[sg] * kappa * my * (-mx * wr[1] + my * wr[0]), denom);
                    dy += g_div(w * pr[sg] * kappa * mx * (mx * wr[1] - my * wr[0]), denom);
                }
                g[2] = dx;
                g[3] = dy;
            }
            else
            {
                float ik; // 1/kappa - coth(kappa)
                if (kappa < 1.0f)
                    ik = 0.000962f + -0.344883f * kappa + 0.030147f * kappa * kappa;
                else
                {
                    const float e = g_exp(-2.0f * kappa);
                    ik = g_div(1.0f, kappa) - g_div(1.0f + e, 1.0f - e);
                }
                float dk = w * p[sg] * (vdot(wi, m.mu[sg]) + ik);
                if (onN)
                    dk += w * pr[sg] * (vdot(wr, m.mu[sg]) + ik);
                g[1] = dk;
                const float mx = m.mu_orig[sg][0], my = m.mu_orig[sg][1], mz = m.mu_orig[sg][D - 1];
                // d(mu_hat . w)/d(mu_orig) = (|m|^2 w - (m . w) m) / |m|^3, written out like the reference
#pragma unroll
                for (int pass = 0; pass < 2; ++pass)
                {
                    if (pass == 1 && !onN)
                        break;
                    const Vec<D> &v = pass == 0 ? wi : wr;
                    const float pv = pass == 0 ? p[sg] : pr[sg];
                    const float x = v[0], y = v[1], z = v[D - 1];
                    const float gx = g_div(w * pv * kappa * (-mx * my * y - mx * mz * z + my * my * x + mz * mz * x), denom);
                    const float gy = g_div(w * pv * kappa * (-mx * my * x - my * mz * z + mx * mx * y + mz * mz * y), denom);
                    const float gz = g_div(w * pv * kappa * (-mx * mz * x - my * mz * y + mx * mx * z + my * my * z), denom);
                    if (pass == 0)
                        g[2] = gx, g[3] = gy, g[D + 1] = gz;
                    else
                        g[2] += gx, g[3] += gy, g[D + 1] += gz;
                }
            }
        }
        return prob;
    }

    // One training sample (GuidedOutput, guideditem.h:44-51): direction taken, its effective pdf, target radiance.
    template <int D>
    struct TrainSample
    {
        Vec<D> dir, nN;
        float dirPdf, Li;
        bool onN;
    };
    // compute_dL_doutput_divergence, KL branch (train.h:491-553). raw: the kOut raw network outputs of the row;
    // dl: kOut gradients w.r.t. the raw outputs (already multiplied by loss_scale / batch). Returns the per-sample loss.
    template <int D>
    WOST_HOSTDEV float kl_dl_doutput(const float *raw, const TrainSample<D> &t, float loss_scale_over_n, float *dl)
    {
        VMM<D> m;
        m.decode(raw);
        const float dirPdf = t.dirPdf + kMEps;
        const float guidePdf = vmm_grad_prob(m, t.dir, t.onN, t.nN, dl) + kMEps;
        const float prefix = g_div(g_div(-t.Li, dirPdf), guidePdf) * loss_scale_over_n;
#pragma unroll
        for (int sg = 0; sg < kLobes; ++sg)
        {
            float *g = dl + sg * (D + 2);
            const float *r = raw + sg * (D + 2);
            // exponential activation: derivative == activated value (also where the clamp saturates, train.h:96-97)
            g[0] = prefix * g[0] * act_exp(r[0]);
            g[1] = prefix * g[1] * act_exp(r[1]);
#pragma unroll
            for (int k = 0; k < D; ++k)
                g[2 + k] = prefix * g[2 + k];
        }
        const float sp = g_logistic(raw[kLobes * (D + 2)]);
        const float up = t.onN ? (D == 2 ? 1.0f / DM_PI : 1.0f / DM_2PI) : (D == 2 ? 1.0f / DM_2PI : 1.0f / DM_4PI);
        dl[kLobes * (D + 2)] = g_div(loss_scale_over_n * -0.2f * t.Li * (guidePdf - up), dirPdf * dirPdf) * (sp * (1.0f - sp));
        return g_div(-t.Li, dirPdf) * g_log(guidePdf);
    }
} // namespace wost
