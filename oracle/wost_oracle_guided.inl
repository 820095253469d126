// TEST INFRASTRUCTURE ONLY -- included by wost_oracle.cpp after wost_oracle_walk.inl.
// Scalar CPU restatement of the guided integrator's arithmetic:
//   * von Mises / vMF lobes and the 8-lobe mixture (util/vonmises.h:64-209, util/vmf.h:21-84,
//     integrator/guided/distribution.h:136-444), output activations (integrator/guided/train.h:49-105)
//   * KL-divergence gradient w.r.t. the raw network outputs (train.h:491-553)
//   * DenseGrid encoding + ReLU MLP forward (tcnn encodings/grid.h:171-350, fully_fused_mlp.cu:499-558; fp16 operands,
//     wide accumulation)
//   * the guided walk itself (integrator/guided/integrator.cu:153-965) with fixed network weights
// Written independently of elaina_b200/csrc/vmm.cuh (different structure: per-lobe structs, explicit reference
// formulas incl. the expanded d log I0 / d kappa expressions) so that agreement is evidence, not identity.
// Pinned by the reference's own known answers (test/vonmises_test.cu:7,13-14,57-59,133,146) in tests/test_oracle_guided.py.
// PARITY UNPINNED for this layer (guided walker / network): the reference ships no golden fields and its integrators cannot run without a GPU
// build. What IS pinned: the LBVH layer underneath (bit-exact against the host-compiled reference headers,
// oracle/_ref) and the special functions (known answers of test/vonmises_test.cu); fields are checked against analytic solutions.

namespace
{
    inline float m_exp(float x) { return g_math_mode ? expf(x) : dm_exp(x); }
    const float G_EPS = 1e-5f; // M_EPSILON (krrmath/constants.h:19)

    const float COEF_S0[7] = {1.0f, 3.5156229f, 3.0899424f, 1.2067492f, 0.2659732f, 0.360768e-1f, 0.45813e-2f};
    const float COEF_S1[7] = {0.5f, 0.87890594f, 0.51498869f, 0.15084934f, 0.2658733e-1f, 0.301532e-2f, 0.32411e-3f};
    const float COEF_L0[9] = {0.39894228f, 0.1328592e-1f, 0.225319e-2f, -0.157565e-2f, 0.916281e-2f, -0.2057706e-1f, 0.2635537e-1f, -0.1647633e-1f, 0.392377e-2f};
    const float COEF_L1[9] = {0.39894228f, -0.3988024e-1f, -0.362018e-2f, 0.163801e-2f, -0.1031555e-1f, 0.2282967e-1f, -0.2895312e-1f, 0.1787654e-1f, -0.420059e-2f};

    // evalPoly (util/vonmises.h:64-73)
    float g_eval_poly(float y, const float *coeff, int n)
    {
        float ret = coeff[n - 1];
        for (int i = n - 2; i >= 0; --i)
            ret = coeff[i] + y * ret;
        return ret;
    }
    // logModifiedBesselFn (util/vonmises.h:75-92)
    float g_log_bessel(float x, int order)
    {
        float y = x / 3.75f;
        y *= y;
        float small = g_eval_poly(y, order ? COEF_S1 : COEF_S0, 7);
        if (order == 1)
            small = fabsf(x) * small;
        small = m_log(small);
        y = 3.75f / x;
        const float large = x - 0.5f * m_log(x) + m_log(g_eval_poly(y, order ? COEF_L1 : COEF_L0, 9));
        return (x < 3.75f) ? small : large;
    }
    struct GVonMises // util/vonmises.h:120-209
    {
        float kappa;
        float log_eval(float c) const { return kappa * c - m_log(DM_2PI) - g_log_bessel(kappa, 0); }
        float eval(float c) const { return kappa < 1e-3f ? 1.0f / DM_2PI : m_exp(log_eval(c)); }
        float d_log_eval_d_kappa(float cosTheta) const
        {
            if (kappa < 3.75f)
            {
                const float *coeff = COEF_S0;
                const float coef = 0.0711111111111111f, c142 = 0.142222222222222f, c010 = 0.0101135802469136f;
                const float kappa2 = kappa * kappa;
                const float term7 = coeff[6] * kappa2;
                const float term6 = coeff[5] + coef * term7;
                const float term5 = coeff[4] + coef * kappa2 * term6;
                const float term4 = coeff[3] + coef * kappa2 * term5;
                const float term3 = coeff[2] + coef * kappa2 * term4;
                const float term2 = coeff[1] + coef * kappa2 * term3;
                const float numerator = coef * kappa2 * (coef * kappa2 * (coef * kappa2 * (coef * kappa2 * (c010 * coeff[6] * kappa * kappa2 + c142 * kappa * term6) + c142 * kappa * term5) + c142 * kappa * term4) + c142 * kappa * term3) + c142 * kappa * term2;
                const float denominator = coeff[0] + coef * kappa2 * term2;
                return cosTheta - numerator / denominator;
            }
            // large branch: d/dx [x - log(x)/2 + log P(3.75/x)], P' by the nested sums of the reference's expression
            const float *K = COEF_L0;
            const double x = kappa;
            double P = 0.0, dP = 0.0; // P(y) = sum K_i y^i ; dP/dx = sum i K_i y^(i-1) * (-3.75/x^2)
            const double y = 3.75 / x;
            double yp = 1.0;
            for (int i = 0; i < 9; ++i)
            {
                P += K[i] * yp;
                if (i + 1 < 9)
                    dP += (double)(i + 1) * K[i + 1] * yp;
                yp *= y;
            }
            dP *= -3.75 / (x * x);
            return (float)(cosTheta - 1.0 - dP / P + 0.5 / x);
        }
        float d_eval_d_kappa(float c) const { return eval(c) * d_log_eval_d_kappa(c); }
    };
    // VMF::eval (util/vmf.h:27-32)
    float g_vmf_eval(float kappa, float c)
    {
        if (kappa < G_EPS)
            return 1.0f / DM_4PI;
        return m_exp(kappa * std::min(0.0f, c - 1.0f)) * kappa / (DM_2PI * (1.0f - m_exp(-2.0f * kappa)));
    }
    float g_act_exp(float v) { return m_exp(std::min(std::max(v, -10.0f), 15.0f)); } // network_to_params, train.h:60-79
    float g_logistic(float v) { return 1.0f / (1.0f + m_exp(-v)); }                  // krrmath functors.h:182

    template <int D>
    struct GLobe
    {
        float lambda, kappa;
        Vec<D> mu, mu_original;
        float pdf(const Vec<D> &w) const
        {
            const float c = dot(w, mu);
            if (D == 2)
                return GVonMises{kappa}.eval(c);
            return g_vmf_eval(kappa, c);
        }
    };
    template <int D>
    struct GMixture // VMM<DIM, 8>
    {
        GLobe<D> sg[8];
        float weight[8], totalWeight;
        explicit GMixture(const float *data)
        {
            totalWeight = 0.0f;
            for (int i = 0; i < 8; ++i)
            {
                const float *p = data + i * (D + 2);
                sg[i].lambda = g_act_exp(p[0]);
                sg[i].kappa = g_act_exp(p[1]);
                float n2 = 0.0f;
                for (int k = 0; k < D; ++k)
                {
                    sg[i].mu_original[k] = p[2 + k];
                    n2 += p[2 + k] * p[2 + k];
                }
                const float n = sqrtf(n2);
                for (int k = 0; k < D; ++k)
                    sg[i].mu[k] = n > 0.0f ? p[2 + k] / n : p[2 + k];
                totalWeight += sg[i].lambda;
            }
            for (int i = 0; i < 8; ++i)
                weight[i] = sg[i].lambda / totalWeight;
        }
        float pdf(const Vec<D> &w) const
        {
            float v = 0.0f;
            for (int i = 0; i < 8; ++i)
                v += weight[i] * sg[i].pdf(w);
            return v;
        }
        int pick(float u) const
        {
            for (int i = 0; i < 8; ++i)
            {
                if (u < weight[i])
                    return i;
                u -= weight[i];
            }
            return 0;
        }
    };
    template <int D>
    Vec<D> g_reflect(const Vec<D> &d, const Vec<D> &n)
    {
        const float s = 2.0f * dot(d, n);
        Vec<D> r;
        for (int k = 0; k < D; ++k)
            r[k] = d[k] - s * n[k];
        return r;
    }
    // VMM::gradients_probability (distribution.h:200-264 / 347-431)
    template <int D>
    float g_grad_prob(const GMixture<D> &m, const Vec<D> &wi, bool onN, const Vec<D> &nN, float *out)
    {
        float probability = 0.0f;
        Vec<D> wr = wi;
        if (onN)
            wr = g_reflect(wi, nN);
        for (int sg = 0; sg < 8; ++sg)
        {
            float *g = out + sg * (D + 2);
            const GLobe<D> &L = m.sg[sg];
            const float vm = L.pdf(wi);
            probability += m.weight[sg] * vm;
            float vmR = 0.0f;
            if (onN)
            {
                vmR = L.pdf(wr);
                probability += m.weight[sg] * vmR;
            }
            float dl = (vm + vmR) * (m.totalWeight - L.lambda) / (m.totalWeight * m.totalWeight);
            for (int k = 0; k < 8; ++k)
                if (k != sg)
                {
                    dl -= m.weight[k] / m.totalWeight * m.sg[k].pdf(wi);
                    if (onN)
                        dl -= m.weight[k] / m.totalWeight * m.sg[k].pdf(wr);
                }
            g[0] = dl;
            float n2 = 0.0f;
            for (int k = 0; k < D; ++k)
                n2 += L.mu_original[k] * L.mu_original[k];
            float denom = powf(n2, 1.5f);
            if (denom < G_EPS)
                denom = G_EPS;
            if (D == 2)
            {
                float dk = m.weight[sg] * GVonMises{L.kappa}.d_eval_d_kappa(dot(wi, L.mu));
                if (onN)
                    dk += m.weight[sg] * GVonMises{L.kappa}.d_eval_d_kappa(dot(wr, L.mu));
                g[1] = dk;
                const float mx = L.mu_original[0], my = L.mu_original[1];
                float dx = m.weight[sg] * vm * L.kappa * my * (-mx * wi[1] + my * wi[0]) / denom;
                float dy = m.weight[sg] * vm * L.kappa * mx * (mx * wi[1] - my * wi[0]) / denom;
                if (onN)
                {
                    dx += m.weight[sg] * vmR * L.kappa * my * (-mx * wr[1] + my * wr[0]) / denom;
                    dy += m.weight[sg] * vmR * L.kappa * mx * (mx * wr[1] - my * wr[0]) / denom;
                }
                g[2] = dx;
                g[3] = dy;
            }
            else
            {
                const float kappa = L.kappa;
                float ik;
                if (kappa < 1)
                    ik = 0.000962f + -0.344883f * kappa + 0.030147f * kappa * kappa;
                else
                    ik = 1 / kappa - (1 + m_exp(-2 * kappa)) / (1 - m_exp(-2 * kappa));
                float dk = m.weight[sg] * vm * (dot(wi, L.mu) + ik);
                if (onN)
                    dk += m.weight[sg] * vmR * (dot(wr, L.mu) + ik);
                g[1] = dk;
                const float mo[3] = {L.mu_original[0], L.mu_original[1], L.mu_original[D - 1]};
                for (int pass = 0; pass < (onN ? 2 : 1); ++pass)
                {
                    const Vec<D> &v = pass ? wr : wi;
                    const float pv = pass ? vmR : vm;
                    const float w3[3] = {v[0], v[1], v[D - 1]};
                    // d (mu_hat . w) / d mu_original_j = (|m|^2 w_j - (m . w) m_j) / |m|^3
                    const float mw = mo[0] * w3[0] + mo[1] * w3[1] + mo[2] * w3[2];
                    for (int j = 0; j < 3; ++j)
                    {
                        const float t = m.weight[sg] * pv * kappa * ((n2 - mo[j] * mo[j]) * w3[j] - (mw - mo[j] * w3[j]) * mo[j]) / denom;
                        if (pass == 0)
                            g[2 + j] = t;
                        else
                            g[2 + j] += t;
                    }
                }
            }
        }
        return probability;
    }
    // compute_dL_doutput_divergence, KL (train.h:491-553). scale = loss_scale / nElements.
    template <int D>
    float g_kl_grad(const float *raw, const Vec<D> &wi, const Vec<D> &nN, bool onN, float dirPdfIn, float Li, float scale, float *dl)
    {
        const GMixture<D> m(raw);
        const float dirPdf = dirPdfIn + G_EPS;
        const float guidePdf = g_grad_prob<D>(m, wi, onN, nN, dl) + G_EPS;
        const float prefix = -Li / dirPdf / guidePdf * scale;
        for (int sg = 0; sg < 8; ++sg)
        {
            float *g = dl + sg * (D + 2);
            const float *r = raw + sg * (D + 2);
            g[0] = prefix * g[0] * g_act_exp(r[0]);
            g[1] = prefix * g[1] * g_act_exp(r[1]);
            for (int k = 0; k < D; ++k)
                g[2 + k] = prefix * g[2 + k];
        }
        const float sp = g_logistic(raw[8 * (D + 2)]);
        float up = D == 2 ? 1.0f / DM_2PI : 1.0f / DM_4PI;
        if (onN)
            up = D == 2 ? 1.0f / DM_PI : 1.0f / DM_2PI;
        dl[8 * (D + 2)] = scale * (-0.2f) * Li * (guidePdf - up) / (dirPdf * dirPdf) * (sp * (1 - sp));
        return -Li / dirPdf * m_log(guidePdf);
    }

    // ---- lobe sampling with the product's Philox draws (elaina_b200/csrc/vmm.cuh documents the stream)
    inline double g_u53(uint32_t hi, uint32_t lo) { return (double)((((uint64_t)hi << 32) | lo) >> 11) * (1.0 / 9007199254740992.0); }
    Vec<2> g_sample_lobe(const GLobe<2> &L, uint32_t seed, uint32_t pixel, uint32_t sample, uint32_t depth)
    {
        float theta;
        if (L.kappa < 1e-3f)
            theta = DM_2PI * u32_to_unit_float(wost_rng(seed, pixel, sample, depth, WOST_RNG_GUIDED, 0).x);
        else
        {
            // rejectionSample (util/vonmises.h:95-118) with proposalR_compute (:196-203), in double like the reference
            const double k = L.kappa;
            const double tau = 1.0 + sqrt(1.0 + 4.0 * k * k);
            const double rho = (tau - sqrt(2.0 * tau)) / (2.0 * k);
            const double pr = (1.0 + rho * rho) / (2.0 * rho);
            double f = 1.0, sgn = 1.0;
            for (uint32_t it = 0; it < 64u; ++it)
            {
                const Philox4 q = wost_rng(seed, pixel, sample, depth, WOST_RNG_GUIDED, it);
                const double u1 = g_u53(q.x, q.y), u2 = g_u53(q.z, q.w);
                sgn = (q.w & 1u) ? 1.0 : -1.0;
                const double z = cos(M_PI * u1);
                f = (1.0 + pr * z) / (pr + z);
                const double c = k * (pr - f);
                if ((c * (2.0 - c) - u2) > 0.0 || (log(c / u2) + 1.0 - c >= 0.0))
                    break;
            }
            f = std::min(1.0, std::max(-1.0, f));
            theta = (float)(sgn * acos(f));
        }
        float s, c;
        dm_sincos(theta, &s, &c);
        return {{L.mu[0] * c + -L.mu[1] * s, L.mu[1] * c + L.mu[0] * s}};
    }
    Vec<3> g_sample_lobe(const GLobe<3> &L, uint32_t seed, uint32_t pixel, uint32_t sample, uint32_t depth)
    {
        const Philox4 q = wost_rng(seed, pixel, sample, depth, WOST_RNG_GUIDED, 0);
        const float u0 = u32_to_unit_float(q.x), u1 = u32_to_unit_float(q.y);
        float s, c;
        dm_sincos(DM_2PI * u1, &s, &c);
        if (L.kappa < G_EPS)
        {
            const float z = 1 - 2 * u0;
            const float rr = sqrtf(std::max(0.0f, 1 - z * z));
            return {{rr * c, rr * s, z}};
        }
        const float cosT = 1.0f + m_log(1.0f + (-u0 + m_exp(-2.0f * L.kappa) * u0)) / L.kappa;
        const float sinT = sqrtf(std::max(0.0f, 1.0f - cosT * cosT));
        return to_world(L.mu, Vec<3>{{c * sinT, s * sinT, cosT}});
    }
}

// ------------------------------------------------------------------------------------------------ network
namespace
{
    inline float g_half(float v) { return (float)(_Float16)v; } // round to nearest fp16, like the device operands

    struct OrcNet
    {
        int dim, n_levels, k0, base_res;
        float per_level_scale;
        std::vector<float> scale;
        std::vector<uint32_t> res, off; // off: n_levels + 1 (grid points)
        float lo[3], hi[3];
        int n_mlp, n_grid;
        std::vector<float> params; // fp32 slab [W0|W1|W2|W3|grid], W row-major [out][in]
        int layer_in(int l) const { return l == 0 ? k0 : 64; }
        int layer_out(int l) const { return l == 3 ? 48 : 64; }
        int layer_off(int l) const { return l == 0 ? 0 : 64 * k0 + (l - 1) * 4096; }
    };
    // normalizeSpatialCoord (train.h:148-155); lo/hi are the already inflated box
    template <int D>
    void g_normalize(const OrcNet &n, const Vec<D> &x, float *xn)
    {
        for (int k = 0; k < D; ++k)
            xn[k] = 0.5f + (x[k] - (n.lo[k] + n.hi[k]) * 0.5f) / (n.hi[k] - n.lo[k]);
    }
    // kernel_grid, Linear + Dense (grid.h:221-350): corner parameter indices and weights of one level
    template <int D>
    void g_cell(const OrcNet &n, int level, const float *xn, uint32_t *idx, float *w)
    {
        const float scale = n.scale[level];
        const uint32_t res = n.res[level], size = n.off[level + 1] - n.off[level];
        float fr[3];
        uint32_t pg[3];
        for (int d = 0; d < D; ++d)
        {
            const float p = xn[d] * scale + 0.5f;
            const float fl = floorf(p);
            pg[d] = (uint32_t)(int)fl;
            fr[d] = p - fl;
        }
        for (int corner = 0; corner < (1 << D); ++corner)
        {
            float ww = 1.0f;
            uint32_t stride = 1, index = 0;
            for (int d = 0; d < D; ++d)
            {
                const bool up = (corner >> d) & 1;
                ww *= up ? fr[d] : 1.0f - fr[d];
                if (stride <= size)
                {
                    index += (pg[d] + (up ? 1u : 0u)) * stride;
                    stride *= res;
                }
            }
            idx[corner] = n.off[level] + index % size;
            w[corner] = ww;
        }
    }
    template <int D>
    void g_encode(const OrcNet &n, const Vec<D> &x, float *enc)
    {
        float xn[3];
        g_normalize<D>(n, x, xn);
        const float *grid = n.params.data() + n.n_mlp;
        for (int l = 0; l < n.n_levels; ++l)
        {
            uint32_t idx[8];
            float w[8];
            g_cell<D>(n, l, xn, idx, w);
            for (int f = 0; f < 4; ++f)
            {
                float s = 0.0f;
                for (int c = 0; c < (1 << D); ++c)
                    s += w[c] * g_half(grid[(size_t)idx[c] * 4 + f]);
                enc[l * 4 + f] = s;
            }
        }
        for (int k = n.n_levels * 4; k < n.k0; ++k)
            enc[k] = 0.0f;
    }
    // acts: optional 4 * 64 floats receiving the fp16-rounded inputs of every layer
    void g_mlp_forward(const OrcNet &n, const float *enc, float *out, float *acts)
    {
        float a[64], b[64];
        for (int k = 0; k < n.k0; ++k)
            a[k] = g_half(enc[k]);
        for (int l = 0; l < 4; ++l)
        {
            const int K = n.layer_in(l), N = n.layer_out(l);
            const float *W = n.params.data() + n.layer_off(l);
            if (acts)
                for (int k = 0; k < K; ++k)
                    acts[l * 64 + k] = a[k];
            for (int o = 0; o < N; ++o)
            {
                double s = 0.0;
                for (int k = 0; k < K; ++k)
                    s += (double)a[k] * (double)g_half(W[o * K + k]);
                b[o] = (float)s;
            }
            if (l == 3)
                for (int o = 0; o < N; ++o)
                    out[o] = b[o];
            else
                for (int o = 0; o < N; ++o)
                    a[o] = g_half(std::max(b[o], 0.0f));
        }
    }
    template <int D>
    bool g_contains(const float *lo, const float *hi, const Vec<D> &x)
    {
        for (int k = 0; k < D; ++k)
            if (!(lo[k] <= x[k] && x[k] <= hi[k]))
                return false;
        return true;
    }
}

extern "C"
{
    void *orc_net_create(int dim, int n_levels, int base_res, float per_level_scale, const float *aabb_min, const float *aabb_max)
    {
        OrcNet *n = new OrcNet();
        n->dim = dim, n->n_levels = n_levels, n->base_res = base_res, n->per_level_scale = per_level_scale;
        const int feat = n_levels * 4;
        n->k0 = feat <= 16 ? 16 : (feat <= 32 ? 32 : 64);
        const float log2s = std::log2(per_level_scale);
        uint32_t off = 0;
        for (int l = 0; l < n_levels; ++l)
        {
            const float scale = std::exp2((float)l * log2s) * (float)base_res - 1.0f; // grid_scale, grid.h:193-198
            const uint32_t res = (uint32_t)std::ceil(scale) + 1u;                      // grid_resolution, :200-202
            uint64_t pts = 1;
            for (int d = 0; d < dim; ++d)
                pts *= res;
            pts = (pts + 7u) / 8u * 8u; // next_multiple(params_in_level, 8), :990
            n->scale.push_back(scale);
            n->res.push_back(res);
            n->off.push_back(off);
            off += (uint32_t)pts;
        }
        n->off.push_back(off);
        n->n_grid = (int)off * 4;
        n->n_mlp = 64 * n->k0 + 2 * 4096 + 48 * 64;
        float d2 = 0.0f;
        for (int k = 0; k < dim; ++k)
            d2 += (aabb_max[k] - aabb_min[k]) * (aabb_max[k] - aabb_min[k]);
        const float infl = std::sqrt(d2) * 0.005f;
        for (int k = 0; k < 3; ++k)
        {
            n->lo[k] = k < dim ? aabb_min[k] - infl : 0.0f;
            n->hi[k] = k < dim ? aabb_max[k] + infl : 1.0f;
        }
        n->params.assign((size_t)n->n_mlp + n->n_grid, 0.0f);
        return n;
    }
    void orc_net_destroy(void *net) { delete (OrcNet *)net; }
    int orc_net_num_params(void *net, int *n_mlp, int *k0)
    {
        OrcNet *n = (OrcNet *)net;
        if (n_mlp)
            *n_mlp = n->n_mlp;
        if (k0)
            *k0 = n->k0;
        return n->n_mlp + n->n_grid;
    }
    void orc_net_set_params(void *net, const float *params)
    {
        OrcNet *n = (OrcNet *)net;
        memcpy(n->params.data(), params, sizeof(float) * n->params.size());
    }
    /* raw outputs (n * 48) at world-space points; acts (n * 4 * 64, optional) = fp16-rounded layer inputs;
     * cells (optional): idx n*levels*2^dim u32 + weights */
    void orc_net_forward(void *net, const float *pts, int n, float *out, float *acts, uint32_t *cell_idx, float *cell_w)
    {
        OrcNet *N = (OrcNet *)net;
        const int D = N->dim, C = 1 << D;
        parallel_for(n, 0, [&](int i)
                     {
            float enc[64];
            if (D == 2)
            {
                Vec<2> x = {{pts[(size_t)i * 2], pts[(size_t)i * 2 + 1]}};
                g_encode<2>(*N, x, enc);
                if (cell_idx)
                {
                    float xn[3];
                    g_normalize<2>(*N, x, xn);
                    for (int l = 0; l < N->n_levels; ++l)
                        g_cell<2>(*N, l, xn, cell_idx + ((size_t)i * N->n_levels + l) * C, cell_w + ((size_t)i * N->n_levels + l) * C);
                }
            }
            else
            {
                Vec<3> x = {{pts[(size_t)i * 3], pts[(size_t)i * 3 + 1], pts[(size_t)i * 3 + 2]}};
                g_encode<3>(*N, x, enc);
                if (cell_idx)
                {
                    float xn[3];
                    g_normalize<3>(*N, x, xn);
                    for (int l = 0; l < N->n_levels; ++l)
                        g_cell<3>(*N, l, xn, cell_idx + ((size_t)i * N->n_levels + l) * C, cell_w + ((size_t)i * N->n_levels + l) * C);
                }
            }
            g_mlp_forward(*N, enc, out + (size_t)i * 48, acts ? acts + (size_t)i * 256 : nullptr); });
    }
    /* VMM probe: mixture pdf (incl. reflected direction on a Neumann boundary), its gradient w.r.t. the activated lobe
     * parameters (n * 8*(dim+2)), one sampled direction per row with Philox key (seed, row) */
    void orc_vmm_eval(int dim, const float *raw, const float *dir, const float *nrm, const uint8_t *onN, int n, uint32_t seed, float *pdf, float *grad,
                      float *sampled)
    {
        const int kOut = 8 * (dim + 2) + 1, kG = 8 * (dim + 2);
        for (int i = 0; i < n; ++i)
        {
            if (dim == 2)
            {
                const GMixture<2> m(raw + (size_t)i * kOut);
                const Vec<2> w = {{dir[i * 2], dir[i * 2 + 1]}}, nn = {{nrm[i * 2], nrm[i * 2 + 1]}};
                pdf[i] = g_grad_prob<2>(m, w, onN[i] != 0, nn, grad + (size_t)i * kG);
                const Philox4 q = wost_rng(seed, (uint32_t)i, 0u, 0u, WOST_RNG_SELECT, 0);
                const Vec<2> s = g_sample_lobe(m.sg[m.pick(u32_to_unit_float(q.y))], seed, (uint32_t)i, 0u, 0u);
                sampled[i * 2] = s[0], sampled[i * 2 + 1] = s[1];
            }
            else
            {
                const GMixture<3> m(raw + (size_t)i * kOut);
                const Vec<3> w = {{dir[i * 3], dir[i * 3 + 1], dir[i * 3 + 2]}}, nn = {{nrm[i * 3], nrm[i * 3 + 1], nrm[i * 3 + 2]}};
                pdf[i] = g_grad_prob<3>(m, w, onN[i] != 0, nn, grad + (size_t)i * kG);
                const Philox4 q = wost_rng(seed, (uint32_t)i, 0u, 0u, WOST_RNG_SELECT, 0);
                const Vec<3> s = g_sample_lobe(m.sg[m.pick(u32_to_unit_float(q.y))], seed, (uint32_t)i, 0u, 0u);
                sampled[i * 3] = s[0], sampled[i * 3 + 1] = s[1], sampled[i * 3 + 2] = s[2];
            }
        }
    }
    /* dL/d(raw output) of the KL loss for n rows (dl: n * 48, padded with zeros), per-row loss */
    void orc_kl_grad(int dim, const float *raw48, const float *dir, const float *nrm, const uint8_t *onN, const float *dirPdf, const float *li, int n,
                     float scale, float *dl, float *loss)
    {
        for (int i = 0; i < n; ++i)
        {
            float *g = dl + (size_t)i * 48;
            for (int k = 0; k < 48; ++k)
                g[k] = 0.0f;
            if (dim == 2)
            {
                const Vec<2> w = {{dir[i * 2], dir[i * 2 + 1]}}, nn = {{nrm[i * 2], nrm[i * 2 + 1]}};
                loss[i] = g_kl_grad<2>(raw48 + (size_t)i * 48, w, nn, onN[i] != 0, dirPdf[i], li[i], scale, g);
            }
            else
            {
                const Vec<3> w = {{dir[i * 3], dir[i * 3 + 1], dir[i * 3 + 2]}}, nn = {{nrm[i * 3], nrm[i * 3 + 1], nrm[i * 3 + 2]}};
                loss[i] = g_kl_grad<3>(raw48 + (size_t)i * 48, w, nn, onN[i] != 0, dirPdf[i], li[i], scale, g);
            }
        }
    }
    /* special-function probes for the reference's known answers (test/vonmises_test.cu) */
    float orc_eval_poly_large0(float y) { return g_eval_poly(y, COEF_L0, 9); }
    float orc_log_bessel(float x, int order) { return g_log_bessel(x, order); }
    float orc_vonmises_log_pdf(float kappa, float cos_theta) { return GVonMises{kappa}.log_eval(cos_theta); }
    float orc_vonmises_pdf(float kappa, float cos_theta) { return GVonMises{kappa}.eval(cos_theta); }
    float orc_vonmises_dlog_dkappa(float kappa, float cos_theta) { return GVonMises{kappa}.d_log_eval_d_kappa(cos_theta); }
    float orc_vmf_pdf(float kappa, float cos_theta) { return g_vmf_eval(kappa, cos_theta); }
}

// ------------------------------------------------------------------------------------------------ guided walker
namespace
{
    // source + Neumann boundary stages of one step (integrator/guided/integrator.cu:277-494 == the uniform stages of
    // wost_oracle_walk.inl plus the record hooks); every contribution goes through add(r, g, b)
    template <int D, class Add>
    void g_stage_contribs(const orc_problem_desc &p, const orc_settings &S, const Scene<D> *sn, const Vec<D> &x, float thp, bool onN, const Vec<D> &nN,
                          uint32_t pixel, uint32_t sample, uint32_t depth, float R_B, Add add)
    {
        const float eps = S.epsilon_shell;
        if (p.source)
        {
            const Philox4 q = wost_rng(S.seed, pixel, sample, depth, WOST_RNG_SOURCE_DIR, 0);
            Vec<D> dir;
            float dirPdf, alpha = 1.0f;
            if (onN)
            {
                dir = to_world(nN, sample_hemisphere_local(u32_to_unit_float(q.x), u32_to_unit_float(q.y), (Vec<D> *)nullptr));
                dirPdf = Dims<D>::hemi_pdf();
                alpha = 0.5f;
            }
            else
            {
                dir = sample_sphere(u32_to_unit_float(q.x), u32_to_unit_float(q.y), (Vec<D> *)nullptr);
                dirPdf = Dims<D>::sphere_pdf();
            }
            float dist = R_B;
            if (sn)
            {
                Vec<D> o;
                for (int k = 0; k < D; ++k)
                    o[k] = x[k] + eps * dir[k];
                const RayResult rr = query_ray(*sn, Ray<D>(o, dir), dist, false);
                if (rr.hit)
                    dist = dm_fmin(rr.t, dist);
            }
            Draws dr; // Philox stream only (the guided walker has no PCG mode): 2D restarts a block per iteration, 3D continues with words z, w of q
            dr.mode = 0;
            dr.seed = S.seed;
            dr.pixel = pixel;
            dr.sample = sample;
            dr.depth = depth;
            dr.pcg = nullptr;
            dr.q = q;
            dr.k = 2;
            const float r = green_sample_radius(R_B, dr, (Vec<D> *)nullptr);
            if (r <= dist)
            {
                float w3[3] = {x[0] + r * dir[0], x[1] + r * dir[1], 0.0f};
                if (D == 3)
                    w3[2] = x[D - 1] + r * dir[D - 1];
                float g[3], c[3];
                source_lookup(p, w3, g);
                for (int ch = 0; ch < 3; ++ch)
                    c[ch] = thp * (g[ch] * p.source_intensity) * Dims<D>::green_norm(R_B) * Dims<D>::cond_pdf(Dims<D>::sphere_pdf(), r) / Dims<D>::cond_pdf(dirPdf, r) / alpha;
                add(c[0], c[1], c[2]);
            }
        }
        if (sn)
        {
            const Philox4 q = wost_rng(S.seed, pixel, sample, depth, WOST_RNG_NEUMANN, 0);
            const float u0 = u32_to_unit_float(q.x), u1 = u32_to_unit_float(q.y), u2 = u32_to_unit_float(q.z);
            int obj;
            float pdf;
            sample_in_sphere(*sn, x, R_B, u0, obj, pdf);
            if (obj != -1 && pdf > 0)
            {
                const Vec<D> y = sample_on_prim(*sn, obj, u1, u2);
                const float r = length(sub(y, x));
                if (r < R_B && r > 0)
                {
                    Vec<D> o = x;
                    if (onN)
                        for (int k = 0; k < D; ++k)
                            o[k] = o[k] + eps * nN[k];
                    Vec<D> rd = sub(y, o);
                    const float cl = length(rd);
                    if (cl > 0)
                        for (int k = 0; k < D; ++k)
                            rd[k] = rd[k] / cl;
                    if (!query_ray(*sn, Ray<D>(o, rd), cl - eps, true).hit)
                    {
                        int side = point_side(*sn, obj, x);
                        float uv[2];
                        proj_ratio(*sn, obj, y, uv);
                        if (onN)
                        {
                            const float dn = dot(prim_normal(*sn, obj), nN);
                            side = (0.0f < dn) - (dn < 0.0f);
                        }
                        if (side != 0)
                        {
                            float c[3];
                            surface_color<D>(p.color_neumann, sn->P(obj), side, uv, c);
                            const float k = thp * Dims<D>::green_eval(R_B, r) / (onN ? 0.5f : 1.0f) / pdf;
                            add(-(c[0] * p.neumann_intensity * k), -(c[1] * p.neumann_intensity * k), -(c[2] * p.neumann_intensity * k));
                        }
                    }
                }
            }
        }
    }

    struct GSample // one training sample (GuidedInput + GuidedOutput, guideditem.h:38-51)
    {
        float pos[3], dir[3], nrm[3], pdf, li;
        uint8_t onN;
    };
    template <int D>
    void g_walk_pixel(const Problem &P, const orc_settings &S, const orc_guided_settings &G, const OrcNet *net, const Scene<D> *sd, const Scene<D> *sn,
                      int pixel, int sample0, int nsamples, float sol[3], uint64_t &steps, std::vector<GSample> *emit)
    {
        const orc_problem_desc &p = P.d;
        const float eps = S.epsilon_shell;
        for (int sample = sample0; sample < sample0 + nsamples; ++sample)
        {
            Vec<D> x = eval_point<D>(p, S.frame_w, S.frame_h, pixel);
            float thp = 1.0f, T[3] = {0.0f, 0.0f, 0.0f};
            bool onN = false;
            Vec<D> nN;
            for (int k = 0; k < D; ++k)
                nN[k] = 0.0f;
            struct Rec
            {
                Vec<D> pos, dir, nrm;
                float pdf, thp, base[3];
                bool onN;
            } rec[4];
            int nrec = 0;
            auto add = [&](float c0, float c1, float c2)
            {
                sol[0] = c0 + sol[0], sol[1] = c1 + sol[1], sol[2] = c2 + sol[2];
                T[0] += c0, T[1] += c1, T[2] += c2;
            };
            for (int depth = 0; depth < S.max_walking_depth; ++depth)
            {
                ++steps;
                // separate + boundary (integrator.cu:153-275)
                float R_D = INF_F;
                if (sd)
                {
                    uint32_t idx;
                    query_nearest(*sd, x, idx, R_D);
                    const int side = point_side(*sd, (int)idx, x);
                    float uv[2];
                    proj_ratio(*sd, (int)idx, x, uv);
                    if (R_D < eps && uv_strictly_inside<D>(uv))
                    {
                        float c[3];
                        surface_color<D>(p.color_dirichlet, sd->P((int)idx), side, uv, c);
                        add(c[0] * p.dirichlet_intensity * thp, c[1] * p.dirichlet_intensity * thp, c[2] * p.dirichlet_intensity * thp);
                        break;
                    }
                }
                float R_N = INF_F;
                if (sn)
                    R_N = query_silhouette(*sn, x, false);
                const float R_B = dm_fmax(1e-4f, dm_fmin(R_D, R_N)); // no 0.99 here (integrator.cu:238-239)
                if (dm_isinf(R_B))
                    break;
                g_stage_contribs<D>(p, S, sn, x, thp, onN, nN, (uint32_t)pixel, (uint32_t)sample, (uint32_t)depth, R_B, add);

                // direction (handleOutShellPoint / handleGuidedSampling / handleUniformSampling / oneStepWalk, :496-965)
                Vec<D> dir;
                float pdf;
                const float uniformPdf = onN ? Dims<D>::hemi_pdf() : Dims<D>::sphere_pdf();
                auto uniform_dir = [&]()
                {
                    const Philox4 q = wost_rng(S.seed, (uint32_t)pixel, (uint32_t)sample, (uint32_t)depth, WOST_RNG_WALK, 0);
                    if (onN)
                        return to_world(nN, sample_hemisphere_local(u32_to_unit_float(q.x), u32_to_unit_float(q.y), (Vec<D> *)nullptr));
                    return sample_sphere(u32_to_unit_float(q.x), u32_to_unit_float(q.y), (Vec<D> *)nullptr);
                };
                if (net && depth < G.max_guided_depth)
                {
                    float enc[64], raw[48];
                    g_encode<D>(*net, x, enc);
                    g_mlp_forward(*net, enc, raw, nullptr);
                    const float p_sel = g_logistic(raw[8 * (D + 2)]);
                    const bool inside = g_contains<D>(p.aabb_min, p.aabb_max, x);
                    const Philox4 qs = wost_rng(S.seed, (uint32_t)pixel, (uint32_t)sample, (uint32_t)depth, WOST_RNG_SELECT, 0);
                    const bool guided = inside && G.uniform_fraction < 1.0f && (G.uniform_fraction == 0.0f || u32_to_unit_float(qs.x) < p_sel);
                    if (guided)
                    {
                        const GMixture<D> m(raw);
                        dir = g_sample_lobe(m.sg[m.pick(u32_to_unit_float(qs.y))], S.seed, (uint32_t)pixel, (uint32_t)sample, (uint32_t)depth);
                        float guidedPdf = m.pdf(dir);
                        if (onN)
                        {
                            const Vec<D> refl = g_reflect(dir, nN);
                            if (dot(nN, dir) <= 0.0f)
                                dir = refl;
                            guidedPdf += m.pdf(refl);
                        }
                        pdf = p_sel * guidedPdf + (1.0f - p_sel) * uniformPdf;
                    }
                    else
                    {
                        dir = uniform_dir();
                        pdf = uniformPdf;
                        if (inside)
                        {
                            const GMixture<D> m(raw);
                            float guidedPdf = m.pdf(dir);
                            if (onN)
                                guidedPdf += m.pdf(g_reflect(dir, nN));
                            pdf = p_sel * guidedPdf + (1.0f - p_sel) * pdf;
                        }
                    }
                }
                else
                {
                    dir = uniform_dir();
                    pdf = uniformPdf;
                }
                if (emit && depth < G.max_train_depth && nrec < 4)
                {
                    Rec &r = rec[nrec++];
                    r.pos = x, r.dir = dir, r.nrm = nN, r.pdf = pdf, r.thp = thp, r.onN = onN;
                    r.base[0] = T[0], r.base[1] = T[1], r.base[2] = T[2];
                }
                if (depth + 1 >= S.max_walking_depth)
                    break;
                // advance (second half of the sampling kernels)
                Vec<D> cur = x, next, normal;
                float alpha = 1.0f;
                if (onN)
                {
                    alpha = 0.5f;
                    for (int k = 0; k < D; ++k)
                        cur[k] = cur[k] + eps * nN[k];
                }
                for (int k = 0; k < D; ++k)
                {
                    next[k] = x[k] + R_B * dir[k];
                    normal[k] = 0.0f;
                }
                bool hit = false;
                if (sn)
                {
                    const RayResult rr = query_ray(*sn, Ray<D>(cur, dir), R_B, false);
                    hit = rr.hit;
                    if (hit)
                    {
                        normal = prim_normal(*sn, (int)rr.obj);
                        if (dot(normal, dir) > 0)
                            for (int k = 0; k < D; ++k)
                                normal[k] = -normal[k];
                        for (int k = 0; k < D; ++k)
                            next[k] = cur[k] + rr.t * dir[k];
                    }
                }
                x = next;
                thp = thp / pdf / alpha / Dims<D>::sphere_measure();
                onN = hit;
                nN = normal;
            }
            // generate_training_data (train.h:422-471)
            for (int i = 0; emit && i < nrec; ++i)
            {
                const Rec &r = rec[i];
                float li = 0.0f;
                bool nan = false;
                for (int ch = 0; ch < 3; ++ch)
                {
                    float v = 0.0f;
                    if (fabsf(r.thp) > G_EPS)
                        v = (T[ch] - r.base[ch]) / r.thp;
                    nan = nan || std::isnan(v);
                    li += fabsf(v);
                }
                li /= 3.0f;
                bool bad = nan || std::isnan(r.pdf) || r.pdf == 0.0f;
                for (int k = 0; k < D; ++k)
                    bad = bad || std::isnan(r.pos[k]) || std::isnan(r.dir[k]);
                if (!g_contains<D>(p.aabb_min, p.aabb_max, r.pos) || bad)
                    continue;
                GSample s;
                memset(&s, 0, sizeof(s));
                for (int k = 0; k < D; ++k)
                    s.pos[k] = r.pos[k], s.dir[k] = r.dir[k], s.nrm[k] = r.nrm[k];
                s.pdf = r.pdf, s.li = li, s.onN = r.onN ? 1 : 0;
                emit->push_back(s);
            }
        }
    }
}

extern "C"
{
    /* Guided WoSt with FIXED network weights over pixels [pixel_begin, pixel_end), samples [sample0, sample0 + nsamples).
     * solution: W*H*3, the SUM of the samples' contributions divided by settings.samples_per_pixel.
     * samples_out (optional, capacity cap rows of 11 floats: pos3 dir3 nrm3 pdf li, + onN in samples_on): the training
     * set the walks would produce, ordered by pixel then record depth; returns the number of samples. */
    int orc_solve_guided(const orc_problem_desc *p, const orc_settings *s, const orc_guided_settings *g, void *net, int pixel_begin, int pixel_end,
                         int sample0, int nsamples, float *solution, uint64_t *steps, float *samples_out, uint8_t *samples_on, int cap, int nthreads)
    {
        Problem P;
        P.d = *p;
        P.dir = (AnyScene *)p->dirichlet;
        P.neu = (AnyScene *)p->neumann;
        const int n = pixel_end - pixel_begin;
        std::atomic<uint64_t> total{0};
        std::vector<std::vector<GSample>> per_pixel(samples_out ? n : 0);
        parallel_for(n, nthreads, [&](int i)
                     {
            const int pixel = pixel_begin + i;
            float *sol = solution + (size_t)pixel * 3;
            sol[0] = sol[1] = sol[2] = 0.0f;
            if (p->mask && !p->mask[pixel])
                return;
            uint64_t st = 0;
            std::vector<GSample> *em = samples_out ? &per_pixel[i] : nullptr;
            if (p->dim == 2)
                g_walk_pixel<2>(P, *s, *g, (OrcNet *)net, P.dir ? P.dir->s2 : nullptr, P.neu ? P.neu->s2 : nullptr, pixel, sample0, nsamples, sol, st, em);
            else
                g_walk_pixel<3>(P, *s, *g, (OrcNet *)net, P.dir ? P.dir->s3 : nullptr, P.neu ? P.neu->s3 : nullptr, pixel, sample0, nsamples, sol, st, em);
            for (int ch = 0; ch < 3; ++ch)
                sol[ch] = sol[ch] / (float)s->samples_per_pixel;
            total.fetch_add(st, std::memory_order_relaxed); });
        if (steps)
            *steps = total.load();
        int count = 0;
        for (auto &v : per_pixel)
            for (const GSample &sm : v)
            {
                if (count < cap)
                {
                    float *o = samples_out + (size_t)count * 11;
                    for (int k = 0; k < 3; ++k)
                        o[k] = sm.pos[k], o[3 + k] = sm.dir[k], o[6 + k] = sm.nrm[k];
                    o[9] = sm.pdf, o[10] = sm.li;
                    samples_on[count] = sm.onN;
                }
                ++count;
            }
        return count;
    }
}
