// TEST INFRASTRUCTURE ONLY -- included by wost_oracle.cpp. Scalar restatement of the uniform
// WoSt integrator (integrator/uniform/integrator.cu:64-623) on top of the oracle LBVH.
// One walk = one (pixel, sample); the per-depth stage order of the reference
// (separate -> boundary -> source -> neumann -> walk, integrator.cu:552-575) is kept inside the loop.
// RNG: mode 0 (default) = the Philox stream of elaina_b200/csrc/philox.h (what the CUDA product draws, bit for bit);
// mode 1 = the reference's per-pixel PCG32 (core/sampler.h) in its consumption order, for field-for-field comparison with
// the reference's own CUDA build (baseline/_ref/elaina-exec-ref, tests/golden/b1_*.npz).
// PARITY UNPINNED for this layer (uniform walker): the reference ships no golden fields and its integrators cannot run without a GPU
// build. What IS pinned: the LBVH layer underneath (bit-exact against the host-compiled reference headers,
// oracle/_ref) and the special functions (known answers of test/vonmises_test.cu); fields are checked against analytic solutions.

namespace
{
    struct Problem
    {
        orc_problem_desc d;
        AnyScene *dir, *neu;
    };

    // core/evaluation_grid.h:27-33 (2D) / 60-64 (3D)
    template <int D>
    inline Vec<D> eval_point(const orc_problem_desc &p, int W, int H, int pixel)
    {
        const int px = pixel % W, py = pixel / W;
        const float nx = 2.0f * (float)px / (float)W + (-1.0f);
        const float ny = 2.0f * (float)py / (float)H + (-1.0f);
        Vec<D> r;
        if (D == 2)
        {
            const float ux = p.grid_up[1], uy = -p.grid_up[0];
            const float vx = p.grid_up[0], vy = p.grid_up[1];
            r[0] = p.grid_scale * (nx * ux + ny * vx) + p.grid_pos[0];
            r[1] = p.grid_scale * (nx * uy + ny * vy) + p.grid_pos[1];
        }
        else
        {
            for (int k = 0; k < 3; ++k)
                r[k] = p.grid_scale * (nx * p.grid_right[k] + ny * p.grid_up[k]) + p.grid_pos[k];
        }
        return r;
    }

    // core/utility.cuh:569-635
    inline int point_side(const Scene<2> &s, int obj, const Vec<2> &x)
    {
        const Vec<2> &p0 = s.V(s.P(obj)[0]), &p1 = s.V(s.P(obj)[1]);
        const float dx = p1[0] - p0[0], dy = p1[1] - p0[1];
        const float nx = -dy, ny = dx;
        const float vx = x[0] - p0[0], vy = x[1] - p0[1];
        const float dp = vx * nx + vy * ny;
        return dp > 0 ? 1 : (dp < 0 ? -1 : 0);
    }
    inline int point_side(const Scene<3> &s, int obj, const Vec<3> &x)
    {
        const Vec<3> &p0 = s.V(s.P(obj)[0]), &p1 = s.V(s.P(obj)[1]), &p2 = s.V(s.P(obj)[2]);
        const Vec<3> n = cross(sub(p1, p0), sub(p2, p0));
        const Vec<3> v = sub(x, p0);
        const float dp = v[0] * n[0] + v[1] * n[1] + v[2] * n[2];
        return dp > 0 ? 1 : (dp < 0 ? -1 : 0);
    }
    // core/utility.cuh:637-708 (unclamped)
    inline void proj_ratio(const Scene<2> &s, int obj, const Vec<2> &x, float uv[2])
    {
        const Vec<2> &p0 = s.V(s.P(obj)[0]), &p1 = s.V(s.P(obj)[1]);
        const float dx = p1[0] - p0[0], dy = p1[1] - p0[1];
        const float l2 = dx * dx + dy * dy;
        uv[1] = 0.0f;
        if (l2 < 1e-8f)
        {
            uv[0] = 0.0f;
            return;
        }
        const float vx = x[0] - p0[0], vy = x[1] - p0[1];
        uv[0] = (vx * dx + vy * dy) / l2;
    }
    inline void proj_ratio(const Scene<3> &s, int obj, const Vec<3> &x, float uv[2])
    {
        const Vec<3> &p0 = s.V(s.P(obj)[0]), &p1 = s.V(s.P(obj)[1]), &p2 = s.V(s.P(obj)[2]);
        const Vec<3> e1 = sub(p1, p0), e2 = sub(p2, p0), v = sub(x, p0);
        const float a = dot(e1, e1), b = dot(e1, e2), c = dot(e2, e2), d = dot(v, e1), e = dot(v, e2);
        const float den = a * c - b * b;
        uv[0] = (d * c - e * b) / den;
        uv[1] = (e * a - d * b) / den;
    }
    template <int D>
    inline bool uv_strictly_inside(const float uv[2]) // integrator/uniform/integrator.cu:159-168
    {
        if (D == 2)
            return uv[0] > 0.0f && uv[0] < 1.0f;
        return uv[0] > 0.0f && uv[1] > 0.0f && uv[0] + uv[1] < 1.0f;
    }
    // integrator/common.h:242-260 + krrmath/functors.h:60-95
    template <int D>
    inline void surface_color(const float *pairs, const int *vidx, int side, const float uv[2], float out[3])
    {
        const float *c[3];
        for (int i = 0; i < D; ++i)
            c[i] = pairs ? pairs + (size_t)vidx[i] * 6 + (side >= 0 ? 0 : 3) : nullptr;
        for (int ch = 0; ch < 3; ++ch)
        {
            if (!pairs)
            {
                out[ch] = 0.0f;
                continue;
            }
            if (D == 2)
                out[ch] = c[0][ch] * (1 - uv[0]) + c[1][ch] * uv[0];
            else
            {
                const float w = 1 - uv[0] - uv[1];
                out[ch] = c[0][ch] * w + c[1][ch] * uv[0] + c[2][ch] * uv[1];
            }
        }
    }
    inline Vec<2> prim_normal(const Scene<2> &s, int obj) // scene.cuh:400-406
    {
        const Vec<2> &p0 = s.V(s.P(obj)[0]), &p1 = s.V(s.P(obj)[1]);
        const Vec<2> ls = sub(p1, p0);
        return normalize(Vec<2>{{-ls[1], ls[0]}});
    }
    inline Vec<3> prim_normal(const Scene<3> &s, int obj) // scene.cuh:842-852
    {
        const Vec<3> &pa = s.V(s.P(obj)[0]), &pb = s.V(s.P(obj)[1]), &pc = s.V(s.P(obj)[2]);
        return normalize(cross(sub(pc, pa), sub(pb, pa)));
    }
    inline Vec<2> sample_on_prim(const Scene<2> &s, int obj, float u, float) // scene.cuh:29-32,614-622
    {
        const Vec<2> &pa = s.V(s.P(obj)[0]), &pb = s.V(s.P(obj)[1]);
        return {{pa[0] + u * (pb[0] - pa[0]), pa[1] + u * (pb[1] - pa[1])}};
    }
    inline Vec<3> sample_on_prim(const Scene<3> &s, int obj, float u, float v) // scene.cuh:14-27
    {
        const Vec<3> &pa = s.V(s.P(obj)[0]), &pb = s.V(s.P(obj)[1]), &pc = s.V(s.P(obj)[2]);
        if (u + v > 1.0f)
        {
            u = 1.0f - u;
            v = 1.0f - v;
        }
        const float w = 1.0f - u - v;
        return {{w * pa[0] + u * pb[0] + v * pc[0], w * pa[1] + u * pb[1] + v * pc[1], w * pa[2] + u * pb[2] + v * pc[2]}};
    }

    // util/sampling.h:18-33 (sphere), 68-85 (hemisphere, local frame)
    inline Vec<2> sample_sphere(float u1, float, Vec<2> *) { return {{m_cos(u1 * DM_2PI), m_sin(u1 * DM_2PI)}}; }
    inline Vec<3> sample_sphere(float u1, float u2, Vec<3> *)
    {
        const float z = 1 - 2 * u1;
        const float r = sqrtf(1 - z * z);
        const float phi = DM_2PI * u2;
        return {{r * m_cos(phi), r * m_sin(phi), z}};
    }
    inline Vec<2> sample_hemisphere_local(float u1, float, Vec<2> *) { return {{m_cos(DM_PI * u1), m_sin(DM_PI * u1)}}; }
    inline Vec<3> sample_hemisphere_local(float u1, float u2, Vec<3> *)
    {
        const float z = u1;
        const float r = sqrtf(dm_max(0.0f, 1.0f - z * z));
        const float phi = 2.0f * DM_PI * u2;
        return {{r * m_cos(phi), r * m_sin(phi), z}};
    }
    // util/transformation.h:52-67 + util/math_utils.h:141-156
    inline Vec<2> to_world(const Vec<2> &n, const Vec<2> &v)
    {
        // N = n, T = -perp(n), perp(n) = normalized(-n.y, n.x); world = T*v[0] + N*v[1]
        Vec<2> perp = {{-n[1], n[0]}};
        const float l2 = perp[0] * perp[0] + perp[1] * perp[1];
        if (l2 > 0.0f)
        {
            const float l = sqrtf(l2);
            perp[0] /= l;
            perp[1] /= l;
        }
        return {{-perp[0] * v[0] + n[0] * v[1], -perp[1] * v[0] + n[1] * v[1]}};
    }
    inline Vec<3> safe_normalized(const Vec<3> &a)
    {
        const float l2 = dot(a, a);
        if (!(l2 > 0.0f))
            return a;
        const float l = sqrtf(l2);
        return {{a[0] / l, a[1] / l, a[2] / l}};
    }
    inline Vec<3> to_world(const Vec<3> &n, const Vec<3> &v)
    {
        const float ax = fabsf(n[0]), ay = fabsf(n[1]), az = fabsf(n[2]);
        const uint32_t uyx = (ax - ay) < 0 ? 1 : 0, uzx = (ax - az) < 0 ? 1 : 0, uzy = (ay - az) < 0 ? 1 : 0;
        const uint32_t xm = uyx & uzx, ym = (1 ^ xm) & uzy, zm = 1 ^ (xm | ym);
        const Vec<3> t = safe_normalized(cross(n, Vec<3>{{(float)xm, (float)ym, (float)zm}}));
        const Vec<3> b = safe_normalized(cross(n, t));
        return {{t[0] * v[0] + b[0] * v[1] + n[0] * v[2], t[1] * v[0] + b[1] * v[1] + n[1] * v[2], t[2] * v[0] + b[2] * v[1] + n[2] * v[2]}};
    }

    template <int D>
    struct Dims;
    template <>
    struct Dims<2>
    {
        static float sphere_pdf() { return 1.0f / DM_2PI; }
        static float hemi_pdf() { return 1.0f / DM_PI; }
        static float sphere_measure() { return DM_2PI; }
        static float cond_pdf(float dirPdf, float r) { return dirPdf / r; }
        static float green_eval(float R, float r) { return m_log(R / r) / DM_2PI; } // util/green.h:25-28
        static float green_norm(float R) { return R * R / 4.0f; }                   // util/green.h:30-33
    };
    template <>
    struct Dims<3>
    {
        static float sphere_pdf() { return 1.0f / DM_4PI; }
        static float hemi_pdf() { return 1.0f / DM_2PI; }
        static float sphere_measure() { return DM_4PI; }
        static float cond_pdf(float dirPdf, float r) { return dirPdf / (r * r); }
        static float green_eval(float R, float r) { return (1.0f / r - 1.0f / R) / DM_4PI; } // util/green.h:84-91
        static float green_norm(float R) { return R * R / 6.0f; }
    };

    // The reference's per-pixel generator (core/sampler.h:10-103): PCG32 (64-bit LCG, XSH-RR output), seeded by
    // setPixelSample(pixelCoord, 0) and advanced by 256 * pixelId (integrator/uniform/integrator.cu:72-77), then consumed
    // in kernel order: sample by sample, depth by depth, source -> Neumann -> walk. Only used in rng mode 1, which exists
    // so that this oracle can be compared field-for-field with the reference's own CUDA build (baseline/_ref).
    struct Pcg32
    {
        uint64_t state = 0, inc = 1;
        static constexpr uint64_t MULT = 0x5851f42d4c957f2dULL;
        uint32_t next_uint()
        {
            const uint64_t old = state;
            state = old * MULT + inc;
            const uint32_t xs = (uint32_t)(((old >> 18u) ^ old) >> 27u);
            const uint32_t rot = (uint32_t)(old >> 59u);
            return (xs >> rot) | (xs << ((~rot + 1u) & 31));
        }
        void set_seed(uint64_t initstate, uint64_t initseq) // sampler.h:20-27
        {
            state = 0u;
            inc = (initseq << 1u) | 1u;
            next_uint();
            state += initstate;
            next_uint();
        }
        void advance(int64_t delta) // sampler.h:47-63
        {
            uint64_t cur_mult = MULT, cur_plus = inc, acc_mult = 1u, acc_plus = 0u;
            while (delta > 0)
            {
                if (delta & 1)
                {
                    acc_mult *= cur_mult;
                    acc_plus = acc_plus * cur_mult + cur_plus;
                }
                cur_plus = (cur_mult + 1) * cur_plus;
                cur_mult *= cur_mult;
                delta /= 2;
            }
            state = acc_mult * state + acc_plus;
        }
        float next_float() // sampler.h:93-103
        {
            union
            {
                uint32_t u;
                float f;
            } x;
            x.u = (next_uint() >> 9) | 0x3f800000u;
            return x.f - 1.0f;
        }
        double next_double() // sampler.h:79-91
        {
            union
            {
                uint64_t u;
                double d;
            } x;
            x.u = ((uint64_t)next_uint() << 20) | 0x3ff0000000000000ULL;
            return x.d - 1.0;
        }
    };
    inline uint32_t interleave_32bit(uint32_t vx, uint32_t vy) // util/hash.h:13-28
    {
        uint32_t x = vx & 0x0000ffff, y = vy & 0x0000ffff;
        x = (x | (x << 8)) & 0x00FF00FF;
        x = (x | (x << 4)) & 0x0F0F0F0F;
        x = (x | (x << 2)) & 0x33333333;
        x = (x | (x << 1)) & 0x55555555;
        y = (y | (y << 8)) & 0x00FF00FF;
        y = (y | (y << 4)) & 0x0F0F0F0F;
        y = (y | (y << 2)) & 0x33333333;
        y = (y | (y << 1)) & 0x55555555;
        return x | (y << 1);
    }
    inline Pcg32 pcg_for_pixel(int pixel, int W) // uniform/integrator.cu:72-77
    {
        Pcg32 g;
        g.set_seed(interleave_32bit((uint32_t)(pixel % W), (uint32_t)(pixel / W)), 0);
        g.advance((int64_t)(256 * pixel));
        return g;
    }

    // Where the walker's uniform numbers come from. Mode 0 (product arithmetic): the counter-based Philox stream of
    // elaina_b200/csrc/philox.h, one 4-word block per (pixel, sample, depth, purpose, iteration), words consumed in order.
    // Mode 1 (reference arithmetic): the pixel's PCG32 sequence, consumed in call order; begin() is a no-op.
    struct Draws
    {
        int mode;
        uint32_t seed, pixel, sample, depth;
        Pcg32 *pcg;
        Philox4 q;
        int k;
        void begin(uint32_t purpose, uint32_t iter)
        {
            if (mode == 0)
            {
                q = wost_rng(seed, pixel, sample, depth, purpose, iter);
                k = 0;
            }
        }
        float next()
        {
            if (mode != 0)
                return pcg->next_float();
            const uint32_t w = k == 0 ? q.x : (k == 1 ? q.y : (k == 2 ? q.z : q.w));
            ++k;
            return u32_to_unit_float(w);
        }
    };

    // util/green.h:45-74 (2D: bounded rejection loop) and :98-113 (3D: closed form)
    inline float green_sample_radius(float R, Draws &dr, Vec<2> *)
    {
        const float bound = 1.5f / R;
        float r = 0.0f;
        int iter = 0;
        do
        {
            dr.begin(WOST_RNG_SOURCE_RAD, (uint32_t)iter);
            const float u = dr.next();
            r = dr.next() * R;
            const float pdf = Dims<2>::green_eval(R, r) / Dims<2>::green_norm(R);
            const float pdfRadius = pdf / Dims<2>::sphere_pdf();
            ++iter;
            if (u < pdfRadius / bound)
                break;
        } while (iter < 1000);
        r = dm_max(1e-4f, r);
        if (r > R)
            r = R / 2.0f;
        return r;
    }
    inline float green_sample_radius(float R, Draws &dr, Vec<3> *) // words z, w of the direction block in Philox mode
    {
        const float u1 = dr.next();
        const float u2 = dr.next();
        const float phi = DM_2PI * u2;
        float r = (1.0f + sqrtf(1.0f - m_cbrt(u1 * u1)) * m_cos(phi)) * R / 2.0f;
        r = dm_max(1e-4f, r);
        if (r > R)
            r = R / 2.0f;
        return r;
    }

    // dense Vec3f source grid, trilinear (NanoVDB SampleFromVoxels<Tree,1,false>:
    // ext/openvdb/nanovdb/nanovdb/math/SampleFromVoxels.h:202-243,413-419): lerp z, then y, then x; a + w*(b-a)
    inline void source_lookup(const orc_problem_desc &p, const float w3[3], float out[3])
    {
        const float inv = 1.0f / p.source_voxel;
        float f[3];
        int ijk[3];
        for (int k = 0; k < 3; ++k)
        {
            const float x = (w3[k] - p.source_origin[k]) * inv;
            const float fl = floorf(x);
            ijk[k] = (int)fminf(fmaxf(fl, -2.0f), 1.0e9f); // far outside the grid: stay clear of INT_MAX + 1
            f[k] = x - fl;
        }
        auto fetch = [&](int i, int j, int k, int ch) -> float
        {
            if (i < 0 || j < 0 || k < 0 || i >= p.source_dims[0] || j >= p.source_dims[1] || k >= p.source_dims[2])
                return 0.0f;
            return p.source[(((size_t)k * p.source_dims[1] + j) * p.source_dims[0] + i) * 3 + ch];
        };
        for (int ch = 0; ch < 3; ++ch)
        {
            float v[2][2][2];
            for (int a = 0; a < 2; ++a)
                for (int b = 0; b < 2; ++b)
                    for (int c = 0; c < 2; ++c)
                        v[a][b][c] = fetch(ijk[0] + a, ijk[1] + b, ijk[2] + c, ch);
            auto lerp = [](float a, float b, float w)
            { return a + w * (b - a); };
            out[ch] = lerp(lerp(lerp(v[0][0][0], v[0][0][1], f[2]), lerp(v[0][1][0], v[0][1][1], f[2]), f[1]),
                           lerp(lerp(v[1][0][0], v[1][0][1], f[2]), lerp(v[1][1][0], v[1][1][1], f[2]), f[1]), f[0]);
        }
    }

    template <int D>
    void walk_pixel(const Problem &P, const orc_settings &S, const Scene<D> *sd, const Scene<D> *sn, int pixel,
                    float sol[3], uint64_t &steps)
    {
        const orc_problem_desc &p = P.d;
        const float eps = S.epsilon_shell;
        sol[0] = sol[1] = sol[2] = 0.0f;
        Pcg32 pcg = pcg_for_pixel(pixel, S.frame_w);
        Draws dr;
        dr.mode = g_rng_mode;
        dr.seed = S.seed;
        dr.pixel = (uint32_t)pixel;
        dr.pcg = &pcg;
        auto draw_dir = [&dr](float &u1, float &u2) // uniformSampleSphere/Hemisphere: one draw in 2D, two in 3D (util/sampling.h:18-33,68-85)
        {
            u1 = dr.next();
            u2 = D == 3 ? dr.next() : 0.0f;
        };
        for (int sample = 0; sample < S.samples_per_pixel; ++sample)
        {
            dr.sample = (uint32_t)sample;
            Vec<D> x = eval_point<D>(p, S.frame_w, S.frame_h, pixel);
            float thp = 1.0f;
            bool onN = false;
            Vec<D> nN;
            for (int k = 0; k < D; ++k)
                nN[k] = 0.0f;
            for (int depth = 0; depth < S.max_walking_depth; ++depth)
            {
                ++steps;
                dr.depth = (uint32_t)depth;
                if (g_trace) // diagnostics: positions of the walk (single-threaded use)
                    for (int k = 0; k < D; ++k)
                        g_trace[(size_t)depth * D + k] = x[k];
                // ---- separate (integrator.cu:102-212)
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
                        // ---- boundary (integrator.cu:214-232)
                        float c[3];
                        surface_color<D>(p.color_dirichlet, sd->P((int)idx), side, uv, c);
                        for (int ch = 0; ch < 3; ++ch)
                            sol[ch] = c[ch] * p.dirichlet_intensity * thp + sol[ch];
                        break;
                    }
                }
                float R_N = INF_F;
                if (sn)
                    R_N = query_silhouette(*sn, x, false);
                float R_B = dm_fmax(1e-4f, dm_fmin(R_D, R_N));
                R_B *= 0.99f;
                if (dm_isinf(R_B))
                    break;

                // ---- source (integrator.cu:234-316)
                if (p.source)
                {
                    dr.begin(WOST_RNG_SOURCE_DIR, 0);
                    float du1, du2;
                    draw_dir(du1, du2);
                    Vec<D> dir;
                    float dirPdf, alpha = 1.0f;
                    if (onN)
                    {
                        dir = to_world(nN, sample_hemisphere_local(du1, du2, (Vec<D> *)nullptr));
                        dirPdf = Dims<D>::hemi_pdf();
                        alpha = 0.5f;
                    }
                    else
                    {
                        dir = sample_sphere(du1, du2, (Vec<D> *)nullptr);
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
                    const float r = green_sample_radius(R_B, dr, (Vec<D> *)nullptr);
                    if (r <= dist)
                    {
                        float w3[3] = {x[0] + r * dir[0], x[1] + r * dir[1], 0.0f};
                        if (D == 3)
                            w3[2] = x[D - 1] + r * dir[D - 1];
                        float g[3];
                        source_lookup(p, w3, g);
                        const float f = thp;
                        for (int ch = 0; ch < 3; ++ch)
                        {
                            const float gv = g[ch] * p.source_intensity;
                            const float c = f * gv * Dims<D>::green_norm(R_B) * Dims<D>::cond_pdf(Dims<D>::sphere_pdf(), r) / Dims<D>::cond_pdf(dirPdf, r) / alpha;
                            sol[ch] = c + sol[ch];
                        }
                    }
                }

                // ---- neumann boundary sample (integrator.cu:318-445)
                if (sn)
                {
                    // float u[DIM] filled by DIM draws; the 2D code reads u[2] past the array, ignored by sample_on_object<2> (:343-365)
                    dr.begin(WOST_RNG_NEUMANN, 0);
                    const float u0 = dr.next();
                    const float u1 = dr.next();
                    const float u2 = D == 3 ? dr.next() : 0.0f;
                    int obj;
                    float pdf;
                    sample_in_sphere(*sn, x, R_B, u0, obj, pdf);
                    if (obj != -1 && pdf > 0)
                    {
                        const Vec<D> y = sample_on_prim(*sn, obj, u1, u2);
                        const float r = length(sub(y, x));
                        if (r < R_B && r > 0)
                        {
                            bool blocked = false;
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
                                blocked = query_ray(*sn, Ray<D>(o, rd), cl - eps, true).hit;
                            }
                            if (!blocked)
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
                                    const float alpha = onN ? 0.5f : 1.0f;
                                    const float k = thp * Dims<D>::green_eval(R_B, r) / alpha / pdf;
                                    for (int ch = 0; ch < 3; ++ch)
                                        sol[ch] = -(c[ch] * p.neumann_intensity * k) + sol[ch];
                                }
                            }
                        }
                    }
                }

                // ---- one-step walk (integrator.cu:447-526)
                {
                    dr.begin(WOST_RNG_WALK, 0);
                    float wu1, wu2;
                    draw_dir(wu1, wu2);
                    Vec<D> cur = x, dir;
                    float pdf, alpha = 1.0f;
                    if (onN)
                    {
                        dir = to_world(nN, sample_hemisphere_local(wu1, wu2, (Vec<D> *)nullptr));
                        pdf = Dims<D>::hemi_pdf();
                        alpha = 0.5f;
                        for (int k = 0; k < D; ++k)
                            cur[k] = cur[k] + eps * nN[k];
                    }
                    else
                    {
                        dir = sample_sphere(wu1, wu2, (Vec<D> *)nullptr);
                        pdf = Dims<D>::sphere_pdf();
                    }
                    Vec<D> next, normal;
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
            }
        }
        const float inv = (float)S.samples_per_pixel;
        for (int ch = 0; ch < 3; ++ch)
            sol[ch] = sol[ch] / inv;
    }
}

extern "C"
{
    void orc_solve_uniform(const orc_problem_desc *p, const orc_settings *s, int pixel_begin, int pixel_end,
                           float *solution, uint64_t *steps, int nthreads)
    {
        Problem P;
        P.d = *p;
        P.dir = (AnyScene *)p->dirichlet;
        P.neu = (AnyScene *)p->neumann;
        const int n = pixel_end - pixel_begin;
        std::atomic<uint64_t> total{0};
        parallel_for(n, nthreads, [&](int i)
                     {
            const int pixel = pixel_begin + i;
            float *sol = solution + (size_t)pixel * 3;
            if (p->mask && !p->mask[pixel])
            {
                sol[0] = sol[1] = sol[2] = 0.0f;
                return;
            }
            uint64_t st = 0;
            if (p->dim == 2)
                walk_pixel<2>(P, *s, P.dir ? P.dir->s2 : nullptr, P.neu ? P.neu->s2 : nullptr, pixel, sol, st);
            else
                walk_pixel<3>(P, *s, P.dir ? P.dir->s3 : nullptr, P.neu ? P.neu->s3 : nullptr, pixel, sol, st);
            if (g_pixel_steps)
                g_pixel_steps[pixel] = (uint32_t)st;
            total.fetch_add(st, std::memory_order_relaxed); });
        if (steps)
            *steps = total.load();
    }

    // diagnostics: when set, orc_solve_uniform also writes the walk-steps of every pixel it evaluates (W*H uint32)
    void orc_set_pixel_steps(uint32_t *buf) { g_pixel_steps = buf; }
    // diagnostics: when set, every walk writes its positions [depth][dim] here (call orc_solve_uniform on ONE pixel, 1 spp, 1 thread)
    void orc_set_trace(float *buf) { g_trace = buf; }
    // known-answer probe: setSeed(initstate, initseq), advance(delta), then n raw 32-bit outputs
    void orc_pcg32_probe(uint64_t initstate, uint64_t initseq, int64_t delta, int n, uint32_t *out)
    {
        Pcg32 g;
        g.set_seed(initstate, initseq);
        g.advance(delta);
        for (int i = 0; i < n; ++i)
            out[i] = g.next_uint();
    }

    void orc_eval_point(const orc_problem_desc *p, const orc_settings *s, int pixel, float *out)
    {
        if (p->dim == 2)
        {
            const Vec<2> v = eval_point<2>(*p, s->frame_w, s->frame_h, pixel);
            out[0] = v[0];
            out[1] = v[1];
        }
        else
        {
            const Vec<3> v = eval_point<3>(*p, s->frame_w, s->frame_h, pixel);
            out[0] = v[0];
            out[1] = v[1];
            out[2] = v[2];
        }
    }
}
