// One walk-on-stars step as a single fused device function (sm_100a).
//
// The reference runs a step as 5-7 separate full-frame kernels that hand work items to each other through
// managed-memory queues (integrator/uniform/integrator.cu:552-575: separate -> handleBoundary -> sampleSource ->
// sampleNeumann -> oneStepWalk). Here the whole step -- closest-point query, epsilon-shell test, Dirichlet read-out,
// silhouette query, star radius, source sample, Neumann boundary sample + visibility ray, direction sample, Neumann
// ray cast and reflection bookkeeping -- runs on register-resident state; only BVH records are fetched from memory.
// Semantics (what is computed, in which order contributions reach the pixel accumulator) follow the reference stage
// by stage; file:line cited at each block. Randomness is the counter-based Philox stream of philox.h.
#pragma once
#include <algorithm>
#include <cstdlib>
#include "lbvh.cuh"
#include "philox.h"
#include "problem.cuh"

namespace wost
{
    template <int D>
    struct WalkState
    {
        Vec<D> x;       // current point
        Vec<D> nN;      // shading normal of the Neumann boundary the point sits on (if onN)
        float thp;      // throughput (a scalar: every factor applied to it is a scalar, oneStepWalk :521)
        uint32_t pixel; // pixelId = x + y*W (keys the Philox stream, locates the evaluation point)
        uint32_t ai;    // rank of the pixel among this shard's active pixels: indexes accumulators / walk records (sized by the shard, not the frame)
        uint32_t sample;
        bool onN;
    };

    // core/evaluation_grid.h:27-33 / 60-64
    template <int D>
    WOST_DEV Vec<D> eval_point(const WalkCtx<D> &c, uint32_t pixel)
    {
        const int px = (int)(pixel % (uint32_t)c.frame_w), py = (int)(pixel / (uint32_t)c.frame_w);
        const float nx = dm_div(2.0f * (float)px, (float)c.frame_w) + (-1.0f);
        const float ny = dm_div(2.0f * (float)py, (float)c.frame_h) + (-1.0f);
        Vec<D> r;
        if (D == 2)
        {
            const float ux = c.grid_up[1], uy = -c.grid_up[0];
            const float vx = c.grid_up[0], vy = c.grid_up[1];
            r[0] = c.grid_scale * (nx * ux + ny * vx) + c.grid_pos[0];
            r[1] = c.grid_scale * (nx * uy + ny * vy) + c.grid_pos[1];
        }
        else
        {
#pragma unroll
            for (int k = 0; k < D; ++k)
                r[k] = c.grid_scale * (nx * c.grid_right[k] + ny * c.grid_up[k]) + c.grid_pos[k];
        }
        return r;
    }

    template <int D>
    WOST_DEV bool uv_strictly_inside(const float uv[2]) // integrator/uniform/integrator.cu:159-168
    {
        if (D == 2)
            return uv[0] > 0.0f && uv[0] < 1.0f;
        return uv[0] > 0.0f && uv[1] > 0.0f && uv[0] + uv[1] < 1.0f;
    }
    // integrator/common.h:242-260 + krrmath/functors.h:60-95
    template <int D>
    WOST_DEV void surface_color(const float *pairs, const int *vidx, int side, const float uv[2], float out[3])
    {
        if (!pairs)
        {
            out[0] = out[1] = out[2] = 0.0f;
            return;
        }
        const int off = side >= 0 ? 0 : 3;
        const float *c0 = pairs + (size_t)vidx[0] * 6 + off;
        const float *c1 = pairs + (size_t)vidx[1] * 6 + off;
        if (D == 2)
        {
#pragma unroll
            for (int ch = 0; ch < 3; ++ch)
                out[ch] = __ldg(c0 + ch) * (1 - uv[0]) + __ldg(c1 + ch) * uv[0];
        }
        else
        {
            const float *c2 = pairs + (size_t)vidx[D - 1] * 6 + off;
            const float w = 1 - uv[0] - uv[1];
#pragma unroll
            for (int ch = 0; ch < 3; ++ch)
                out[ch] = __ldg(c0 + ch) * w + __ldg(c1 + ch) * uv[0] + __ldg(c2 + ch) * uv[1];
        }
    }

    // util/sampling.h:18-33, 68-85
    WOST_DEV Vec<2> sample_sphere(float u1, float, Vec<2> *)
    {
        Vec<2> r;
        dm_sincos(u1 * DM_2PI, &r[1], &r[0]);
        return r;
    }
    WOST_DEV Vec<3> sample_sphere(float u1, float u2, Vec<3> *)
    {
        const float z = 1 - 2 * u1;
        const float rr = dm_sqrt(1 - z * z);
        float s, c;
        dm_sincos(DM_2PI * u2, &s, &c);
        Vec<3> r;
        r[0] = rr * c, r[1] = rr * s, r[2] = z;
        return r;
    }
    WOST_DEV Vec<2> sample_hemisphere_local(float u1, float, Vec<2> *)
    {
        Vec<2> r;
        dm_sincos(DM_PI * u1, &r[1], &r[0]);
        return r;
    }
    WOST_DEV Vec<3> sample_hemisphere_local(float u1, float u2, Vec<3> *)
    {
        const float z = u1;
        const float rr = dm_sqrt(dm_max(0.0f, 1.0f - z * z));
        float s, c;
        dm_sincos(2.0f * DM_PI * u2, &s, &c);
        Vec<3> r;
        r[0] = rr * c, r[1] = rr * s, r[2] = z;
        return r;
    }
    // util/transformation.h:52-67 + util/math_utils.h:141-156
    WOST_DEV Vec<2> to_world(const Vec<2> &n, const Vec<2> &v)
    {
        Vec<2> perp;
        perp[0] = -n[1], perp[1] = n[0];
        const float l2 = perp[0] * perp[0] + perp[1] * perp[1];
        if (l2 > 0.0f)
        {
            const float l = dm_sqrt(l2);
            perp[0] = dm_div(perp[0], l);
            perp[1] = dm_div(perp[1], l);
        }
        Vec<2> r;
        r[0] = -perp[0] * v[0] + n[0] * v[1];
        r[1] = -perp[1] * v[0] + n[1] * v[1];
        return r;
    }
    WOST_DEV Vec<3> safe_normalized(const Vec<3> &a)
    {
        const float l2 = vdot(a, a);
        if (!(l2 > 0.0f))
            return a;
        const float l = dm_sqrt(l2);
        Vec<3> r;
        r[0] = dm_div(a[0], l), r[1] = dm_div(a[1], l), r[2] = dm_div(a[2], l);
        return r;
    }
    WOST_DEV Vec<3> to_world(const Vec<3> &n, const Vec<3> &v)
    {
        const float ax = dm_abs(n[0]), ay = dm_abs(n[1]), az = dm_abs(n[2]);
        const uint32_t uyx = (ax - ay) < 0 ? 1 : 0, uzx = (ax - az) < 0 ? 1 : 0, uzy = (ay - az) < 0 ? 1 : 0;
        const uint32_t xm = uyx & uzx, ym = (1 ^ xm) & uzy, zm = 1 ^ (xm | ym);
        Vec<3> a;
        a[0] = (float)xm, a[1] = (float)ym, a[2] = (float)zm;
        const Vec<3> t = safe_normalized(vcross(n, a));
        const Vec<3> b = safe_normalized(vcross(n, t));
        Vec<3> r;
#pragma unroll
        for (int k = 0; k < 3; ++k)
            r[k] = t[k] * v[0] + b[k] * v[1] + n[k] * v[2];
        return r;
    }

    template <int D>
    struct Dims;
    template <>
    struct Dims<2>
    {
        static WOST_DEV float sphere_pdf() { return 1.0f / DM_2PI; }
        static WOST_DEV float hemi_pdf() { return 1.0f / DM_PI; }
        static WOST_DEV float sphere_measure() { return DM_2PI; }
        static WOST_DEV float cond_pdf(float dirPdf, float r) { return dm_div(dirPdf, r); }
        static WOST_DEV float green_eval(float R, float r) { return dm_div(dm_log(dm_div(R, r)), DM_2PI); } // util/green.h:25-28
        static WOST_DEV float green_norm(float R) { return dm_div(R * R, 4.0f); }
    };
    template <>
    struct Dims<3>
    {
        static WOST_DEV float sphere_pdf() { return 1.0f / DM_4PI; }
        static WOST_DEV float hemi_pdf() { return 1.0f / DM_2PI; }
        static WOST_DEV float sphere_measure() { return DM_4PI; }
        static WOST_DEV float cond_pdf(float dirPdf, float r) { return dm_div(dirPdf, r * r); }
        static WOST_DEV float green_eval(float R, float r) { return dm_div(dm_div(1.0f, r) - dm_div(1.0f, R), DM_4PI); } // util/green.h:84-91
        static WOST_DEV float green_norm(float R) { return dm_div(R * R, 6.0f); }
    };

    // util/green.h:45-74 (2D, bounded rejection) / :98-113 (3D, closed form)
    WOST_DEV float green_sample_radius(float R, uint32_t seed, uint32_t pixel, uint32_t sample, uint32_t depth, const Philox4 &, Vec<2> *)
    {
        const float bound = dm_div(1.5f, R);
        float r = 0.0f;
        int iter = 0;
        do
        {
            const Philox4 q = wost_rng(seed, pixel, sample, depth, WOST_RNG_SOURCE_RAD, (uint32_t)iter);
            const float u = u32_to_unit_float(q.x);
            r = u32_to_unit_float(q.y) * R;
            const float pdf = dm_div(Dims<2>::green_eval(R, r), Dims<2>::green_norm(R));
            const float pdfRadius = dm_div(pdf, Dims<2>::sphere_pdf());
            ++iter;
            if (u < dm_div(pdfRadius, bound))
                break;
        } while (iter < 1000);
        r = dm_max(1e-4f, r);
        if (r > R)
            r = dm_div(R, 2.0f);
        return r;
    }
    WOST_DEV float green_sample_radius(float R, uint32_t, uint32_t, uint32_t, uint32_t, const Philox4 &dirdraw, Vec<3> *)
    {
        const float u1 = u32_to_unit_float(dirdraw.z), u2 = u32_to_unit_float(dirdraw.w);
        const float phi = DM_2PI * u2;
        float r = dm_div((1.0f + dm_sqrt(1.0f - dm_cbrt(u1 * u1)) * dm_cos(phi)) * R, 2.0f);
        r = dm_max(1e-4f, r);
        if (r > R)
            r = dm_div(R, 2.0f);
        return r;
    }

    // dense Vec3f grid, trilinear like NanoVDB SampleFromVoxels<Tree,1,false> (SampleFromVoxels.h:202-243,413-419)
    template <int D>
    WOST_DEV void source_lookup(const WalkCtx<D> &c, const float w3[3], float out[3])
    {
        float f[3];
        int ijk[3];
#pragma unroll
        for (int k = 0; k < 3; ++k)
        {
            const float x = (w3[k] - c.source_origin[k]) * c.source_inv_voxel;
            const float fl = floorf(x);
            // clamp before the conversion: walks that escape an exterior problem reach |x| ~ 1e19, where (int)fl saturates to
            // INT_MAX and `ijk + 1` wraps (signed overflow: the bounds test below may then be folded into `true`)
            ijk[k] = (int)fminf(fmaxf(fl, -2.0f), 1.0e9f);
            f[k] = x - fl;
        }
#pragma unroll
        for (int ch = 0; ch < 3; ++ch)
        {
            float v[2][2][2];
#pragma unroll
            for (int a = 0; a < 2; ++a)
#pragma unroll
                for (int b = 0; b < 2; ++b)
#pragma unroll
                    for (int d = 0; d < 2; ++d)
                    {
                        const int i = ijk[0] + a, j = ijk[1] + b, k = ijk[2] + d;
                        const bool in = i >= 0 && j >= 0 && k >= 0 && i < c.source_dims[0] && j < c.source_dims[1] && k < c.source_dims[2];
                        v[a][b][d] = in ? __ldg(c.source + (((size_t)k * c.source_dims[1] + j) * c.source_dims[0] + i) * 3 + ch) : 0.0f;
                    }
            const float z00 = v[0][0][0] + f[2] * (v[0][0][1] - v[0][0][0]);
            const float z01 = v[0][1][0] + f[2] * (v[0][1][1] - v[0][1][0]);
            const float z10 = v[1][0][0] + f[2] * (v[1][0][1] - v[1][0][0]);
            const float z11 = v[1][1][0] + f[2] * (v[1][1][1] - v[1][1][0]);
            const float y0 = z00 + f[1] * (z01 - z00);
            const float y1 = z10 + f[1] * (z11 - z10);
            out[ch] = y0 + f[0] * (y1 - y0);
        }
    }

    // Pixel accumulator handled like the reference's addColor (integrator/uniform/workqueue.h:25-29):
    // loaded on first use, every contribution applied in stage order, stored once at the end of the step.
    // TrackTotal: also keep the running sum of the walk's own contributions (tot) -- the guided integrator derives
    // its training targets from it (GuidedPixelStateBuffer::recordSolution / recordSourceContribution, guided.h:46-67).
    template <bool TrackTotal>
    struct PixelAccT
    {
        float *ptr;
        float v[3];
        float tot[3];
        bool loaded;
        WOST_DEV void init(float *p)
        {
            ptr = p;
            loaded = false;
        }
        WOST_DEV void add(float c0, float c1, float c2)
        {
            if (TrackTotal)
            {
                tot[0] += c0;
                tot[1] += c1;
                tot[2] += c2;
            }
            if (!loaded)
            {
                v[0] = ptr[0];
                v[1] = ptr[1];
                v[2] = ptr[2];
                loaded = true;
            }
            v[0] = c0 + v[0];
            v[1] = c1 + v[1];
            v[2] = c2 + v[2];
        }
        WOST_DEV void flush()
        {
            if (loaded)
            {
                ptr[0] = v[0];
                ptr[1] = v[1];
                ptr[2] = v[2];
            }
        }
    };
    using PixelAcc = PixelAccT<false>;

    // Problem features a kernel instantiation is compiled for: the host picks the leanest variant the problem allows
    // (no Neumann mesh -> no silhouette query, no ray casts, no hemisphere sampling; no source grid -> no Green's-function
    // sample), which cuts registers and the instruction footprint of the walk kernels.
    constexpr int kFeatNeumann = 1, kFeatSource = 2, kFeatFlux = 4 /* non-zero Neumann data: boundary-sample stage */, kFeatAll = 7;

    // ---------------------------------------------------------------------------------------------------------
    // One step of the UNIFORM integrator. Returns true when the walk goes on (w holds the next evaluation point).
    // Everything of a walk step that follows the closest-point query on the Dirichlet mesh (R_D, leaf position pos).
    // Stages separate -> boundary -> source sample -> Neumann sample of one step. Returns false when the walk ended
    // (epsilon-shell hit or no star radius); otherwise R_B holds the star radius of the step.
    // rb_scale: 0.99 in the uniform integrator (integrator/uniform/integrator.cu:193-195), 1 in the guided one
    // (integrator/guided/integrator.cu:238-239).
    template <int D, int F = kFeatAll, class Acc>
    WOST_DEV bool walk_step_contrib(const WalkCtx<D> &c, const WalkState<D> &w, uint32_t depth, SmemStack &st, Acc &acc, float R_D, uint32_t pos,
                                    float rb_scale, float &R_B)
    {
        const float eps = c.eps;
        const bool has_neu = (F & kFeatNeumann) && c.neu.nprims > 0, has_src = (F & kFeatSource) && c.source != nullptr;
        const bool onN = (F & kFeatNeumann) && w.onN;
        // ---- separate (integrator/uniform/integrator.cu:102-212)
        if (c.dir.nprims > 0)
        {
            if (R_D < eps)
            {
                const Prim<D> pr = load_prim(c.dir.leaf_prim + pos);
                float uv[2];
                proj_ratio(pr, w.x, uv);
                if (uv_strictly_inside<D>(uv))
                {
                    // ---- boundary (integrator.cu:214-232)
                    const int side = point_side(pr, w.x);
                    const uint32_t obj = __ldg(c.dir.leaf_obj + pos);
                    int vidx[D];
#pragma unroll
                    for (int k = 0; k < D; ++k)
                        vidx[k] = __ldg(c.dir_prims + (size_t)obj * D + k);
                    float col[3];
                    surface_color<D>(c.color_dir, vidx, side, uv, col);
                    acc.add(col[0] * c.dirichlet_intensity * w.thp, col[1] * c.dirichlet_intensity * w.thp, col[2] * c.dirichlet_intensity * w.thp);
                    return false;
                }
            }
        }
        float R_N = dm_inf();
        if (has_neu)
            R_N = query_silhouette(c.neu, w.x, false, st, R_D); // only min(R_D, R_N) is used
        R_B = dm_fmax(1e-4f, dm_fmin(R_D, R_N));
        R_B *= rb_scale;
        if (dm_isinf(R_B))
            return false;

        // ---- source sample (integrator.cu:234-316)
        if (has_src)
        {
            const Philox4 q = wost_rng(c.seed, w.pixel, w.sample, depth, WOST_RNG_SOURCE_DIR, 0);
            Vec<D> dir;
            float dirPdf, alpha = 1.0f;
            if (onN)
            {
                dir = to_world(w.nN, sample_hemisphere_local(u32_to_unit_float(q.x), u32_to_unit_float(q.y), (Vec<D> *)nullptr));
                dirPdf = Dims<D>::hemi_pdf();
                alpha = 0.5f;
            }
            else
            {
                dir = sample_sphere(u32_to_unit_float(q.x), u32_to_unit_float(q.y), (Vec<D> *)nullptr);
                dirPdf = Dims<D>::sphere_pdf();
            }
            float dist = R_B;
            if (has_neu)
            {
                Vec<D> o;
#pragma unroll
                for (int k = 0; k < D; ++k)
                    o[k] = w.x[k] + eps * dir[k];
                const RayResult rr = query_ray<D, false>(c.neu, Ray<D>(o, dir), dist, st);
                if (rr.hit)
                    dist = dm_fmin(rr.t, dist);
            }
            const float r = green_sample_radius(R_B, c.seed, w.pixel, w.sample, depth, q, (Vec<D> *)nullptr);
            if (r <= dist)
            {
                float w3[3];
                w3[0] = w.x[0] + r * dir[0];
                w3[1] = w.x[1] + r * dir[1];
                w3[2] = (D == 3) ? w.x[D - 1] + r * dir[D - 1] : 0.0f;
                float g[3];
                source_lookup(c, w3, g);
                float cc[3];
#pragma unroll
                for (int ch = 0; ch < 3; ++ch)
                {
                    const float gv = g[ch] * c.source_intensity;
                    cc[ch] = dm_div(dm_div(w.thp * gv * Dims<D>::green_norm(R_B) * Dims<D>::cond_pdf(Dims<D>::sphere_pdf(), r), Dims<D>::cond_pdf(dirPdf, r)), alpha);
                }
                acc.add(cc[0], cc[1], cc[2]);
            }
        }

        // ---- Neumann boundary sample (integrator.cu:318-445). Skipped when every Neumann colour is zero:
        // the stage could then only add -0 to the pixel ("pure overhead there").
        if ((F & kFeatFlux) && has_neu && c.neumann_flux_enabled)
        {
            const Philox4 q = wost_rng(c.seed, w.pixel, w.sample, depth, WOST_RNG_NEUMANN, 0);
            const float u0 = u32_to_unit_float(q.x), u1 = u32_to_unit_float(q.y), u2 = u32_to_unit_float(q.z);
            int pos;
            float pdf;
            sample_in_sphere(c.neu, w.x, R_B, u0, pos, pdf);
            if (pos != -1 && pdf > 0)
            {
                const Prim<D> pr = load_prim(c.neu.leaf_prim + pos);
                const Vec<D> y = sample_on_prim(pr, u1, u2);
                const float r = vlen(vsub(y, w.x));
                if (r < R_B && r > 0)
                {
                    Vec<D> o = w.x;
                    if (onN)
                    {
#pragma unroll
                        for (int k = 0; k < D; ++k)
                            o[k] = o[k] + eps * w.nN[k];
                    }
                    Vec<D> rd = vsub(y, o);
                    const float cl = vlen(rd);
                    if (cl > 0)
                    {
#pragma unroll
                        for (int k = 0; k < D; ++k)
                            rd[k] = dm_div(rd[k], cl);
                    }
                    const bool blocked = query_ray<D, true>(c.neu, Ray<D>(o, rd), cl - eps, st).hit;
                    if (!blocked)
                    {
                        int side = point_side(pr, w.x);
                        float uv[2];
                        proj_ratio(pr, y, uv);
                        if (onN)
                        {
                            const float dn = vdot(prim_normal(pr), w.nN);
                            side = (0.0f < dn) - (dn < 0.0f);
                        }
                        if (side != 0)
                        {
                            const uint32_t obj = __ldg(c.neu.leaf_obj + pos);
                            int vidx[D];
#pragma unroll
                            for (int k = 0; k < D; ++k)
                                vidx[k] = __ldg(c.neu_prims + (size_t)obj * D + k);
                            float col[3];
                            surface_color<D>(c.color_neu, vidx, side, uv, col);
                            const float alpha = w.onN ? 0.5f : 1.0f;
                            const float kf = dm_div(dm_div(w.thp * Dims<D>::green_eval(R_B, r), alpha), pdf);
                            acc.add(-(col[0] * c.neumann_intensity * kf), -(col[1] * c.neumann_intensity * kf), -(col[2] * c.neumann_intensity * kf));
                        }
                    }
                }
            }
        }

        return true;
    }
    // uniform direction of oneStepWalk (integrator.cu:447-470): sphere, or hemisphere about the Neumann normal
    template <int D, int F = kFeatAll>
    WOST_DEV Vec<D> uniform_direction(const WalkCtx<D> &c, const WalkState<D> &w, uint32_t depth, float &pdf)
    {
        const bool onN = (F & kFeatNeumann) && w.onN;
        const Philox4 q = wost_rng(c.seed, w.pixel, w.sample, depth, WOST_RNG_WALK, 0);
        if (onN)
        {
            pdf = Dims<D>::hemi_pdf();
            return to_world(w.nN, sample_hemisphere_local(u32_to_unit_float(q.x), u32_to_unit_float(q.y), (Vec<D> *)nullptr));
        }
        pdf = Dims<D>::sphere_pdf();
        return sample_sphere(u32_to_unit_float(q.x), u32_to_unit_float(q.y), (Vec<D> *)nullptr);
    }
    // second half of oneStepWalk (integrator.cu:472-526): move along dir by R_B or to the first Neumann hit, update thp
    template <int D, int F = kFeatAll>
    WOST_DEV void walk_advance(const WalkCtx<D> &c, WalkState<D> &w, const Vec<D> &dir, float pdf, float R_B, SmemStack &st)
    {
        const float eps = c.eps;
        const bool has_neu = (F & kFeatNeumann) && c.neu.nprims > 0, onN = (F & kFeatNeumann) && w.onN;
        {
            Vec<D> cur = w.x;
            float alpha = 1.0f;
            if (onN)
            {
                alpha = 0.5f;
#pragma unroll
                for (int k = 0; k < D; ++k)
                    cur[k] = cur[k] + eps * w.nN[k];
            }
            Vec<D> next, normal;
#pragma unroll
            for (int k = 0; k < D; ++k)
            {
                next[k] = w.x[k] + R_B * dir[k];
                normal[k] = 0.0f;
            }
            bool hit = false;
            if (has_neu)
            {
                const RayResult rr = query_ray<D, false>(c.neu, Ray<D>(cur, dir), R_B, st);
                hit = rr.hit;
                if (hit)
                {
                    normal = prim_normal(load_prim(c.neu.leaf_prim + rr.leaf_pos));
                    if (vdot(normal, dir) > 0)
                    {
#pragma unroll
                        for (int k = 0; k < D; ++k)
                            normal[k] = -normal[k];
                    }
#pragma unroll
                    for (int k = 0; k < D; ++k)
                        next[k] = cur[k] + rr.t * dir[k];
                }
            }
            w.x = next;
            w.thp = dm_div(dm_div(dm_div(w.thp, pdf), alpha), Dims<D>::sphere_measure());
            w.onN = hit;
            w.nN = normal;
        }
    }
    // Everything of a UNIFORM walk step that follows the closest-point query on the Dirichlet mesh (R_D, leaf position pos).
    template <int D, int F = kFeatAll, class Acc>
    WOST_DEV bool walk_step_tail(const WalkCtx<D> &c, WalkState<D> &w, uint32_t depth, SmemStack &st, Acc &acc, float R_D, uint32_t pos, float rb_scale = 0.99f)
    {
        float R_B;
        if (!walk_step_contrib<D, F>(c, w, depth, st, acc, R_D, pos, rb_scale, R_B))
            return false;
        float pdf;
        const Vec<D> dir = uniform_direction<D, F>(c, w, depth, pdf);
        walk_advance<D, F>(c, w, dir, pdf, R_B, st);
        return true;
    }
    // One step of the UNIFORM integrator. Returns true when the walk goes on (w holds the next evaluation point).
    // Warp-synchronous (the closest-point driver is): every lane of the warp calls it, lanes without a walk pass valid = false.
    // Lanes that must be waiting (query finished) before a warp of the persistent kernels runs a tail phase. The tail runs at the
    // width of the waiting lanes, the traversal at the width of the rest: the costlier the tail (silhouette query + ray cast with a
    // Neumann mesh, source lookups and triangle distances in 3D), the more lanes are worth waiting for. Round 1 found 6 for every
    // variant; the oriented boxes made the queries ~3x shorter and moved the optimum (profiles/refill_sweep_r2.json: C3 2 373 M at 6,
    // 2 455 M at 12; c1 1 388 M at 6, 1 759 M at 26, 1 546 M at 32; c4-small 409 / 577 / 487 M). WOST_REFILL_THRESHOLD overrides it (tuning).
    inline int default_refill_threshold(int D, int F)
    {
        static const int env = []
        {
            const char *e = getenv("WOST_REFILL_THRESHOLD");
            return e ? std::max(1, std::min(32, atoi(e))) : 0;
        }();
        if (env > 0)
            return env;
        return (D == 2 && F == 0) ? 12 : 24;
    }
    // Shared-memory levels of the traversal stack in the persistent walk kernels (deeper levels spill to thread-local memory). The
    // stacks of 9 resident CTAs x 20 levels take 180 KB of the SM's 256 KB and leave ~50 KB of L1 for the node records; with the
    // oriented boxes few far siblings are ever pending, so a shorter shared stack trades rare spills for L1 hits on the top levels.
    // WOST_STACK_LEVELS overrides it (tuning).
    inline int walk_stack_levels(int D, int F)
    {
        static const int env = []
        {
            const char *e = getenv("WOST_STACK_LEVELS");
            return e ? std::max(kMinSmemLevels, std::min(kSmemLevels, atoi(e))) : 0;
        }();
        (void)F;
        return env > 0 ? env : (D == 2 ? 12 : kSmemLevels); // sweep (profiles/stack_levels_r2.json): 2D 12 (+1-2 %), 3D the full 20
    }
    // RefOrder: closest-point queries visit the tree in the reference's own order -- same results, kept as the
    // on-device stand-in for the reference's traversal cost (bench.py --walk-mode reforder).
    template <int D, bool RefOrder>
    WOST_DEV bool walk_step_uniform(const WalkCtx<D> &c, WalkState<D> &w, uint32_t depth, SmemStack &st, PixelAcc &acc, bool valid)
    {
        float R_D = dm_inf();
        uint32_t pos = kInvalid;
        if (c.dir.nprims > 0)
            query_nearest_sync<D, RefOrder>(c.dir, w.x, st, pos, R_D, valid, c.eps);
        if (!valid)
            return false;
        return walk_step_tail<D>(c, w, depth, st, acc, R_D, pos);
    }
    // Closest-point query of every active evaluation point, once per solve. All samples of a pixel start their walk at
    // the same point (generateEvaluationPointsImpl, integrator.cu:80-100), and that first query is the most expensive of
    // the walk (deep interior, no a-priori bound): the walk kernels read its result instead of repeating it per sample.
    template <int D, bool RefOrder>
    __global__ void __launch_bounds__(128) k_first_hit(const __grid_constant__ WalkCtx<D> c, const uint32_t *active, int n_active, uint2 *first, int *overflow)
    {
        extern __shared__ uint2 smem_stack[];
        SmemStack st;
        uint2 spill_mem[kSpillLevels];
        st.init(smem_stack, overflow, spill_mem);
        for (int base = blockIdx.x * blockDim.x; base < n_active; base += gridDim.x * blockDim.x)
        {
            const int i = base + threadIdx.x;
            const bool valid = i < n_active;
            const Vec<D> x = eval_point(c, valid ? __ldg(active + i) : 0u);
            NearestTrav<D> t;
            t.win = kInvalid, t.best = dm_inf();
            bool trav = false;
            if (valid)
                trav = nearest_begin<D, RefOrder>(c.dir, x, st, t);
            while (__any_sync(kFullMask, trav))
            {
                if (trav)
                    trav = nearest_trip<D, RefOrder>(c.dir, x, st, t);
            }
            if (valid)
                first[i] = make_uint2(t.win, __float_as_uint(t.best));
        }
    }

} // namespace wost
