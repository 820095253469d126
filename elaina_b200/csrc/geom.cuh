// Device geometry for the SNCH-LBVH queries (sm_100a). IEEE fp32 without FMA contraction: this
// translation-unit family is compiled with -fmad=false so that every result equals the host-compiled
// reference / oracle bit for bit (see DESIGN.md "one arithmetic").
//
// Semantics follow ext/lbvh/include/snch_lbvh (file:line cited per function); the data layout does not:
// primitives and silhouette records are stored inline per leaf in Morton order, as 128-bit-loadable records.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "detmath.h"

namespace wost
{
#define WOST_DEV __device__ __forceinline__
#define WOST_HOSTDEV __host__ __device__ __forceinline__

    constexpr float kPi = 3.14159265358979323846f;
    constexpr float kPi2 = 1.57079632679489661923f;
    constexpr float kEps = 1.1920928955078125e-07f;
    constexpr float kBvhOffset = 1e-3f; // scene.cuh:12
    constexpr uint32_t kLeafFlag = 0x80000000u;
    constexpr uint32_t kInvalid = 0xFFFFFFFFu;

    template <int D>
    struct Vec
    {
        float v[D];
        WOST_HOSTDEV float &operator[](int i) { return v[i]; }
        WOST_HOSTDEV float operator[](int i) const { return v[i]; }
    };
    template <int D>
    WOST_HOSTDEV Vec<D> vsub(const Vec<D> &a, const Vec<D> &b)
    {
        Vec<D> r;
#pragma unroll
        for (int i = 0; i < D; ++i)
            r[i] = a[i] - b[i];
        return r;
    }
    WOST_HOSTDEV float vdot(const Vec<2> &a, const Vec<2> &b) { return a[0] * b[0] + a[1] * b[1]; }
    WOST_HOSTDEV float vdot(const Vec<3> &a, const Vec<3> &b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
    template <int D>
    WOST_HOSTDEV float vlen(const Vec<D> &a) { return dm_sqrt(vdot(a, a)); }
    template <int D>
    WOST_HOSTDEV Vec<D> vnormalize(const Vec<D> &a)
    {
        const float n = vlen(a);
        Vec<D> r;
#pragma unroll
        for (int i = 0; i < D; ++i)
            r[i] = dm_div(a[i], n);
        return r;
    }
    WOST_HOSTDEV Vec<3> vcross(const Vec<3> &u, const Vec<3> &v)
    {
        Vec<3> r;
        r[0] = u[1] * v[2] - u[2] * v[1];
        r[1] = u[2] * v[0] - u[0] * v[2];
        r[2] = u[0] * v[1] - u[1] * v[0];
        return r;
    }
    template <int D>
    WOST_HOSTDEV Vec<D> vabs(const Vec<D> &a)
    {
        Vec<D> r;
#pragma unroll
        for (int i = 0; i < D; ++i)
            r[i] = dm_abs(a[i]);
        return r;
    }

    template <int D>
    struct Box
    {
        Vec<D> up, lo;
    };
    template <int D>
    struct Cone
    {
        Vec<D> axis;
        float half_angle, radius;
    };
    struct Node // canonical (reference) layout, core/bvh.cuh:27-33
    {
        uint32_t parent, left, right, object;
    };

    template <int D>
    WOST_HOSTDEV Vec<D> box_centroid(const Box<D> &b) // aabb.cuh:262-279
    {
        Vec<D> c;
#pragma unroll
        for (int i = 0; i < D; ++i)
            c[i] = (b.up[i] + b.lo[i]) * 0.5f;
        return c;
    }
    template <int D>
    WOST_HOSTDEV Box<D> box_merge(const Box<D> &a, const Box<D> &b) // aabb.cuh:103-124
    {
        Box<D> m;
#pragma unroll
        for (int i = 0; i < D; ++i)
        {
            m.up[i] = dm_fmax_strict(a.up[i], b.up[i]);
            m.lo[i] = dm_fmin_strict(a.lo[i], b.lo[i]);
        }
        return m;
    }
    // For finite (non-NaN) boxes and points -- the only case the traversal sees -- fmin/fmax reduce to compares.
    WOST_HOSTDEV float fmin_nn(float a, float b) { return (b < a) ? b : a; }
    WOST_HOSTDEV float fmax_nn(float a, float b) { return (a < b) ? b : a; }

    template <int D>
    WOST_HOSTDEV float box_mindist(const Box<D> &b, const Vec<D> &p) // aabb.cuh:130-150
    {
        float s = 0.0f;
#pragma unroll
        for (int i = 0; i < D; ++i)
        {
            const float d = dm_fmin(b.up[i], dm_fmax(b.lo[i], p[i])) - p[i];
            s = (i == 0) ? d * d : s + d * d;
        }
        return s;
    }
    template <int D>
    WOST_HOSTDEV float box_minmaxdist(const Box<D> &b, const Vec<D> &p) // aabb.cuh:160-209
    {
        float rm[D], rM[D];
#pragma unroll
        for (int i = 0; i < D; ++i)
        {
            const float a = (b.lo[i] - p[i]) * (b.lo[i] - p[i]);
            const float c = (b.up[i] - p[i]) * (b.up[i] - p[i]);
            const bool sw = (b.up[i] + b.lo[i]) * 0.5f < p[i];
            rm[i] = sw ? c : a;
            rM[i] = sw ? a : c;
        }
        if (D == 2)
        {
            const float dx = rm[0] + rM[1];
            const float dy = rM[0] + rm[1];
            return dm_fmin(dx, dy);
        }
        else
        {
            const float dx = rm[0] + rM[1] + rM[D - 1];
            const float dy = rM[0] + rm[1] + rM[D - 1];
            const float dz = rM[0] + rM[1] + rm[D - 1];
            return dm_fmin(dx, dm_fmin(dy, dz));
        }
    }

    // ------------------------------------------------------------------ primitives
    // scene.cuh:176-201 / 230-255
    // *_d2 variants return the dot product the reference feeds to sqrt (length()), so that callers can defer the sqrt.
    template <int D>
    WOST_HOSTDEV float vlen2(const Vec<D> &a) { return vdot(a, a); }
    template <int D>
    WOST_HOSTDEV float closest_point_segment_d2(const Vec<D> &pa, const Vec<D> &pb, const Vec<D> &x, Vec<D> *pt)
    {
        const Vec<D> u = vsub(pb, pa), v = vsub(x, pa);
        const float c1 = vdot(u, v);
        if (c1 <= 0.0f)
        {
            *pt = pa;
            return vlen2(vsub(x, pa));
        }
        const float c2 = vdot(u, u);
        if (c2 <= c1)
        {
            *pt = pb;
            return vlen2(vsub(x, pb));
        }
        const float t = dm_div(c1, c2);
        Vec<D> q;
#pragma unroll
        for (int k = 0; k < D; ++k)
            q[k] = pa[k] + u[k] * t;
        *pt = q;
        return vlen2(vsub(x, q));
    }
    template <int D>
    WOST_HOSTDEV float closest_point_segment(const Vec<D> &pa, const Vec<D> &pb, const Vec<D> &x, Vec<D> *pt)
    {
        return dm_sqrt(closest_point_segment_d2(pa, pb, x, pt));
    }
    // scene.cuh:34-110
    WOST_HOSTDEV float closest_point_triangle_d2(const Vec<3> &pa, const Vec<3> &pb, const Vec<3> &pc, const Vec<3> &x)
    {
        const Vec<3> ab = vsub(pb, pa), ac = vsub(pc, pa), ax = vsub(x, pa);
        const float d1 = vdot(ab, ax), d2 = vdot(ac, ax);
        if (d1 <= 0.0f && d2 <= 0.0f)
            return vlen2(vsub(x, pa));
        const Vec<3> bx = vsub(x, pb);
        const float d3 = vdot(ab, bx), d4 = vdot(ac, bx);
        if (d3 >= 0.0f && d4 <= d3)
            return vlen2(vsub(x, pb));
        const Vec<3> cx = vsub(x, pc);
        const float d5 = vdot(ab, cx), d6 = vdot(ac, cx);
        if (d6 >= 0.0f && d5 <= d6)
            return vlen2(vsub(x, pc));
        Vec<3> pt;
        const float vc = d1 * d4 - d3 * d2;
        if (vc <= 0.0f && d1 >= 0.0f && d3 <= 0.0f)
        {
            const float v = dm_div(d1, d1 - d3);
            pt[0] = pa[0] + ab[0] * v, pt[1] = pa[1] + ab[1] * v, pt[2] = pa[2] + ab[2] * v;
            return vlen2(vsub(x, pt));
        }
        const float vb = d5 * d2 - d1 * d6;
        if (vb <= 0.0f && d2 >= 0.0f && d6 <= 0.0f)
        {
            const float w = dm_div(d2, d2 - d6);
            pt[0] = pa[0] + ac[0] * w, pt[1] = pa[1] + ac[1] * w, pt[2] = pa[2] + ac[2] * w;
            return vlen2(vsub(x, pt));
        }
        const float va = d3 * d6 - d5 * d4;
        if (va <= 0.0f && (d4 - d3) >= 0.0f && (d5 - d6) >= 0.0f)
        {
            const float w = dm_div(d4 - d3, (d4 - d3) + (d5 - d6));
            pt[0] = pb[0] + (pc[0] - pb[0]) * w, pt[1] = pb[1] + (pc[1] - pb[1]) * w, pt[2] = pb[2] + (pc[2] - pb[2]) * w;
            return vlen2(vsub(x, pt));
        }
        const float denom = dm_div(1.0f, va + vb + vc);
        const float v = vb * denom, w = vc * denom;
        pt[0] = pa[0] + ab[0] * v + ac[0] * w;
        pt[1] = pa[1] + ab[1] * v + ac[1] * w;
        pt[2] = pa[2] + ab[2] * v + ac[2] * w;
        return vlen2(vsub(x, pt));
    }

    WOST_HOSTDEV float closest_point_triangle(const Vec<3> &pa, const Vec<3> &pb, const Vec<3> &pc, const Vec<3> &x)
    {
        return dm_sqrt(closest_point_triangle_d2(pa, pb, pc, x));
    }

    // Primitive with its vertices resolved (what a leaf stores inline)
    template <int D>
    struct Prim
    {
        Vec<D> p[D];
    };
    WOST_HOSTDEV float prim_distance(const Prim<2> &pr, const Vec<2> &x) // scene.cuh:504-515
    {
        Vec<2> pt;
        return closest_point_segment(pr.p[0], pr.p[1], x, &pt);
    }
    WOST_HOSTDEV float prim_distance(const Prim<3> &pr, const Vec<3> &x) // scene.cuh:963-976
    {
        return closest_point_triangle(pr.p[0], pr.p[1], pr.p[2], x);
    }
    WOST_HOSTDEV float prim_distance_d2(const Prim<2> &pr, const Vec<2> &x)
    {
        Vec<2> pt;
        return closest_point_segment_d2(pr.p[0], pr.p[1], x, &pt);
    }
    WOST_HOSTDEV float prim_distance_d2(const Prim<3> &pr, const Vec<3> &x)
    {
        return closest_point_triangle_d2(pr.p[0], pr.p[1], pr.p[2], x);
    }

    // core/utility.cuh:569-635
    WOST_HOSTDEV int point_side(const Prim<2> &pr, const Vec<2> &x)
    {
        const float dx = pr.p[1][0] - pr.p[0][0], dy = pr.p[1][1] - pr.p[0][1];
        const float nx = -dy, ny = dx;
        const float vx = x[0] - pr.p[0][0], vy = x[1] - pr.p[0][1];
        const float dp = vx * nx + vy * ny;
        return dp > 0 ? 1 : (dp < 0 ? -1 : 0);
    }
    WOST_HOSTDEV int point_side(const Prim<3> &pr, const Vec<3> &x)
    {
        const Vec<3> n = vcross(vsub(pr.p[1], pr.p[0]), vsub(pr.p[2], pr.p[0]));
        const Vec<3> v = vsub(x, pr.p[0]);
        const float dp = v[0] * n[0] + v[1] * n[1] + v[2] * n[2];
        return dp > 0 ? 1 : (dp < 0 ? -1 : 0);
    }
    // core/utility.cuh:637-708
    WOST_HOSTDEV void proj_ratio(const Prim<2> &pr, const Vec<2> &x, float uv[2])
    {
        const float dx = pr.p[1][0] - pr.p[0][0], dy = pr.p[1][1] - pr.p[0][1];
        const float l2 = dx * dx + dy * dy;
        uv[1] = 0.0f;
        if (l2 < 1e-8f)
        {
            uv[0] = 0.0f;
            return;
        }
        const float vx = x[0] - pr.p[0][0], vy = x[1] - pr.p[0][1];
        uv[0] = dm_div(vx * dx + vy * dy, l2);
    }
    WOST_HOSTDEV void proj_ratio(const Prim<3> &pr, const Vec<3> &x, float uv[2])
    {
        const Vec<3> e1 = vsub(pr.p[1], pr.p[0]), e2 = vsub(pr.p[2], pr.p[0]), v = vsub(x, pr.p[0]);
        const float a = vdot(e1, e1), b = vdot(e1, e2), c = vdot(e2, e2), d = vdot(v, e1), e = vdot(v, e2);
        const float den = a * c - b * b;
        uv[0] = dm_div(d * c - e * b, den);
        uv[1] = dm_div(e * a - d * b, den);
    }
    WOST_HOSTDEV Vec<2> prim_normal(const Prim<2> &pr) // scene.cuh:400-406
    {
        const Vec<2> ls = vsub(pr.p[1], pr.p[0]);
        Vec<2> n;
        n[0] = -ls[1], n[1] = ls[0];
        return vnormalize(n);
    }
    WOST_HOSTDEV Vec<3> prim_normal(const Prim<3> &pr) // scene.cuh:842-852
    {
        return vnormalize(vcross(vsub(pr.p[2], pr.p[0]), vsub(pr.p[1], pr.p[0])));
    }
    WOST_HOSTDEV float prim_measure(const Prim<2> &pr) { return vlen(vsub(pr.p[0], pr.p[1])); } // scene.cuh:409-418
    WOST_HOSTDEV float prim_measure(const Prim<3> &pr)                                            // scene.cuh:855-868
    {
        return dm_div(vlen(vcross(vsub(pr.p[2], pr.p[0]), vsub(pr.p[1], pr.p[0]))), 2.0f);
    }
    WOST_HOSTDEV Vec<2> sample_on_prim(const Prim<2> &pr, float u, float) // scene.cuh:29-32
    {
        Vec<2> r;
        r[0] = pr.p[0][0] + u * (pr.p[1][0] - pr.p[0][0]);
        r[1] = pr.p[0][1] + u * (pr.p[1][1] - pr.p[0][1]);
        return r;
    }
    WOST_HOSTDEV Vec<3> sample_on_prim(const Prim<3> &pr, float u, float v) // scene.cuh:14-27
    {
        if (u + v > 1.0f)
        {
            u = 1.0f - u;
            v = 1.0f - v;
        }
        const float w = 1.0f - u - v;
        Vec<3> r;
#pragma unroll
        for (int k = 0; k < 3; ++k)
            r[k] = w * pr.p[0][k] + u * pr.p[1][k] + v * pr.p[2][k];
        return r;
    }

    // ------------------------------------------------------------------ rays
    template <int D>
    struct Ray // aabb.cuh:298-313
    {
        Vec<D> o, d, inv;
        WOST_HOSTDEV Ray(const Vec<D> &o_, const Vec<D> &d_) : o(o_), d(d_)
        {
#pragma unroll
            for (int k = 0; k < D; ++k)
                inv[k] = dm_div(1.0f, d[k]);
        }
    };
    template <int D>
    WOST_HOSTDEV bool ray_box(const Ray<D> &r, const Box<D> &b, float max_dist, float *dist) // aabb.cuh:366-431
    {
        float tmin = 0.0f, tmax = 0.0f;
#pragma unroll
        for (int k = 0; k < D; ++k)
        {
            const float t1 = (b.lo[k] - r.o[k]) * r.inv[k];
            const float t2 = (b.up[k] - r.o[k]) * r.inv[k];
            if (k == 0)
            {
                tmin = dm_fmin(t1, t2);
                tmax = dm_fmax(t1, t2);
            }
            else
            {
                tmin = dm_fmax(tmin, dm_fmin(t1, t2));
                tmax = dm_fmin(tmax, dm_fmax(t1, t2));
            }
        }
        if (tmax >= tmin && tmax >= 0.0f && tmin <= max_dist)
        {
            *dist = tmin >= 0.0f ? tmin : 0.0f;
            return true;
        }
        return false;
    }
    struct Hit
    {
        bool hit;
        float t, u, v;
    };
    // scene.cuh:543-577. The reference's device branch uses __frcp_rn(D), which is the correctly rounded
    // reciprocal == the host branch's 1.0f / D: one arithmetic on both sides.
    WOST_HOSTDEV Hit ray_prim(const Prim<2> &pr, const Ray<2> &r)
    {
        Hit h;
        h.hit = false, h.t = 0.0f, h.u = 0.0f, h.v = 0.0f;
        const Vec<2> sd = vsub(pr.p[1], pr.p[0]);
        const float Dn = r.d[0] * (-sd[1]) + r.d[1] * sd[0];
        if (dm_abs(Dn) < kEps)
            return h;
        const float inv = dm_div(1.0f, Dn);
        const float t = ((pr.p[0][0] - r.o[0]) * (-sd[1]) - (pr.p[0][1] - r.o[1]) * (-sd[0])) * inv;
        const float s = (r.d[0] * (pr.p[0][1] - r.o[1]) - r.d[1] * (pr.p[0][0] - r.o[0])) * inv;
        if (s >= -1e-3f && s <= 1.0f + 1e-3f && t >= 0.0f)
        {
            h.hit = true;
            h.t = t;
            h.u = s;
        }
        return h;
    }
    // scene.cuh:1005-1052
    WOST_HOSTDEV Hit ray_prim(const Prim<3> &pr, const Ray<3> &r)
    {
        Hit h;
        h.hit = false, h.t = 0.0f, h.u = 0.0f, h.v = 0.0f;
        const Vec<3> e1 = vsub(pr.p[1], pr.p[0]), e2 = vsub(pr.p[2], pr.p[0]);
        Vec<3> hh;
        hh[0] = r.d[1] * e2[2] - r.d[2] * e2[1];
        hh[1] = r.d[2] * e2[0] - r.d[0] * e2[2];
        hh[2] = r.d[0] * e2[1] - r.d[1] * e2[0];
        const float det = e1[0] * hh[0] + e1[1] * hh[1] + e1[2] * hh[2];
        if (dm_abs(det) < kEps)
            return h;
        const float inv = dm_div(1.0f, det);
        const Vec<3> sv = vsub(r.o, pr.p[0]);
        const float u = (sv[0] * hh[0] + sv[1] * hh[1] + sv[2] * hh[2]) * inv;
        if (u < 0.0f || u > 1.0f)
            return h;
        Vec<3> q;
        q[0] = sv[1] * e1[2] - sv[2] * e1[1];
        q[1] = sv[2] * e1[0] - sv[0] * e1[2];
        q[2] = sv[0] * e1[1] - sv[1] * e1[0];
        const float v = (r.d[0] * q[0] + r.d[1] * q[1] + r.d[2] * q[2]) * inv;
        if (v < 0.0f || u + v > 1.0f)
            return h;
        const float t = (e2[0] * q[0] + e2[1] * q[1] + e2[2] * q[2]) * inv;
        if (t >= 0.0f)
        {
            h.hit = true;
            h.t = t;
            h.u = u;
            h.v = v;
        }
        return h;
    }

    // ------------------------------------------------------------------ spheres (boundary sampling)
    template <int D>
    WOST_HOSTDEV bool sphere_box(const Vec<D> &c, float radius, const Box<D> &b) // aabb.cuh:433-449
    {
        float d2 = 0.0f;
#pragma unroll
        for (int k = 0; k < D; ++k)
        {
            const float cl = dm_max(b.lo[k], dm_min(c[k], b.up[k]));
            const float d = cl - c[k];
            d2 = (k == 0) ? d * d : d2 + d * d;
        }
        return d2 <= radius * radius;
    }
    WOST_HOSTDEV bool sphere_prim(const Prim<2> &pr, const Vec<2> &c, float radius) // scene.cuh:579-603
    {
        const Vec<2> d = vsub(pr.p[1], pr.p[0]);
        const float len2 = d[0] * d[0] + d[1] * d[1];
        float t = dm_div((c[0] - pr.p[0][0]) * d[0] + (c[1] - pr.p[0][1]) * d[1], len2);
        t = dm_max(0.0f, dm_min(1.0f, t));
        const float cx = pr.p[0][0] + t * d[0], cy = pr.p[0][1] + t * d[1];
        const float dx = cx - c[0], dy = cy - c[1];
        return dx * dx + dy * dy <= radius * radius;
    }
    WOST_HOSTDEV bool sphere_prim(const Prim<3> &pr, const Vec<3> &c, float radius) // scene.cuh:1054-1117
    {
        const Vec<3> &p1 = pr.p[0], &p2 = pr.p[1], &p3 = pr.p[2];
        const Vec<3> e1 = vsub(p2, p1), e2 = vsub(p3, p1);
        Vec<3> n = vcross(e1, e2);
        const float nl = dm_sqrt(n[0] * n[0] + n[1] * n[1] + n[2] * n[2]);
        n[0] = dm_div(n[0], nl), n[1] = dm_div(n[1], nl), n[2] = dm_div(n[2], nl);
        const float dd = n[0] * p1[0] + n[1] * p1[1] + n[2] * p1[2];
        const float dp = n[0] * c[0] + n[1] * c[1] + n[2] * c[2] - dd;
        Vec<3> proj;
        proj[0] = c[0] - dp * n[0], proj[1] = c[1] - dp * n[1], proj[2] = c[2] - dp * n[2];
        const Vec<3> v2 = vsub(proj, p1);
        const float d00 = vdot(e2, e2), d01 = vdot(e2, e1), d02 = vdot(e2, v2), d11 = vdot(e1, e1), d12 = vdot(e1, v2);
        const float invd = dm_div(1.0f, d00 * d11 - d01 * d01);
        const float u = (d11 * d02 - d01 * d12) * invd;
        const float v = (d00 * d12 - d01 * d02) * invd;
        if (u >= 0 && v >= 0 && u + v <= 1)
            return dm_abs(dp) <= radius;
        Vec<3> cp = proj;
        if (u < 0)
            cp = p1;
        else if (v < 0)
            cp = p3;
        else if (u + v > 1)
            cp = p2;
        const float dx = cp[0] - c[0], dy = cp[1] - c[1], dz = cp[2] - c[2];
        return dx * dx + dy * dy + dz * dz <= radius * radius;
    }
    WOST_HOSTDEV float green_weight(const Vec<2> &x, const Vec<2> &y) // scene.cuh:605-612
    {
        const float r = dm_max(vlen(vsub(x, y)), 1e-2f);
        return dm_abs(dm_div(dm_log(r), kPi * 2.0f));
    }
    WOST_HOSTDEV float green_weight(const Vec<3> &x, const Vec<3> &y) // scene.cuh:1119-1126
    {
        const float r = dm_max(vlen(vsub(x, y)), 1e-4f);
        return dm_div(1.0f, kPi * 4.0f * r);
    }

    // ------------------------------------------------------------------ silhouettes
    // One silhouette element as a leaf stores it: position(s), the two unit face normals (precomputed with the
    // very arithmetic of scene.cuh:326-336 / 747-772) and which faces exist.
    template <int D>
    struct SilRec;
    template <>
    struct SilRec<2>
    {
        Vec<2> p;
        Vec<2> n0, n1;
        uint32_t flags; // bit0: valid, bit1: has_face(0), bit2: has_face(1)
    };
    template <>
    struct SilRec<3>
    {
        Vec<3> pa, pb;
        Vec<3> n0, n1;
        uint32_t flags;
    };

    // scene.cuh:112-141
    WOST_HOSTDEV bool is_silhouette_vertex(const Vec<2> &n0, const Vec<2> &n1, const Vec<2> &view, float d, bool flip)
    {
        const float prec = 1e-3f;
        const float sign = flip ? 1.0f : -1.0f;
        if (d <= prec)
        {
            const float det = n0[0] * n1[1] - n0[1] * n1[0];
            return sign * det > prec;
        }
        Vec<2> vu;
        vu[0] = dm_div(view[0], d), vu[1] = dm_div(view[1], d);
        const float dot0 = vdot(vu, n0), dot1 = vdot(vu, n1);
        if (dm_abs(dot0) <= prec)
            return sign * dot1 > prec;
        if (dm_abs(dot1) <= prec)
            return sign * dot0 > prec;
        return dot0 * dot1 < 0.0f;
    }
    // scene.cuh:143-174
    WOST_HOSTDEV bool is_silhouette_edge(const Vec<3> &pa, const Vec<3> &pb, const Vec<3> &n0, const Vec<3> &n1,
                                         const Vec<3> &view, float d, bool flip)
    {
        const float prec = 1e-3f;
        const float sign = flip ? 1.0f : -1.0f;
        if (d <= prec)
        {
            const Vec<3> ed = vnormalize(vsub(pb, pa));
            const float ang = dm_atan2(vdot(ed, vcross(n0, n1)), vdot(n0, n1));
            return sign * ang > prec;
        }
        const float dot0 = vdot(view, n0), dot1 = vdot(view, n1);
        if (dm_abs(dot0) <= prec)
            return sign * dot1 > prec;
        if (dm_abs(dot1) <= prec)
            return sign * dot0 > prec;
        return dot0 * dot1 < 0.0f;
    }
    // scene.cuh:354-386
    WOST_HOSTDEV bool sil_closest(const SilRec<2> &s, const Vec<2> &o, float max_r2, float &dist, bool flip)
    {
        if (0.0f >= max_r2)
            return false;
        const Vec<2> view = vsub(o, s.p);
        const float d = vlen(view);
        if (d * d > max_r2)
            return false;
        bool is_sil = (s.flags & 6u) != 6u;
        if (!is_sil)
            is_sil = is_silhouette_vertex(s.n0, s.n1, view, d, flip);
        if (is_sil && d * d <= max_r2)
        {
            dist = d;
            return true;
        }
        return false;
    }
    // scene.cuh:788-824
    WOST_HOSTDEV bool sil_closest(const SilRec<3> &s, const Vec<3> &o, float max_r2, float &dist, bool flip)
    {
        if (0.0f >= max_r2)
            return false;
        Vec<3> cp;
        const float d = closest_point_segment(s.pa, s.pb, o, &cp);
        if (d * d > max_r2)
            return false;
        bool is_sil = (s.flags & 6u) != 6u;
        if (!is_sil)
            is_sil = is_silhouette_edge(s.pa, s.pb, s.n0, s.n1, vsub(o, cp), d, flip);
        if (is_sil && d * d <= max_r2)
        {
            dist = d;
            return true;
        }
        return false;
    }

    // core/cone.cuh:24-66
    WOST_HOSTDEV float project_to_plane(const Vec<2> &n, const Vec<2> &e)
    {
        Vec<2> b;
        b[0] = -n[1], b[1] = n[0];
        return dm_abs(vdot(e, vabs(b)));
    }
    WOST_HOSTDEV float project_to_plane(const Vec<3> &n, const Vec<3> &e)
    {
        const float sign = dm_copysign(1.0f, n[2]);
        const float a = dm_div(-1.0f, sign + n[2]);
        const float b = n[0] * n[1] * a;
        Vec<3> b1, b2;
        b1[0] = 1.0f + sign * n[0] * n[0] * a, b1[1] = sign * b, b1[2] = -sign * n[0];
        b2[0] = b, b2[1] = sign + n[1] * n[1] * a, b2[2] = -n[1];
        const float r1 = vdot(e, vabs(b1)), r2 = vdot(e, vabs(b2));
        return dm_sqrt(r1 * r1 + r2 * r2);
    }
    WOST_HOSTDEV bool inrange(float v, float lo, float hi) { return v >= lo && v <= hi; }
    // core/cone.cuh:78-121 / 168-212
    template <int D>
    WOST_HOSTDEV bool cone_overlap(const Cone<D> &bc, const Vec<D> &o, const Box<D> &b, float dist_to_box)
    {
        if (bc.half_angle >= kPi2 || dist_to_box < kEps)
            return true;
        const Vec<D> c = box_centroid(b);
        Vec<D> va = vsub(c, o);
        const float l = vlen(va);
#pragma unroll
        for (int k = 0; k < D; ++k)
            va[k] = dm_div(va[k], l);
        const float daa = dm_acos(dm_max(-1.0f, dm_min(1.0f, vdot(bc.axis, va))));
        if (inrange(kPi2, daa - bc.half_angle, daa + bc.half_angle))
            return true;
        if (l > bc.radius)
        {
            const float vh = dm_asin(dm_div(bc.radius, l));
            const float hs = bc.half_angle + vh;
            return hs >= kPi2 ? true : inrange(kPi2, daa - hs, daa + hs);
        }
        const Vec<D> e = vsub(b.up, c);
        float d = vdot(e, vabs(va));
        const float sdist = l - d;
        if (sdist <= 0.0f)
            return true;
        d = project_to_plane(va, e);
        const float vh = dm_atan2(d, sdist);
        const float hs = bc.half_angle + vh;
        return hs >= kPi2 ? true : inrange(kPi2, daa - hs, daa + hs);
    }

    // core/cone.cuh:266-275, 288-302
    WOST_HOSTDEV Vec<2> cone_rotate(const Vec<2> &u, const Vec<2> &v, float theta)
    {
        const float det = u[0] * v[1] - u[1] * v[0];
        theta *= dm_copysign(1.0f, det);
        float s, c;
        dm_sincos(theta, &s, &c);
        Vec<2> r;
        r[0] = c * u[0] - s * u[1];
        r[1] = s * u[0] + c * u[1];
        return r;
    }
    WOST_HOSTDEV Vec<3> cone_rotate(const Vec<3> &u, const Vec<3> &v, float theta)
    {
        float s, c;
        dm_sincos(theta, &s, &c);
        const Vec<3> w = vnormalize(vcross(u, v));
        Vec<3> o;
        o[0] = (1.0f - c) * w[0], o[1] = (1.0f - c) * w[1], o[2] = (1.0f - c) * w[2];
        const float R00 = c + o[0] * w[0], R01 = o[1] * w[0] - s * w[2], R02 = o[2] * w[0] + s * w[1];
        const float R10 = o[0] * w[1] + s * w[2], R11 = c + o[1] * w[1], R12 = o[2] * w[1] - s * w[0];
        const float R20 = o[0] * w[2] - s * w[1], R21 = o[1] * w[2] + s * w[0], R22 = c + o[2] * w[2];
        Vec<3> r;
        r[0] = R00 * u[0] + R01 * u[1] + R02 * u[2];
        r[1] = R10 * u[0] + R11 * u[1] + R12 * u[2];
        r[2] = R20 * u[0] + R21 * u[1] + R22 * u[2];
        return r;
    }
    // core/cone.cuh:320-372 / 427-480; half_angle = pi where the reference leaves it uninitialised (:347-352)
    template <int D>
    WOST_HOSTDEV Cone<D> cone_merge(const Cone<D> &A, const Cone<D> &B, const Vec<D> &oa, const Vec<D> &ob, const Vec<D> &on)
    {
        Cone<D> r;
#pragma unroll
        for (int k = 0; k < D; ++k)
            r.axis[k] = 0.0f;
        r.radius = 0.0f;
        r.half_angle = kPi;
        const bool va = A.half_angle >= 0.0f, vb = B.half_angle >= 0.0f;
        if (va && vb)
        {
            Vec<D> axa = A.axis, axb = B.axis;
            float ha = A.half_angle, hb = B.half_angle;
            const Vec<D> da = vsub(on, oa), db = vsub(on, ob);
            r.radius = dm_sqrt(dm_max(A.radius * A.radius + vdot(da, da), B.radius * B.radius + vdot(db, db)));
            if (hb > ha)
            {
                const Vec<D> t = axa;
                axa = axb;
                axb = t;
                const float th = ha;
                ha = hb;
                hb = th;
            }
            const float theta = dm_acos(dm_max(-1.0f, dm_min(1.0f, vdot(axa, axb))));
            if (dm_min(theta + hb, kPi) <= ha)
            {
                r.axis = axa;
                r.half_angle = ha;
                return r;
            }
            const float ot = dm_div(ha + theta + hb, 2.0f);
            if (ot >= kPi)
            {
                r.axis = axa;
                r.half_angle = kPi;
                return r;
            }
            r.axis = cone_rotate(axa, axb, ot - ha);
            r.half_angle = ot;
        }
        else if (va)
            r = A;
        else if (vb)
            r = B;
        else
            r.half_angle = -kPi;
        return r;
    }
} // namespace wost
