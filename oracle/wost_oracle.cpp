// TEST INFRASTRUCTURE ONLY -- scalar CPU oracle for the WoSt walker hot path (see wost_oracle.h).
// Every function cites the reference file:line it restates (paths relative to /root/reference).
// Written from the algorithm, not copied: plain arrays, no Thrust, no templates over iterators.
#include "wost_oracle.h"
#include "detmath.h"
#include "philox.h"

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstring>
#include <map>
#include <thread>
#include <unordered_map>
#include <utility>
#include <vector>

namespace
{
    // ------------------------------------------------------------------ math mode
    // mode 0: detmath (== CUDA product arithmetic); mode 1: libm (== host-compiled reference)
    int g_math_mode = 0;
    // uniform walker draws: 0 = Philox (product), 1 = the reference's per-pixel PCG32 (wost_oracle_walk.inl)
    int g_rng_mode = 0;
    uint32_t *g_pixel_steps = nullptr; // diagnostics (orc_set_pixel_steps)
    float *g_trace = nullptr;          // diagnostics (orc_set_trace)
    inline float m_acos(float x) { return g_math_mode ? acosf(x) : dm_acos(x); }
    inline float m_asin(float x) { return g_math_mode ? asinf(x) : dm_asin(x); }
    inline float m_atan2(float y, float x) { return g_math_mode ? atan2f(y, x) : dm_atan2(y, x); }
    inline float m_sin(float x) { return g_math_mode ? sinf(x) : dm_sin(x); }
    inline float m_cos(float x) { return g_math_mode ? cosf(x) : dm_cos(x); }
    inline float m_log(float x) { return g_math_mode ? logf(x) : dm_log(x); }
    inline float m_cbrt(float x) { return g_math_mode ? cbrtf(x) : dm_cbrt(x); }

    const float PI_F = 3.14159265358979323846f; // static_cast<float>(M_PI)
    const float PI_2_F = 1.57079632679489661923f;
    const float EPS_F = 1.1920928955078125e-07f; // std::numeric_limits<float>::epsilon()
    const float BVH_OFFSET = 1e-3f;               // scene.cuh:12
    const float INF_F = std::numeric_limits<float>::infinity();

    template <int D>
    struct Vec
    {
        float v[D];
        float &operator[](int i) { return v[i]; }
        float operator[](int i) const { return v[i]; }
    };
    template <int D>
    inline Vec<D> sub(const Vec<D> &a, const Vec<D> &b)
    {
        Vec<D> r;
        for (int i = 0; i < D; ++i)
            r[i] = a[i] - b[i];
        return r;
    }
    // utility.cuh:236-239 -- left-associated sums of products
    inline float dot(const Vec<2> &a, const Vec<2> &b) { return a[0] * b[0] + a[1] * b[1]; }
    inline float dot(const Vec<3> &a, const Vec<3> &b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
    template <int D>
    inline float sqlen(const Vec<D> &a) { return dot(a, a); }
    template <int D>
    inline float length(const Vec<D> &a) { return sqrtf(dot(a, a)); } // utility.cuh:382-398
    template <int D>
    inline Vec<D> normalize(const Vec<D> &a) // utility.cuh:427-447
    {
        const float n = length(a);
        Vec<D> r;
        for (int i = 0; i < D; ++i)
            r[i] = a[i] / n;
        return r;
    }
    inline Vec<3> cross(const Vec<3> &u, const Vec<3> &v) // utility.cuh:449-455
    {
        return {{u[1] * v[2] - u[2] * v[1], u[2] * v[0] - u[0] * v[2], u[0] * v[1] - u[1] * v[0]}};
    }
    template <int D>
    inline Vec<D> cwiseabs(const Vec<D> &a)
    {
        Vec<D> r;
        for (int i = 0; i < D; ++i)
            r[i] = fabsf(a[i]);
        return r;
    }

    struct Node // core/bvh.cuh:27-33
    {
        uint32_t parent, left, right, object;
    };
    template <int D>
    struct Box // core/aabb.cuh:11-16
    {
        Vec<D> up, lo;
    };
    template <int D>
    struct Cone // core/cone.cuh:10-16
    {
        Vec<D> axis;
        float half_angle, radius;
    };

    template <int D>
    inline Vec<D> centroid(const Box<D> &b) // aabb.cuh:262-279 ((u+l) * 0.5)
    {
        Vec<D> c;
        for (int i = 0; i < D; ++i)
            c[i] = (b.up[i] + b.lo[i]) * 0.5f;
        return c;
    }
    template <int D>
    inline Box<D> merge(const Box<D> &a, const Box<D> &b) // aabb.cuh:103-124 (fmax/fmin)
    {
        Box<D> m;
        for (int i = 0; i < D; ++i)
        {
            m.up[i] = dm_fmax(a.up[i], b.up[i]);
            m.lo[i] = dm_fmin(a.lo[i], b.lo[i]);
        }
        return m;
    }
    template <int D>
    inline float mindist(const Box<D> &b, const Vec<D> &p) // aabb.cuh:130-150
    {
        float s = 0.0f;
        for (int i = 0; i < D; ++i)
        {
            const float d = dm_fmin(b.up[i], dm_fmax(b.lo[i], p[i])) - p[i];
            s = (i == 0) ? d * d : s + d * d;
        }
        return s;
    }
    template <int D>
    inline float minmaxdist(const Box<D> &b, const Vec<D> &p) // aabb.cuh:160-209
    {
        float rm[D], rM[D];
        for (int i = 0; i < D; ++i)
        {
            rm[i] = (b.lo[i] - p[i]) * (b.lo[i] - p[i]);
            rM[i] = (b.up[i] - p[i]) * (b.up[i] - p[i]);
            if ((b.up[i] + b.lo[i]) * 0.5f < p[i])
                std::swap(rm[i], rM[i]);
        }
        if (D == 2)
        {
            const float dx = rm[0] + rM[1];
            const float dy = rM[0] + rm[1];
            return dm_fmin(dx, dy);
        }
        const float dx = rm[0] + rM[1] + rM[2];
        const float dy = rM[0] + rm[1] + rM[2];
        const float dz = rM[0] + rM[1] + rm[2];
        return dm_fmin(dx, dm_fmin(dy, dz));
    }

    // ------------------------------------------------------------------ scene
    template <int D>
    struct Scene
    {
        int nverts = 0, nprims = 0;
        std::vector<Vec<D>> verts;
        std::vector<int> prims;    // nprims * D
        std::vector<int> sil_rec;  // nsil * 4        (scene.cuh:291-294 / 709-712 "indices")
        std::vector<int> prim_sil; // nprims * D      (line_segment / triangle ::silhouette_indices)
        // bvh
        std::vector<Node> nodes;
        std::vector<Box<D>> boxes;
        std::vector<Cone<D>> cones;
        std::vector<uint32_t> morton_sorted, obj_sorted;
        bool used64 = false;

        const Vec<D> &V(int i) const { return verts[i]; }
        const int *P(int i) const { return &prims[(size_t)i * D]; }
        const int *SR(int i) const { return &sil_rec[(size_t)i * 4]; }
        int nsil() const { return (int)sil_rec.size() / 4; }
    };

    // scene.cuh:322-325 / 743-746
    template <int D>
    inline bool sil_has_face(const int *rec, int f)
    {
        if (D == 2)
            return f == 0 ? rec[2] != -1 : rec[0] != -1;
        return f == 0 ? rec[3] != -1 : rec[0] != -1;
    }
    // scene.cuh:326-336 (2D): n = (s.y, -s.x), s = pb - pa
    inline Vec<2> sil_face_normal(const Scene<2> &s, const int *rec, int f, bool do_norm)
    {
        const int i = f == 0 ? 1 : 0;
        const Vec<2> &pa = s.V(rec[i]), &pb = s.V(rec[i + 1]);
        const Vec<2> d = sub(pb, pa);
        const Vec<2> n = {{d[1], -d[0]}};
        return do_norm ? normalize(n) : n;
    }
    // scene.cuh:747-772 (3D): cross(pb - pa, pc - pa) with (i,j,k) = (3,1,2) or (0,2,1)
    inline Vec<3> sil_face_normal(const Scene<3> &s, const int *rec, int f, bool do_norm)
    {
        int i, j, k;
        if (f == 0)
            i = 3, j = 1, k = 2;
        else
            i = 0, j = 2, k = 1;
        const Vec<3> &pa = s.V(rec[j]), &pb = s.V(rec[k]), &pc = s.V(rec[i]);
        const Vec<3> n = cross(sub(pb, pa), sub(pc, pa));
        return do_norm ? normalize(n) : n;
    }
    // scene.cuh:337-353 / 773-787: normalised sum of the unnormalised face normals
    template <int D>
    inline Vec<D> sil_normal(const Scene<D> &s, const int *rec)
    {
        Vec<D> n;
        for (int i = 0; i < D; ++i)
            n[i] = 0.0f;
        for (int f = 0; f < 2; ++f)
            if (sil_has_face<D>(rec, f))
            {
                const Vec<D> nf = sil_face_normal(s, rec, f, false);
                for (int i = 0; i < D; ++i)
                    n[i] = n[i] + nf[i];
            }
        return normalize(n);
    }
    inline Vec<2> sil_centroid(const Scene<2> &s, const int *rec) { return s.V(rec[1]); } // scene.cuh:317-321
    inline Vec<3> sil_centroid(const Scene<3> &s, const int *rec)                          // scene.cuh:736-742
    {
        const Vec<3> &pa = s.V(rec[1]), &pb = s.V(rec[2]);
        return {{(pa[0] + pb[0]) / 2, (pa[1] + pb[1]) / 2, (pa[2] + pb[2]) / 2}};
    }

    // scene.cuh:631-652: one silhouette record per VERTEX: (prev, self, next, -)
    void compute_silhouettes(Scene<2> &s)
    {
        s.sil_rec.assign((size_t)s.nverts * 4, -1);
        for (int i = 0; i < s.nprims; ++i)
        {
            const int x = s.P(i)[0], y = s.P(i)[1];
            int *a = &s.sil_rec[(size_t)x * 4];
            a[1] = x;
            a[2] = y;
            int *b = &s.sil_rec[(size_t)y * 4];
            b[0] = x;
            b[1] = y;
        }
    }
    // scene.cuh:1135-1204: one record per undirected EDGE in first-seen order: (opp. vertex of face 1, a, b, opp. vertex of face 0)
    void compute_silhouettes(Scene<3> &s, std::vector<int> &edge_of_tri)
    {
        std::map<std::pair<int, int>, int> index_map;
        edge_of_tri.assign((size_t)s.nprims * 3, -1);
        int E = 0;
        for (int i = 0; i < s.nprims; ++i)
            for (int j = 0; j < 3; ++j)
            {
                const int k = (j + 1) % 3;
                int I = s.P(i)[j], J = s.P(i)[k];
                if (I > J)
                    std::swap(I, J);
                auto it = index_map.find({I, J});
                if (it == index_map.end())
                    it = index_map.emplace(std::make_pair(I, J), E++).first;
                edge_of_tri[(size_t)i * 3 + j] = it->second;
            }
        s.sil_rec.assign((size_t)E * 4, -1);
        for (int i = 0; i < s.nprims; ++i)
            for (int j = 0; j < 3; ++j)
            {
                const int I = j - 1 < 0 ? 2 : j - 1;
                int J = j, K = j + 1 > 2 ? 0 : j + 1;
                int orientation = 1;
                if (s.P(i)[J] > s.P(i)[K])
                {
                    std::swap(J, K);
                    orientation = -1;
                }
                int *rec = &s.sil_rec[(size_t)edge_of_tri[(size_t)i * 3 + j] * 4];
                rec[orientation == 1 ? 0 : 3] = s.P(i)[I];
                rec[1] = s.P(i)[J];
                rec[2] = s.P(i)[K];
            }
    }

    // scene.cuh:654-672 / 1205-1225: each silhouette element belongs to the FIRST primitive that references it
    void assign_silhouettes(Scene<2> &s)
    {
        std::vector<char> seen(s.nverts, 0);
        s.prim_sil.assign((size_t)s.nprims * 2, -1);
        for (int i = 0; i < s.nprims; ++i)
        {
            int ptr = 0;
            for (int j = 0; j < 2; ++j)
            {
                const int v = s.P(i)[j];
                if (!seen[v])
                {
                    seen[v] = 1;
                    s.prim_sil[(size_t)i * 2 + ptr++] = v;
                }
            }
        }
    }
    void assign_silhouettes(Scene<3> &s, const std::vector<int> &edge_of_tri)
    {
        std::vector<char> seen(s.nsil(), 0);
        s.prim_sil.assign((size_t)s.nprims * 3, -1);
        for (int i = 0; i < s.nprims; ++i)
        {
            int ptr = 0;
            for (int j = 0; j < 3; ++j)
            {
                const int e = edge_of_tri[(size_t)i * 3 + j];
                if (!seen[e])
                {
                    seen[e] = 1;
                    s.prim_sil[(size_t)i * 3 + ptr++] = e;
                }
            }
        }
    }

    // scene.cuh:420-433 (2D: pad by bvh_offset) / 870-885 + aabb.cuh:21-50,93-101 (3D: pad by float epsilon)
    inline Box<2> prim_box(const Scene<2> &s, int i)
    {
        const Vec<2> &p0 = s.V(s.P(i)[0]), &p1 = s.V(s.P(i)[1]);
        Box<2> b;
        for (int k = 0; k < 2; ++k)
        {
            b.up[k] = dm_max(p0[k], p1[k]) + BVH_OFFSET;
            b.lo[k] = dm_min(p0[k], p1[k]) - BVH_OFFSET;
        }
        return b;
    }
    inline Box<3> prim_box(const Scene<3> &s, int i)
    {
        const Vec<3> &pa = s.V(s.P(i)[0]);
        Box<3> b;
        for (int k = 0; k < 3; ++k)
        {
            b.up[k] = pa[k] + EPS_F;
            b.lo[k] = pa[k] - EPS_F;
        }
        for (int v = 1; v < 3; ++v)
        {
            const Vec<3> &p = s.V(s.P(i)[v]);
            for (int k = 0; k < 3; ++k)
            {
                b.lo[k] = dm_fmin(b.lo[k], p[k] - EPS_F);
                b.up[k] = dm_fmax(b.up[k], p[k] + EPS_F);
            }
        }
        return b;
    }

    // scene.cuh:435-502 / 887-961
    template <int D>
    Cone<D> prim_cone(const Scene<D> &s, int i)
    {
        const Box<D> box = prim_box(s, i);
        const Vec<D> bc = centroid(box);
        Cone<D> c;
        for (int k = 0; k < D; ++k)
            c.axis[k] = 0.0f;
        c.half_angle = PI_F;
        c.radius = 0.0f;
        bool any = false, all_two = true;
        for (int j = 0; j < D; ++j)
        {
            const int si = s.prim_sil[(size_t)i * D + j];
            if (si == -1)
                continue;
            const int *rec = s.SR(si);
            const Vec<D> n = sil_normal(s, rec);
            for (int k = 0; k < D; ++k)
                c.axis[k] += n[k];
            c.radius = dm_max(c.radius, length(sub(sil_centroid(s, rec), bc)));
            all_two = all_two && sil_has_face<D>(rec, 0) && sil_has_face<D>(rec, 1);
            any = true;
        }
        if (!any)
            c.half_angle = -PI_F;
        else if (!all_two)
            c.half_angle = PI_F;
        else
        {
            const float an = length(c.axis);
            if (an > EPS_F)
            {
                for (int k = 0; k < D; ++k)
                    c.axis[k] /= an;
                c.half_angle = 0.0f;
                for (int j = 0; j < D; ++j)
                {
                    const int si = s.prim_sil[(size_t)i * D + j];
                    if (si == -1)
                        continue;
                    const int *rec = s.SR(si);
                    for (int f = 0; f < 2; ++f)
                    {
                        Vec<D> n;
                        if (sil_has_face<D>(rec, f))
                            n = sil_face_normal(s, rec, f, true);
                        else
                            for (int k = 0; k < D; ++k)
                                n[k] = 0.0f;
                        const float ang = m_acos(dm_max(-1.0f, dm_min(1.0f, dot(c.axis, n))));
                        c.half_angle = dm_max(c.half_angle, ang);
                    }
                }
            }
        }
        return c;
    }

    // core/morton_code.cuh:19-38
    inline uint32_t expand_bits(uint32_t v)
    {
        v = (v | (v << 16)) & 0x070000FFu;
        v = (v | (v << 8)) & 0x0700F00Fu;
        v = (v | (v << 4)) & 0x430C30C3u;
        v = (v | (v << 2)) & 0x49249249u;
        return v;
    }
    // core/bvh.cuh:258-304 + core/morton_code.cuh:61-70,85-92
    template <int D>
    inline uint32_t morton_of(const Box<D> &box, const Box<D> &whole)
    {
        Vec<D> p = centroid(box);
        uint32_t code = 0;
        for (int k = 0; k < D; ++k)
        {
            float x = p[k] - whole.lo[k];
            x = x / (whole.up[k] - whole.lo[k]);
            x = dm_fmin(dm_fmax(x * 1024.0f, 0.0f), 1024.0f - 1.0f);
            const uint32_t e = expand_bits((uint32_t)x);
            code = (D == 2) ? (k == 0 ? e * 2 : code + e) : (k == 0 ? e * 4 : (k == 1 ? code + e * 2 : code + e));
        }
        return code;
    }

    inline int clz32(uint32_t x) { return x ? __builtin_clz(x) : 32; }
    inline int clz64(uint64_t x) { return x ? __builtin_clzll(x) : 64; }
    inline int common_upper_bits(uint32_t a, uint32_t b) { return clz32(a ^ b); } // morton_code.cuh:144-167
    inline int common_upper_bits(uint64_t a, uint64_t b) { return clz64(a ^ b); }

    // core/bvh.cuh:110-167
    template <typename U>
    inline void determine_range(const U *code, int n, int idx, int &first, int &last)
    {
        if (idx == 0)
        {
            first = 0;
            last = n - 1;
            return;
        }
        const U self = code[idx];
        const int Ld = common_upper_bits(self, code[idx - 1]);
        const int Rd = common_upper_bits(self, code[idx + 1]);
        const int d = (Rd > Ld) ? 1 : -1;
        const int dmin = std::min(Ld, Rd);
        int lmax = 2;
        int delta = -1;
        int it = idx + d * lmax;
        if (0 <= it && it < n)
            delta = common_upper_bits(self, code[it]);
        while (delta > dmin)
        {
            lmax <<= 1;
            it = idx + d * lmax;
            delta = -1;
            if (0 <= it && it < n)
                delta = common_upper_bits(self, code[it]);
        }
        int l = 0;
        int t = lmax >> 1;
        while (t > 0)
        {
            it = idx + (l + t) * d;
            delta = -1;
            if (0 <= it && it < n)
                delta = common_upper_bits(self, code[it]);
            if (delta > dmin)
                l += t;
            t >>= 1;
        }
        int j = idx + l * d;
        first = std::min(idx, j);
        last = std::max(idx, j);
    }
    // core/bvh.cuh:170-199
    template <typename U>
    inline int find_split(const U *code, int first, int last)
    {
        const U fc = code[first], lc = code[last];
        if (fc == lc)
            return (first + last) >> 1;
        const int dn = common_upper_bits(fc, lc);
        int split = first;
        int stride = last - first;
        do
        {
            stride = (stride + 1) >> 1;
            const int mid = split + stride;
            if (mid < last)
            {
                const int delta = common_upper_bits(fc, code[mid]);
                if (delta > dn)
                    split = mid;
            }
        } while (stride > 1);
        return split;
    }

    // core/cone.cuh:266-275 (2D) and :288-302 (3D)
    inline Vec<2> rotate(const Vec<2> &u, const Vec<2> &v, float theta)
    {
        const float det = u[0] * v[1] - u[1] * v[0];
        theta *= copysignf(1.0f, det);
        const float c = m_cos(theta), s = m_sin(theta);
        return {{c * u[0] - s * u[1], s * u[0] + c * u[1]}};
    }
    inline Vec<3> rotate(const Vec<3> &u, const Vec<3> &v, float theta)
    {
        const float c = m_cos(theta), s = m_sin(theta);
        const Vec<3> w = normalize(cross(u, v));
        const Vec<3> o = {{(1.0f - c) * w[0], (1.0f - c) * w[1], (1.0f - c) * w[2]}};
        const float R[3][3] = {
            {c + o[0] * w[0], o[1] * w[0] - s * w[2], o[2] * w[0] + s * w[1]},
            {o[0] * w[1] + s * w[2], c + o[1] * w[1], o[2] * w[1] - s * w[0]},
            {o[0] * w[2] - s * w[1], o[1] * w[2] + s * w[0], c + o[2] * w[2]}};
        return {{R[0][0] * u[0] + R[0][1] * u[1] + R[0][2] * u[2],
                 R[1][0] * u[0] + R[1][1] * u[1] + R[1][2] * u[2],
                 R[2][0] * u[0] + R[2][1] * u[1] + R[2][2] * u[2]}};
    }

    // core/cone.cuh:320-372 / 427-480. The reference leaves half_angle UNINITIALISED when the merged
    // cone would exceed pi (cone.cuh:347-352); the only meaningful value there is "everything" = pi,
    // which is what the oracle (and the product) store. Documented in DESIGN.md ("reference bugs").
    template <int D>
    Cone<D> merge_cones(const Cone<D> &A, const Cone<D> &B, const Vec<D> &oa, const Vec<D> &ob, const Vec<D> &on)
    {
        Cone<D> r;
        for (int k = 0; k < D; ++k)
            r.axis[k] = 0.0f;
        r.radius = 0.0f;
        r.half_angle = PI_F;
        const bool va = A.half_angle >= 0.0f, vb = B.half_angle >= 0.0f;
        if (va && vb)
        {
            Vec<D> axa = A.axis, axb = B.axis;
            float ha = A.half_angle, hb = B.half_angle;
            const Vec<D> da = sub(on, oa), db = sub(on, ob);
            r.radius = sqrtf(dm_max(A.radius * A.radius + sqlen(da), B.radius * B.radius + sqlen(db)));
            if (hb > ha)
            {
                std::swap(axa, axb);
                std::swap(ha, hb);
            }
            const float theta = m_acos(dm_max(-1.0f, dm_min(1.0f, dot(axa, axb))));
            if (dm_min(theta + hb, PI_F) <= ha)
            {
                r.axis = axa;
                r.half_angle = ha;
                return r;
            }
            const float ot = (ha + theta + hb) / 2.0f;
            if (ot >= PI_F)
            {
                r.axis = axa;
                r.half_angle = PI_F; // see note above
                return r;
            }
            const float rt = ot - ha;
            r.axis = rotate(axa, axb, rt);
            r.half_angle = ot;
        }
        else if (va)
            r = A;
        else if (vb)
            r = B;
        else
        {
            // reference leaves axis/radius uninitialised for invalid cones; only half_angle < 0 is ever read
            r.half_angle = -PI_F;
        }
        return r;
    }

    // core/bvh.cuh:380-613
    template <int D>
    void build_bvh(Scene<D> &s)
    {
        const int N = s.nprims;
        if (N == 0)
            return;
        const int NI = N - 1, NN = 2 * N - 1;
        std::vector<Box<D>> lb(N);
        std::vector<Cone<D>> lc(N);
        for (int i = 0; i < N; ++i)
        {
            lb[i] = prim_box(s, i);
            lc[i] = prim_cone(s, i);
        }
        Box<D> whole;
        for (int k = 0; k < D; ++k)
        {
            whole.up[k] = -INF_F;
            whole.lo[k] = INF_F;
        }
        for (int i = 0; i < N; ++i)
            whole = merge(whole, lb[i]);
        std::vector<uint32_t> code(N);
        for (int i = 0; i < N; ++i)
            code[i] = morton_of(lb[i], whole);
        std::vector<uint32_t> order(N);
        for (int i = 0; i < N; ++i)
            order[i] = i;
        std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b)
                         { return code[a] < code[b]; });
        s.morton_sorted.resize(N);
        s.obj_sorted = order;
        for (int i = 0; i < N; ++i)
            s.morton_sorted[i] = code[order[i]];
        bool unique = true;
        for (int i = 1; i < N; ++i)
            if (s.morton_sorted[i] == s.morton_sorted[i - 1])
            {
                unique = false;
                break;
            }
        s.used64 = !unique;

        Node dn = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu};
        s.nodes.assign(NN, dn);
        s.boxes.resize(NN);
        s.cones.resize(NN);
        for (int i = 0; i < NN; ++i)
        {
            for (int k = 0; k < D; ++k)
            {
                s.boxes[i].up[k] = -INF_F;
                s.boxes[i].lo[k] = INF_F;
                s.cones[i].axis[k] = 0.0f;
            }
            s.cones[i].half_angle = PI_F;
            s.cones[i].radius = 0.0f;
        }
        for (int i = 0; i < N; ++i)
        {
            s.nodes[NI + i].object = order[i];
            s.boxes[NI + i] = lb[order[i]];
            s.cones[NI + i] = lc[order[i]];
        }
        if (N == 1)
            return;
        std::vector<uint64_t> code64;
        if (!unique)
        {
            code64.resize(N);
            for (int i = 0; i < N; ++i)
                code64[i] = ((uint64_t)s.morton_sorted[i] << 32) | (uint64_t)order[i];
        }
        for (int idx = 0; idx < NI; ++idx) // core/bvh.cuh:201-229
        {
            int first, last, gamma;
            if (unique)
            {
                determine_range(s.morton_sorted.data(), N, idx, first, last);
                gamma = find_split(s.morton_sorted.data(), first, last);
            }
            else
            {
                determine_range(code64.data(), N, idx, first, last);
                gamma = find_split(code64.data(), first, last);
            }
            uint32_t L = gamma, R = gamma + 1;
            if (first == gamma)
                L += NI;
            if (last == gamma + 1)
                R += NI;
            s.nodes[idx].left = L;
            s.nodes[idx].right = R;
            s.nodes[L].parent = idx;
            s.nodes[R].parent = idx;
        }
        // bottom-up refit (bvh.cuh:518-604): children before parents via explicit post-order
        std::vector<std::pair<uint32_t, int>> st;
        st.push_back({0u, 0});
        while (!st.empty())
        {
            auto &top = st.back();
            const uint32_t n = top.first;
            if (s.nodes[n].object != 0xFFFFFFFFu)
            {
                st.pop_back();
                continue;
            }
            if (top.second == 0)
            {
                top.second = 1;
                st.push_back({s.nodes[n].left, 0});
            }
            else if (top.second == 1)
            {
                top.second = 2;
                st.push_back({s.nodes[n].right, 0});
            }
            else
            {
                const uint32_t l = s.nodes[n].left, r = s.nodes[n].right;
                s.boxes[n] = merge(s.boxes[l], s.boxes[r]);
                s.cones[n] = merge_cones(s.cones[l], s.cones[r], centroid(s.boxes[l]), centroid(s.boxes[r]), centroid(s.boxes[n]));
                st.pop_back();
            }
        }
    }

    // ------------------------------------------------------------------ primitive tests
    // scene.cuh:176-201 and :230-255 (same code for float2 / float3)
    template <int D>
    inline float closest_point_segment(const Vec<D> &pa, const Vec<D> &pb, const Vec<D> &x, Vec<D> *pt, float *t)
    {
        const Vec<D> u = sub(pb, pa), v = sub(x, pa);
        const float c1 = dot(u, v);
        if (c1 <= 0.0f)
        {
            *pt = pa;
            *t = 0.0f;
            return length(sub(x, *pt));
        }
        const float c2 = dot(u, u);
        if (c2 <= c1)
        {
            *pt = pb;
            *t = 1.0f;
            return length(sub(x, *pt));
        }
        *t = c1 / c2;
        for (int k = 0; k < D; ++k)
            (*pt)[k] = pa[k] + u[k] * (*t);
        return length(sub(x, *pt));
    }
    // scene.cuh:34-110
    inline float closest_point_triangle(const Vec<3> &pa, const Vec<3> &pb, const Vec<3> &pc, const Vec<3> &x, Vec<3> *pt)
    {
        const Vec<3> ab = sub(pb, pa), ac = sub(pc, pa), ax = sub(x, pa);
        const float d1 = dot(ab, ax), d2 = dot(ac, ax);
        if (d1 <= 0.0f && d2 <= 0.0f)
        {
            *pt = pa;
            return length(sub(x, *pt));
        }
        const Vec<3> bx = sub(x, pb);
        const float d3 = dot(ab, bx), d4 = dot(ac, bx);
        if (d3 >= 0.0f && d4 <= d3)
        {
            *pt = pb;
            return length(sub(x, *pt));
        }
        const Vec<3> cx = sub(x, pc);
        const float d5 = dot(ab, cx), d6 = dot(ac, cx);
        if (d6 >= 0.0f && d5 <= d6)
        {
            *pt = pc;
            return length(sub(x, *pt));
        }
        const float vc = d1 * d4 - d3 * d2;
        if (vc <= 0.0f && d1 >= 0.0f && d3 <= 0.0f)
        {
            const float v = d1 / (d1 - d3);
            *pt = {{pa[0] + ab[0] * v, pa[1] + ab[1] * v, pa[2] + ab[2] * v}};
            return length(sub(x, *pt));
        }
        const float vb = d5 * d2 - d1 * d6;
        if (vb <= 0.0f && d2 >= 0.0f && d6 <= 0.0f)
        {
            const float w = d2 / (d2 - d6);
            *pt = {{pa[0] + ac[0] * w, pa[1] + ac[1] * w, pa[2] + ac[2] * w}};
            return length(sub(x, *pt));
        }
        const float va = d3 * d6 - d5 * d4;
        if (va <= 0.0f && (d4 - d3) >= 0.0f && (d5 - d6) >= 0.0f)
        {
            const float w = (d4 - d3) / ((d4 - d3) + (d5 - d6));
            *pt = {{pb[0] + (pc[0] - pb[0]) * w, pb[1] + (pc[1] - pb[1]) * w, pb[2] + (pc[2] - pb[2]) * w}};
            return length(sub(x, *pt));
        }
        const float denom = 1.0f / (va + vb + vc);
        const float v = vb * denom, w = vc * denom;
        *pt = {{pa[0] + ab[0] * v + ac[0] * w, pa[1] + ab[1] * v + ac[1] * w, pa[2] + ab[2] * v + ac[2] * w}};
        return length(sub(x, *pt));
    }
    inline float prim_distance(const Scene<2> &s, int obj, const Vec<2> &x) // scene.cuh:504-515
    {
        Vec<2> pt;
        float t;
        return closest_point_segment(s.V(s.P(obj)[0]), s.V(s.P(obj)[1]), x, &pt, &t);
    }
    inline float prim_distance(const Scene<3> &s, int obj, const Vec<3> &x) // scene.cuh:963-976
    {
        Vec<3> pt;
        return closest_point_triangle(s.V(s.P(obj)[0]), s.V(s.P(obj)[1]), s.V(s.P(obj)[2]), x, &pt);
    }

    // scene.cuh:112-141
    inline bool is_silhouette_vertex(const Vec<2> &n0, const Vec<2> &n1, const Vec<2> &view, float d, bool flip)
    {
        const float prec = 1e-3f;
        const float sign = flip ? 1.0f : -1.0f;
        if (d <= prec)
        {
            const float det = n0[0] * n1[1] - n0[1] * n1[0];
            return sign * det > prec;
        }
        const Vec<2> vu = {{view[0] / d, view[1] / d}};
        const float dot0 = dot(vu, n0), dot1 = dot(vu, n1);
        if (fabsf(dot0) <= prec)
            return sign * dot1 > prec;
        if (fabsf(dot1) <= prec)
            return sign * dot0 > prec;
        return dot0 * dot1 < 0.0f;
    }
    // scene.cuh:143-174 (note: view_dir is NOT normalised in 3D -- restated as is)
    inline bool is_silhouette_edge(const Vec<3> &pa, const Vec<3> &pb, const Vec<3> &n0, const Vec<3> &n1,
                                   const Vec<3> &view, float d, bool flip)
    {
        const float prec = 1e-3f;
        const float sign = flip ? 1.0f : -1.0f;
        if (d <= prec)
        {
            const Vec<3> ed = normalize(sub(pb, pa));
            const float ang = m_atan2(dot(ed, cross(n0, n1)), dot(n0, n1));
            return sign * ang > prec;
        }
        const float dot0 = dot(view, n0), dot1 = dot(view, n1);
        if (fabsf(dot0) <= prec)
            return sign * dot1 > prec;
        if (fabsf(dot1) <= prec)
            return sign * dot0 > prec;
        return dot0 * dot1 < 0.0f;
    }
    // scene.cuh:354-386
    inline bool sil_closest(const Scene<2> &s, const int *rec, const Vec<2> &o, float max_r2, float &dist, bool flip, float min_r2)
    {
        if (min_r2 >= max_r2)
            return false;
        const Vec<2> &p = s.V(rec[1]);
        const Vec<2> view = sub(o, p);
        const float d = length(view);
        if (d * d > max_r2)
            return false;
        bool is_sil = !sil_has_face<2>(rec, 0) || !sil_has_face<2>(rec, 1);
        if (!is_sil)
            is_sil = is_silhouette_vertex(sil_face_normal(s, rec, 0, true), sil_face_normal(s, rec, 1, true), view, d, flip);
        if (is_sil && d * d <= max_r2)
        {
            dist = d;
            return true;
        }
        return false;
    }
    // scene.cuh:788-824
    inline bool sil_closest(const Scene<3> &s, const int *rec, const Vec<3> &o, float max_r2, float &dist, bool flip, float min_r2)
    {
        if (min_r2 >= max_r2)
            return false;
        const Vec<3> &pa = s.V(rec[1]), &pb = s.V(rec[2]);
        Vec<3> cp;
        float t;
        const float d = closest_point_segment(pa, pb, o, &cp, &t);
        if (d * d > max_r2)
            return false;
        bool is_sil = !sil_has_face<3>(rec, 0) || !sil_has_face<3>(rec, 1);
        if (!is_sil)
            is_sil = is_silhouette_edge(pa, pb, sil_face_normal(s, rec, 0, true), sil_face_normal(s, rec, 1, true), sub(o, cp), d, flip);
        if (is_sil && d * d <= max_r2)
        {
            dist = d;
            return true;
        }
        return false;
    }
    // scene.cuh:517-541 / 978-1003
    template <int D>
    inline bool prim_sil_distance(const Scene<D> &s, int obj, const Vec<D> &o, float max_r2, float &dist, bool flip, float min_r2)
    {
        float maxd = max_r2;
        bool ret = false;
        for (int j = 0; j < D; ++j)
        {
            const int si = s.prim_sil[(size_t)obj * D + j];
            if (si == -1)
                continue;
            if (sil_closest(s, s.SR(si), o, maxd, dist, flip, min_r2))
            {
                ret = true;
                maxd = dist * dist;
            }
        }
        return ret;
    }

    // core/cone.cuh:24-66
    inline float project_to_plane(const Vec<2> &n, const Vec<2> &e)
    {
        const Vec<2> b = {{-n[1], n[0]}};
        return fabsf(dot(e, cwiseabs(b)));
    }
    inline float project_to_plane(const Vec<3> &n, const Vec<3> &e)
    {
        const float sign = copysignf(1.0f, n[2]);
        const float a = -1.0f / (sign + n[2]);
        const float b = n[0] * n[1] * a;
        const Vec<3> b1 = {{1.0f + sign * n[0] * n[0] * a, sign * b, -sign * n[0]}};
        const Vec<3> b2 = {{b, sign + n[1] * n[1] * a, -n[1]}};
        const float r1 = dot(e, cwiseabs(b1)), r2 = dot(e, cwiseabs(b2));
        return sqrtf(r1 * r1 + r2 * r2);
    }
    inline bool inrange(float v, float lo, float hi) { return v >= lo && v <= hi; }
    // core/cone.cuh:78-121 / 168-212
    template <int D>
    inline bool cone_overlap(const Cone<D> &bc, const Vec<D> &o, const Box<D> &b, float dist_to_box)
    {
        if (bc.half_angle >= PI_2_F || dist_to_box < EPS_F)
            return true;
        const Vec<D> c = centroid(b);
        Vec<D> va = sub(c, o);
        const float l = length(va);
        for (int k = 0; k < D; ++k)
            va[k] /= l;
        const float daa = m_acos(dm_max(-1.0f, dm_min(1.0f, dot(bc.axis, va))));
        if (inrange(PI_2_F, daa - bc.half_angle, daa + bc.half_angle))
            return true;
        if (l > bc.radius)
        {
            const float vh = m_asin(bc.radius / l);
            const float hs = bc.half_angle + vh;
            return hs >= PI_2_F ? true : inrange(PI_2_F, daa - hs, daa + hs);
        }
        const Vec<D> e = sub(b.up, c);
        float d = dot(e, cwiseabs(va));
        const float sdist = l - d;
        if (sdist <= 0.0f)
            return true;
        d = project_to_plane(va, e);
        const float vh = m_atan2(d, sdist);
        const float hs = bc.half_angle + vh;
        return hs >= PI_2_F ? true : inrange(PI_2_F, daa - hs, daa + hs);
    }

    // ------------------------------------------------------------------ queries
    struct StackEntry
    {
        uint32_t node;
        float d;
    };

    // core/query.cuh:238-318. N == 1 (root is the leaf) is handled explicitly; the reference would read
    // nodes[0].left_idx = 0xFFFFFFFF there (SURVEY "reference bugs").
    template <int D>
    inline void query_nearest(const Scene<D> &s, const Vec<D> &q, uint32_t &nearest, float &dist)
    {
        nearest = 0xFFFFFFFFu;
        float best = INF_F;
        if (s.nprims == 0)
        {
            dist = INF_F;
            return;
        }
        if (s.nprims == 1)
        {
            nearest = s.nodes[0].object;
            dist = prim_distance(s, nearest, q);
            dist = sqrtf(dist * dist);
            return;
        }
        StackEntry stack[128];
        int sp = 0;
        stack[sp++] = {0u, mindist(s.boxes[0], q)};
        do
        {
            const StackEntry e = stack[--sp];
            if (e.d > best)
                continue;
            const uint32_t L = s.nodes[e.node].left, R = s.nodes[e.node].right;
            const float Lmin = mindist(s.boxes[L], q), Rmin = mindist(s.boxes[R], q);
            const float Lmm = minmaxdist(s.boxes[L], q), Rmm = minmaxdist(s.boxes[R], q);
            if (Lmin <= Rmm)
            {
                const uint32_t obj = s.nodes[L].object;
                if (obj != 0xFFFFFFFFu)
                {
                    float d = prim_distance(s, obj, q);
                    d *= d;
                    if (d <= best)
                    {
                        best = d;
                        nearest = obj;
                    }
                }
                else
                    stack[sp++] = {L, Lmin};
            }
            if (Rmin <= Lmm)
            {
                const uint32_t obj = s.nodes[R].object;
                if (obj != 0xFFFFFFFFu)
                {
                    float d = prim_distance(s, obj, q);
                    d *= d;
                    if (d <= best)
                    {
                        best = d;
                        nearest = obj;
                    }
                }
                else
                    stack[sp++] = {R, Rmin};
            }
        } while (sp > 0);
        dist = sqrtf(best);
    }

    // core/query.cuh:325-423
    template <int D>
    inline float query_silhouette(const Scene<D> &s, const Vec<D> &q, bool flip)
    {
        float best = INF_F;
        if (s.nprims == 0)
            return best;
        if (s.nprims == 1)
        {
            float d = INF_F;
            if (s.cones[0].half_angle >= 0.0f && prim_sil_distance(s, (int)s.nodes[0].object, q, INF_F, d, flip, 0.0f))
                best = d;
            return best;
        }
        StackEntry stack[128];
        int sp = 0;
        stack[sp++] = {0u, mindist(s.boxes[0], q)};
        do
        {
            const StackEntry e = stack[--sp];
            if (e.d > best * best)
                continue;
            const uint32_t L = s.nodes[e.node].left, R = s.nodes[e.node].right;
            const float Lmin = mindist(s.boxes[L], q), Rmin = mindist(s.boxes[R], q);
            const bool hitL = s.cones[L].half_angle >= 0.0f && cone_overlap(s.cones[L], q, s.boxes[L], Lmin);
            const bool hitR = s.cones[R].half_angle >= 0.0f && cone_overlap(s.cones[R], q, s.boxes[R], Rmin);
            if (hitL)
            {
                const uint32_t obj = s.nodes[L].object;
                if (obj != 0xFFFFFFFFu)
                {
                    float d = INF_F;
                    const bool found = prim_sil_distance(s, (int)obj, q, best * best, d, flip, 0.0f);
                    if (found && d <= best)
                        best = d;
                }
                else
                    stack[sp++] = {L, Lmin};
            }
            if (hitR)
            {
                const uint32_t obj = s.nodes[R].object;
                if (obj != 0xFFFFFFFFu)
                {
                    float d = INF_F;
                    const bool found = prim_sil_distance(s, (int)obj, q, best * best, d, flip, 0.0f);
                    if (found && d <= best)
                        best = d;
                }
                else
                    stack[sp++] = {R, Rmin};
            }
        } while (sp > 0);
        return best;
    }

    template <int D>
    struct Ray // aabb.cuh:298-313
    {
        Vec<D> o, d, inv;
        Ray(const Vec<D> &o_, const Vec<D> &d_) : o(o_), d(d_)
        {
            for (int k = 0; k < D; ++k)
                inv[k] = 1 / d[k];
        }
    };
    // aabb.cuh:366-431 (std::fmin/std::fmax: NaN-suppressing)
    template <int D>
    inline bool ray_box(const Ray<D> &r, const Box<D> &b, float max_dist, float *dist)
    {
        float tmin = 0, tmax = 0;
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
        float t;
        float u, v;
    };
    // scene.cuh:543-577 (host branch: invD = 1/D; the product uses the same IEEE reciprocal)
    inline Hit ray_prim(const Scene<2> &s, int obj, const Ray<2> &r)
    {
        const Vec<2> &p0 = s.V(s.P(obj)[0]), &p1 = s.V(s.P(obj)[1]);
        const Vec<2> sd = sub(p1, p0);
        const float Dn = r.d[0] * (-sd[1]) + r.d[1] * sd[0];
        if (fabsf(Dn) < EPS_F)
            return {false, 0, 0, 0};
        const float inv = 1.0f / Dn;
        const float t = ((p0[0] - r.o[0]) * (-sd[1]) - (p0[1] - r.o[1]) * (-sd[0])) * inv;
        const float sp = (r.d[0] * (p0[1] - r.o[1]) - r.d[1] * (p0[0] - r.o[0])) * inv;
        if (sp >= -1e-3f && sp <= 1.0f + 1e-3f && t >= 0.0f)
            return {true, t, sp, 0};
        return {false, 0, 0, 0};
    }
    // scene.cuh:1005-1052 (Moller-Trumbore)
    inline Hit ray_prim(const Scene<3> &s, int obj, const Ray<3> &r)
    {
        const Vec<3> &v0 = s.V(s.P(obj)[0]), &v1 = s.V(s.P(obj)[1]), &v2 = s.V(s.P(obj)[2]);
        const Vec<3> e1 = sub(v1, v0), e2 = sub(v2, v0);
        const Vec<3> h = {{r.d[1] * e2[2] - r.d[2] * e2[1], r.d[2] * e2[0] - r.d[0] * e2[2], r.d[0] * e2[1] - r.d[1] * e2[0]}};
        const float det = e1[0] * h[0] + e1[1] * h[1] + e1[2] * h[2];
        if (fabsf(det) < EPS_F)
            return {false, 0, 0, 0};
        const float inv = 1.0f / det;
        const Vec<3> sv = sub(r.o, v0);
        const float u = (sv[0] * h[0] + sv[1] * h[1] + sv[2] * h[2]) * inv;
        if (u < 0.0f || u > 1.0f)
            return {false, 0, 0, 0};
        const Vec<3> q = {{sv[1] * e1[2] - sv[2] * e1[1], sv[2] * e1[0] - sv[0] * e1[2], sv[0] * e1[1] - sv[1] * e1[0]}};
        const float v = (r.d[0] * q[0] + r.d[1] * q[1] + r.d[2] * q[2]) * inv;
        if (v < 0.0f || u + v > 1.0f)
            return {false, 0, 0, 0};
        const float t = (e2[0] * q[0] + e2[1] * q[1] + e2[2] * q[2]) * inv;
        if (t >= 0.0f)
            return {true, t, u, v};
        return {false, 0, 0, 0};
    }

    struct RayResult
    {
        bool hit;
        float t;
        float u, v;
        uint32_t obj;
    };
    // core/query.cuh:79-169
    template <int D>
    inline RayResult query_ray(const Scene<D> &s, const Ray<D> &r, float max_dist, bool test_only)
    {
        RayResult res = {false, INF_F, 0, 0, 0xFFFFFFFFu};
        if (s.nprims == 0)
            return res;
        StackEntry stack[128];
        int sp = 0;
        stack[sp++] = {0u, INF_F};
        float best = INF_F;
        do
        {
            const StackEntry e = stack[--sp];
            if (e.d > best)
                continue;
            const uint32_t obj = s.nodes[e.node].object;
            if (obj != 0xFFFFFFFFu)
            {
                const Hit h = ray_prim(s, (int)obj, r);
                if (h.hit && h.t < max_dist && h.t < best)
                {
                    best = h.t;
                    res.hit = true;
                    if (test_only)
                        return res;
                    res.u = h.u;
                    res.v = h.v;
                    res.obj = obj;
                }
            }
            else
            {
                const uint32_t L = s.nodes[e.node].left, R = s.nodes[e.node].right;
                float Ld = 0, Rd = 0;
                const bool Lh = ray_box(r, s.boxes[L], max_dist, &Ld);
                const bool Rh = ray_box(r, s.boxes[R], max_dist, &Rd);
                if (Lh && Rh)
                {
                    uint32_t closer = L, other = R;
                    if (Rd < Ld)
                    {
                        std::swap(Ld, Rd);
                        std::swap(closer, other);
                    }
                    stack[sp++] = {other, Rd};
                    stack[sp++] = {closer, Ld};
                }
                else if (Lh)
                    stack[sp++] = {L, Ld};
                else if (Rh)
                    stack[sp++] = {R, Rd};
            }
        } while (sp > 0);
        res.t = best;
        return res;
    }

    // aabb.cuh:433-449 (std::max/std::min)
    template <int D>
    inline bool sphere_box(const Vec<D> &c, float radius, const Box<D> &b)
    {
        float d2 = 0.0f;
        for (int k = 0; k < D; ++k)
        {
            const float cl = dm_max(b.lo[k], dm_min(c[k], b.up[k]));
            const float d = cl - c[k];
            d2 = (k == 0) ? d * d : d2 + d * d;
        }
        return d2 <= radius * radius;
    }
    // scene.cuh:579-603
    inline bool sphere_prim(const Scene<2> &s, int obj, const Vec<2> &c, float radius)
    {
        const Vec<2> &p1 = s.V(s.P(obj)[0]), &p2 = s.V(s.P(obj)[1]);
        const Vec<2> d = sub(p2, p1);
        const float len2 = d[0] * d[0] + d[1] * d[1];
        float t = ((c[0] - p1[0]) * d[0] + (c[1] - p1[1]) * d[1]) / len2;
        t = dm_max(0.0f, dm_min(1.0f, t));
        const float cx = p1[0] + t * d[0], cy = p1[1] + t * d[1];
        const float dx = cx - c[0], dy = cy - c[1];
        return dx * dx + dy * dy <= radius * radius;
    }
    // scene.cuh:1054-1117
    inline bool sphere_prim(const Scene<3> &s, int obj, const Vec<3> &c, float radius)
    {
        const Vec<3> &p1 = s.V(s.P(obj)[0]), &p2 = s.V(s.P(obj)[1]), &p3 = s.V(s.P(obj)[2]);
        const Vec<3> e1 = sub(p2, p1), e2 = sub(p3, p1);
        Vec<3> n = cross(e1, e2);
        const float nl = sqrtf(n[0] * n[0] + n[1] * n[1] + n[2] * n[2]);
        n = {{n[0] / nl, n[1] / nl, n[2] / nl}};
        const float dd = n[0] * p1[0] + n[1] * p1[1] + n[2] * p1[2];
        const float dp = n[0] * c[0] + n[1] * c[1] + n[2] * c[2] - dd;
        const Vec<3> proj = {{c[0] - dp * n[0], c[1] - dp * n[1], c[2] - dp * n[2]}};
        const Vec<3> v0 = e2, v1 = e1, v2 = sub(proj, p1);
        const float d00 = dot(v0, v0), d01 = dot(v0, v1), d02 = dot(v0, v2), d11 = dot(v1, v1), d12 = dot(v1, v2);
        const float invd = 1.0f / (d00 * d11 - d01 * d01);
        const float u = (d11 * d02 - d01 * d12) * invd;
        const float v = (d00 * d12 - d01 * d02) * invd;
        if (u >= 0 && v >= 0 && u + v <= 1)
            return fabsf(dp) <= radius;
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
    // scene.cuh:409-418 / 855-868
    inline float prim_measure(const Scene<2> &s, int obj) { return length(sub(s.V(s.P(obj)[0]), s.V(s.P(obj)[1]))); }
    inline float prim_measure(const Scene<3> &s, int obj)
    {
        const Vec<3> &pa = s.V(s.P(obj)[0]), &pb = s.V(s.P(obj)[1]), &pc = s.V(s.P(obj)[2]);
        return length(cross(sub(pc, pa), sub(pb, pa))) / 2;
    }
    // scene.cuh:605-612 / 1119-1126 (device overload: float log)
    inline float green_weight(const Vec<2> &x, const Vec<2> &y)
    {
        const float r = dm_max(length(sub(x, y)), 1e-2f);
        return fabsf(m_log(r) / (PI_F * 2.0f));
    }
    inline float green_weight(const Vec<3> &x, const Vec<3> &y)
    {
        const float r = dm_max(length(sub(x, y)), 1e-4f);
        return 1.0f / (PI_F * 4.0f * r);
    }
    // core/sample.cuh:23-92
    template <int D>
    inline void sample_in_sphere(const Scene<D> &s, const Vec<D> &c, float radius, float u, int &obj_out, float &pdf_out)
    {
        obj_out = -1;
        pdf_out = 0.0f;
        if (s.nprims == 0)
            return;
        uint32_t node = 0;
        float pdf = 1.0f;
        for (;;)
        {
            const uint32_t obj = s.nodes[node].object;
            if (obj != 0xFFFFFFFFu)
            {
                if (sphere_prim(s, (int)obj, c, radius))
                {
                    obj_out = (int)obj;
                    pdf_out = pdf / prim_measure(s, (int)obj);
                }
                return;
            }
            const uint32_t L = s.nodes[node].left, R = s.nodes[node].right;
            const float Lw = sphere_box(c, radius, s.boxes[L]) ? green_weight(c, centroid(s.boxes[L])) : 0.0f;
            const float Rw = sphere_box(c, radius, s.boxes[R]) ? green_weight(c, centroid(s.boxes[R])) : 0.0f;
            const float tw = Lw + Rw;
            if (!(tw > 0))
                return;
            const float Lp = Lw / tw;
            if (u < Lp)
            {
                u /= Lp;
                node = L;
                pdf = Lp * pdf;
            }
            else
            {
                const float Rp = 1.0f - Lp;
                u = (u - Lp) / Rp;
                node = R;
                pdf = Rp * pdf;
            }
        }
    }

    struct AnyScene
    {
        int dim;
        Scene<2> *s2;
        Scene<3> *s3;
    };

    template <typename F>
    void parallel_for(int n, int nthreads, F &&f)
    {
        unsigned nt = nthreads > 0 ? (unsigned)nthreads : std::max(1u, std::thread::hardware_concurrency());
        if (n < 512 || nt == 1)
        {
            for (int i = 0; i < n; ++i)
                f(i);
            return;
        }
        std::atomic<int> next{0};
        std::vector<std::thread> pool;
        for (unsigned t = 0; t < nt; ++t)
            pool.emplace_back([&]()
                              {
                for (;;)
                {
                    const int b = next.fetch_add(128);
                    if (b >= n)
                        break;
                    const int e = std::min(n, b + 128);
                    for (int i = b; i < e; ++i)
                        f(i);
                } });
        for (auto &th : pool)
            th.join();
    }

    template <int D>
    inline Vec<D> ldv(const float *p, int i)
    {
        Vec<D> v;
        for (int k = 0; k < D; ++k)
            v[k] = p[(size_t)i * D + k];
        return v;
    }
}

#include "wost_oracle_walk.inl"
#include "wost_oracle_guided.inl"

extern "C"
{
    void orc_set_math_mode(int mode) { g_math_mode = mode; }
    void orc_set_rng_mode(int mode) { g_rng_mode = mode; }

    void *orc_scene_create(int dim, const float *verts, int nverts, const int *idx, int nprims)
    {
        AnyScene *a = new AnyScene{dim, nullptr, nullptr};
        if (dim == 2)
        {
            Scene<2> *s = new Scene<2>();
            s->nverts = nverts;
            s->nprims = nprims;
            s->verts.resize(nverts);
            for (int i = 0; i < nverts; ++i)
                s->verts[i] = ldv<2>(verts, i);
            s->prims.assign(idx, idx + (size_t)nprims * 2);
            compute_silhouettes(*s);
            assign_silhouettes(*s);
            build_bvh(*s);
            a->s2 = s;
        }
        else
        {
            Scene<3> *s = new Scene<3>();
            s->nverts = nverts;
            s->nprims = nprims;
            s->verts.resize(nverts);
            for (int i = 0; i < nverts; ++i)
                s->verts[i] = ldv<3>(verts, i);
            s->prims.assign(idx, idx + (size_t)nprims * 3);
            std::vector<int> eot;
            compute_silhouettes(*s, eot);
            assign_silhouettes(*s, eot);
            build_bvh(*s);
            a->s3 = s;
        }
        return a;
    }
    void orc_scene_destroy(void *h)
    {
        AnyScene *a = (AnyScene *)h;
        delete a->s2;
        delete a->s3;
        delete a;
    }
    int orc_num_nodes(void *h)
    {
        AnyScene *a = (AnyScene *)h;
        return a->dim == 2 ? (int)a->s2->nodes.size() : (int)a->s3->nodes.size();
    }
    int orc_used_morton64(void *h)
    {
        AnyScene *a = (AnyScene *)h;
        return a->dim == 2 ? a->s2->used64 : a->s3->used64;
    }
} // extern "C"
    template <int D>
    static void dump_bvh(const Scene<D> &s, uint32_t *nodes, float *aabbs, float *cones)
    {
        for (size_t i = 0; i < s.nodes.size(); ++i)
        {
            nodes[4 * i + 0] = s.nodes[i].parent;
            nodes[4 * i + 1] = s.nodes[i].left;
            nodes[4 * i + 2] = s.nodes[i].right;
            nodes[4 * i + 3] = s.nodes[i].object;
            for (int k = 0; k < D; ++k)
            {
                aabbs[2 * D * i + k] = s.boxes[i].up[k];
                aabbs[2 * D * i + D + k] = s.boxes[i].lo[k];
                cones[(D + 2) * i + k] = s.cones[i].axis[k];
            }
            cones[(D + 2) * i + D] = s.cones[i].half_angle;
            cones[(D + 2) * i + D + 1] = s.cones[i].radius;
        }
    }
extern "C"
{
    void orc_dump_bvh(void *h, uint32_t *nodes, float *aabbs, float *cones)
    {
        AnyScene *a = (AnyScene *)h;
        if (a->dim == 2)
            dump_bvh(*a->s2, nodes, aabbs, cones);
        else
            dump_bvh(*a->s3, nodes, aabbs, cones);
    }
    void orc_dump_sorted(void *h, uint32_t *morton, uint32_t *object_idx)
    {
        AnyScene *a = (AnyScene *)h;
        const auto &m = a->dim == 2 ? a->s2->morton_sorted : a->s3->morton_sorted;
        const auto &o = a->dim == 2 ? a->s2->obj_sorted : a->s3->obj_sorted;
        std::copy(m.begin(), m.end(), morton);
        std::copy(o.begin(), o.end(), object_idx);
    }
    void orc_dump_silhouette_assignment(void *h, int *sil)
    {
        AnyScene *a = (AnyScene *)h;
        const auto &v = a->dim == 2 ? a->s2->prim_sil : a->s3->prim_sil;
        std::copy(v.begin(), v.end(), sil);
    }
    int orc_num_silhouettes(void *h)
    {
        AnyScene *a = (AnyScene *)h;
        return a->dim == 2 ? a->s2->nsil() : a->s3->nsil();
    }
    void orc_dump_silhouettes(void *h, int *rec)
    {
        AnyScene *a = (AnyScene *)h;
        const auto &v = a->dim == 2 ? a->s2->sil_rec : a->s3->sil_rec;
        std::copy(v.begin(), v.end(), rec);
    }

    void orc_query_nearest(void *h, const float *pts, int n, uint32_t *idx, float *dist)
    {
        AnyScene *a = (AnyScene *)h;
        parallel_for(n, 0, [&](int i)
                     {
            if (a->dim == 2)
                query_nearest(*a->s2, ldv<2>(pts, i), idx[i], dist[i]);
            else
                query_nearest(*a->s3, ldv<3>(pts, i), idx[i], dist[i]); });
    }
    void orc_query_silhouette(void *h, const float *pts, int n, int flip, float *dist)
    {
        AnyScene *a = (AnyScene *)h;
        parallel_for(n, 0, [&](int i)
                     {
            if (a->dim == 2)
                dist[i] = query_silhouette(*a->s2, ldv<2>(pts, i), flip != 0);
            else
                dist[i] = query_silhouette(*a->s3, ldv<3>(pts, i), flip != 0); });
    }
    void orc_query_ray(void *h, const float *org, const float *dir, const float *max_dist, int n,
                       int *hit, float *t, float *uv, uint32_t *obj)
    {
        AnyScene *a = (AnyScene *)h;
        parallel_for(n, 0, [&](int i)
                     {
            RayResult r;
            if (a->dim == 2)
                r = query_ray(*a->s2, Ray<2>(ldv<2>(org, i), ldv<2>(dir, i)), max_dist[i], false);
            else
                r = query_ray(*a->s3, Ray<3>(ldv<3>(org, i), ldv<3>(dir, i)), max_dist[i], false);
            hit[i] = r.hit ? 1 : 0;
            t[i] = r.t;
            uv[2 * i] = r.hit ? r.u : 0.0f;
            uv[2 * i + 1] = r.hit ? r.v : 0.0f;
            obj[i] = r.obj; });
    }
    void orc_query_ray_any(void *h, const float *org, const float *dir, const float *max_dist, int n, int *hit)
    {
        AnyScene *a = (AnyScene *)h;
        parallel_for(n, 0, [&](int i)
                     {
            if (a->dim == 2)
                hit[i] = query_ray(*a->s2, Ray<2>(ldv<2>(org, i), ldv<2>(dir, i)), max_dist[i], true).hit ? 1 : 0;
            else
                hit[i] = query_ray(*a->s3, Ray<3>(ldv<3>(org, i), ldv<3>(dir, i)), max_dist[i], true).hit ? 1 : 0; });
    }
    void orc_sample_in_sphere(void *h, const float *ctr, const float *radius, const float *u, int n, int *obj, float *pdf)
    {
        AnyScene *a = (AnyScene *)h;
        for (int i = 0; i < n; ++i)
        {
            if (a->dim == 2)
                sample_in_sphere(*a->s2, ldv<2>(ctr, i), radius[i], u[i], obj[i], pdf[i]);
            else
                sample_in_sphere(*a->s3, ldv<3>(ctr, i), radius[i], u[i], obj[i], pdf[i]);
        }
    }

    void orc_detmath_eval(int fn, const float *x, const float *y, int n, float *out)
    {
        for (int i = 0; i < n; ++i)
        {
            switch (fn)
            {
            case 0: out[i] = dm_sin(x[i]); break;
            case 1: out[i] = dm_cos(x[i]); break;
            case 2: out[i] = dm_log(x[i]); break;
            case 3: out[i] = dm_exp(x[i]); break;
            case 4: out[i] = dm_atan(x[i]); break;
            case 5: out[i] = dm_atan2(y[i], x[i]); break;
            case 6: out[i] = dm_asin(x[i]); break;
            case 7: out[i] = dm_acos(x[i]); break;
            case 8: out[i] = dm_cbrt(x[i]); break;
            case 9: out[i] = dm_logistic(x[i]); break;
            default: out[i] = 0.0f;
            }
        }
    }
    void orc_philox(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1, uint32_t *out4)
    {
        const Philox4 r = philox4x32_10(c0, c1, c2, c3, k0, k1);
        out4[0] = r.x;
        out4[1] = r.y;
        out4[2] = r.z;
        out4[3] = r.w;
    }
}
