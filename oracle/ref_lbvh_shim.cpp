// TEST INFRASTRUCTURE ONLY -- never linked into, imported by, or executed from the product.
//
// Thin C wrapper that compiles the reference's OWN snch-lbvh headers, unmodified and from where
// they lie (/root/reference/ext/lbvh/include), with plain g++ and Thrust's CPP (host) backend.
// The result (oracle/_ref/libref_lbvh.so) is the pinned ground truth the restated oracle
// (oracle/wost_oracle.cpp) and the CUDA product are checked against. Built by oracle/Makefile;
// outputs only under oracle/_ref/ (git-ignored, travels to the GPU box with the snapshot).
//
// No reference source is copied here: this file only #includes the headers and forwards calls.
// Entry points used: lbvh::scene<2|3> (scene.cuh:288-703, 706-1268), lbvh::bvh::construct
// (core/bvh.cuh:380-613), lbvh::query_device nearest / nearest_silhouette / ray_intersect
// (core/query.cuh:79-169, 238-423), lbvh::sample_object_in_sphere (core/sample.cuh:23-92).
#include <algorithm>
#include <cassert>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <iostream>
#include <unordered_map>
#include <vector>

// ---- shim for the few device intrinsics the headers reference when not compiled by nvcc ----
using std::max;
using std::min;
#ifndef M_PIf
#define M_PIf 3.14159265358979323846f
#endif
static inline int atomicCAS(int *addr, int compare, int val)
{
    int old = *addr;
    if (old == compare)
        *addr = val;
    return old;
}
static inline void __threadfence() {}

#include <vector_types.h>
#include <vector_functions.h>
#include <thrust/pair.h>
#include <thrust/tuple.h>
#include "snch_lbvh/lbvh.cuh"
#include "snch_lbvh/scene.cuh"

#include <thread>
#include <atomic>
// simple chunked thread pool loop (OpenMP is not reliably linkable in this image)
template <typename F>
static void parallel_for(int n, F &&f)
{
    unsigned nt = std::max(1u, std::thread::hardware_concurrency());
    if (n < 1024 || nt == 1)
    {
        for (int i = 0; i < n; ++i)
            f(i);
        return;
    }
    std::atomic<int> next{0};
    std::vector<std::thread> pool;
    for (unsigned t = 0; t < nt; ++t)
        pool.emplace_back([&]() {
            for (;;)
            {
                int b = next.fetch_add(256);
                if (b >= n)
                    break;
                int e = std::min(n, b + 256);
                for (int i = b; i < e; ++i)
                    f(i);
            } });
    for (auto &th : pool)
        th.join();
}

namespace
{
    template <unsigned DIM>
    struct Handle
    {
        lbvh::scene<DIM> *scene;
    };
    struct AnyHandle
    {
        int dim;
        void *scene;
    };
    inline lbvh::scene<2> &S2(void *h) { return *static_cast<lbvh::scene<2> *>(static_cast<AnyHandle *>(h)->scene); }
    inline lbvh::scene<3> &S3(void *h) { return *static_cast<lbvh::scene<3> *>(static_cast<AnyHandle *>(h)->scene); }
}

extern "C"
{
    // verts: nverts*dim floats; idx: nprims*dim ints (0-based)
    void *ref_scene_create(int dim, const float *verts, int nverts, const int *idx, int nprims)
    {
        std::streambuf *old = std::cout.rdbuf(nullptr); // silence "Morton code collision detected."
        AnyHandle *h = new AnyHandle{dim, nullptr};
        if (dim == 2)
        {
            std::vector<float2> v(nverts);
            std::vector<int2> I(nprims);
            for (int i = 0; i < nverts; ++i)
                v[i] = make_float2(verts[2 * i], verts[2 * i + 1]);
            for (int i = 0; i < nprims; ++i)
                I[i] = make_int2(idx[2 * i], idx[2 * i + 1]);
            auto *s = new lbvh::scene<2>(v.begin(), v.end(), I.begin(), I.end());
            s->compute_silhouettes();
            s->build_bvh();
            h->scene = s;
        }
        else
        {
            std::vector<float3> v(nverts);
            std::vector<int3> I(nprims);
            for (int i = 0; i < nverts; ++i)
                v[i] = make_float3(verts[3 * i], verts[3 * i + 1], verts[3 * i + 2]);
            for (int i = 0; i < nprims; ++i)
                I[i] = make_int3(idx[3 * i], idx[3 * i + 1], idx[3 * i + 2]);
            auto *s = new lbvh::scene<3>(v.begin(), v.end(), I.begin(), I.end());
            s->compute_silhouettes();
            s->build_bvh();
            h->scene = s;
        }
        std::cout.rdbuf(old);
        return h;
    }

    void ref_scene_destroy(void *h)
    {
        AnyHandle *a = static_cast<AnyHandle *>(h);
        if (a->dim == 2)
            delete static_cast<lbvh::scene<2> *>(a->scene);
        else
            delete static_cast<lbvh::scene<3> *>(a->scene);
        delete a;
    }

    int ref_num_nodes(void *h)
    {
        AnyHandle *a = static_cast<AnyHandle *>(h);
        return a->dim == 2 ? (int)S2(h).bvh_dev.num_nodes : (int)S3(h).bvh_dev.num_nodes;
    }

    // nodes: num_nodes*4 u32 (parent,left,right,object); aabbs: num_nodes*2*dim floats (upper then lower);
    // cones: num_nodes*(dim+2) floats (axis, half_angle, radius)
    void ref_dump_bvh(void *h, uint32_t *nodes, float *aabbs, float *cones)
    {
        AnyHandle *a = static_cast<AnyHandle *>(h);
        if (a->dim == 2)
        {
            auto &b = S2(h).bvh_dev;
            for (unsigned i = 0; i < b.num_nodes; ++i)
            {
                nodes[4 * i + 0] = b.nodes[i].parent_idx;
                nodes[4 * i + 1] = b.nodes[i].left_idx;
                nodes[4 * i + 2] = b.nodes[i].right_idx;
                nodes[4 * i + 3] = b.nodes[i].object_idx;
                aabbs[4 * i + 0] = b.aabbs[i].upper.x;
                aabbs[4 * i + 1] = b.aabbs[i].upper.y;
                aabbs[4 * i + 2] = b.aabbs[i].lower.x;
                aabbs[4 * i + 3] = b.aabbs[i].lower.y;
                cones[4 * i + 0] = b.cones[i].axis.x;
                cones[4 * i + 1] = b.cones[i].axis.y;
                cones[4 * i + 2] = b.cones[i].half_angle;
                cones[4 * i + 3] = b.cones[i].radius;
            }
        }
        else
        {
            auto &b = S3(h).bvh_dev;
            for (unsigned i = 0; i < b.num_nodes; ++i)
            {
                nodes[4 * i + 0] = b.nodes[i].parent_idx;
                nodes[4 * i + 1] = b.nodes[i].left_idx;
                nodes[4 * i + 2] = b.nodes[i].right_idx;
                nodes[4 * i + 3] = b.nodes[i].object_idx;
                aabbs[6 * i + 0] = b.aabbs[i].upper.x;
                aabbs[6 * i + 1] = b.aabbs[i].upper.y;
                aabbs[6 * i + 2] = b.aabbs[i].upper.z;
                aabbs[6 * i + 3] = b.aabbs[i].lower.x;
                aabbs[6 * i + 4] = b.aabbs[i].lower.y;
                aabbs[6 * i + 5] = b.aabbs[i].lower.z;
                cones[5 * i + 0] = b.cones[i].axis.x;
                cones[5 * i + 1] = b.cones[i].axis.y;
                cones[5 * i + 2] = b.cones[i].axis.z;
                cones[5 * i + 3] = b.cones[i].half_angle;
                cones[5 * i + 4] = b.cones[i].radius;
            }
        }
    }

    // per primitive: dim silhouette indices (-1 = none)
    void ref_dump_silhouette_assignment(void *h, int *sil)
    {
        AnyHandle *a = static_cast<AnyHandle *>(h);
        if (a->dim == 2)
        {
            auto &s = S2(h);
            for (size_t i = 0; i < s.lines.size(); ++i)
            {
                sil[2 * i + 0] = s.lines[i].silhouette_indices.x;
                sil[2 * i + 1] = s.lines[i].silhouette_indices.y;
            }
        }
        else
        {
            auto &s = S3(h);
            for (size_t i = 0; i < s.triangles.size(); ++i)
            {
                sil[3 * i + 0] = s.triangles[i].silhouette_indices.x;
                sil[3 * i + 1] = s.triangles[i].silhouette_indices.y;
                sil[3 * i + 2] = s.triangles[i].silhouette_indices.z;
            }
        }
    }

    int ref_num_silhouettes(void *h)
    {
        AnyHandle *a = static_cast<AnyHandle *>(h);
        return a->dim == 2 ? (int)S2(h).silhouettes_h.size() : (int)S3(h).silhouettes_h.size();
    }
    // per silhouette record: 4 ints
    void ref_dump_silhouettes(void *h, int *rec)
    {
        AnyHandle *a = static_cast<AnyHandle *>(h);
        if (a->dim == 2)
        {
            auto &s = S2(h);
            for (size_t i = 0; i < s.silhouettes_h.size(); ++i)
            {
                int4 v = static_cast<const lbvh::scene<2>::silhouette_vertex &>(s.silhouettes_h[i]).indices;
                rec[4 * i + 0] = v.x, rec[4 * i + 1] = v.y, rec[4 * i + 2] = v.z, rec[4 * i + 3] = v.w;
            }
        }
        else
        {
            auto &s = S3(h);
            for (size_t i = 0; i < s.silhouettes_h.size(); ++i)
            {
                int4 v = static_cast<const lbvh::scene<3>::silhouette_edge &>(s.silhouettes_h[i]).indices;
                rec[4 * i + 0] = v.x, rec[4 * i + 1] = v.y, rec[4 * i + 2] = v.z, rec[4 * i + 3] = v.w;
            }
        }
    }

    // closest primitive: idx[n], dist[n]
    void ref_query_nearest(void *h, const float *pts, int n, uint32_t *idx, float *dist)
    {
        AnyHandle *a = static_cast<AnyHandle *>(h);
        parallel_for(n, [&](int i)
        {
            if (a->dim == 2)
            {
                auto r = lbvh::query_device(S2(h).bvh_dev, lbvh::nearest(make_float2(pts[2 * i], pts[2 * i + 1])),
                                            lbvh::scene<2>::distance_calculator());
                idx[i] = r.first;
                dist[i] = r.second;
            }
            else
            {
                auto r = lbvh::query_device(S3(h).bvh_dev, lbvh::nearest(make_float3(pts[3 * i], pts[3 * i + 1], pts[3 * i + 2])),
                                            lbvh::scene<3>::distance_calculator());
                idx[i] = r.first;
                dist[i] = r.second;
            }
        });
    }

    void ref_query_silhouette(void *h, const float *pts, int n, int flip, float *dist)
    {
        AnyHandle *a = static_cast<AnyHandle *>(h);
        parallel_for(n, [&](int i)
        {
            if (a->dim == 2)
                dist[i] = lbvh::query_device(S2(h).bvh_dev, lbvh::nearest_silhouette(make_float2(pts[2 * i], pts[2 * i + 1]), flip != 0),
                                             lbvh::scene<2>::silhouette_distance_calculator());
            else
                dist[i] = lbvh::query_device(S3(h).bvh_dev, lbvh::nearest_silhouette(make_float3(pts[3 * i], pts[3 * i + 1], pts[3 * i + 2]), flip != 0),
                                             lbvh::scene<3>::silhouette_distance_calculator());
        });
    }

    // closest-hit ray: hit[n] (0/1), t[n], uv[n*2] (2D: uv[2i] = s), obj[n]
    void ref_query_ray(void *h, const float *org, const float *dir, const float *max_dist, int n,
                       int *hit, float *t, float *uv, uint32_t *obj)
    {
        AnyHandle *a = static_cast<AnyHandle *>(h);
        parallel_for(n, [&](int i)
        {
            if (a->dim == 2)
            {
                lbvh::ray<float, 2> r(make_float2(org[2 * i], org[2 * i + 1]), make_float2(dir[2 * i], dir[2 * i + 1]));
                auto res = lbvh::query_device(S2(h).bvh_dev, lbvh::ray_intersect(r, max_dist[i]), lbvh::scene<2>::intersect_test());
                hit[i] = thrust::get<0>(res) ? 1 : 0;
                t[i] = thrust::get<1>(res);
                uv[2 * i] = hit[i] ? thrust::get<2>(res) : 0.0f;
                uv[2 * i + 1] = 0.0f;
                obj[i] = thrust::get<3>(res);
            }
            else
            {
                lbvh::ray<float, 3> r(make_float3(org[3 * i], org[3 * i + 1], org[3 * i + 2]), make_float3(dir[3 * i], dir[3 * i + 1], dir[3 * i + 2]));
                auto res = lbvh::query_device(S3(h).bvh_dev, lbvh::ray_intersect(r, max_dist[i]), lbvh::scene<3>::intersect_test());
                hit[i] = thrust::get<0>(res) ? 1 : 0;
                t[i] = thrust::get<1>(res);
                float2 q = thrust::get<2>(res);
                uv[2 * i] = hit[i] ? q.x : 0.0f;
                uv[2 * i + 1] = hit[i] ? q.y : 0.0f;
                obj[i] = thrust::get<3>(res);
            }
        });
    }

    // any-hit ray (TestOnly = true)
    void ref_query_ray_any(void *h, const float *org, const float *dir, const float *max_dist, int n, int *hit)
    {
        AnyHandle *a = static_cast<AnyHandle *>(h);
        parallel_for(n, [&](int i)
        {
            if (a->dim == 2)
            {
                lbvh::ray<float, 2> r(make_float2(org[2 * i], org[2 * i + 1]), make_float2(dir[2 * i], dir[2 * i + 1]));
                hit[i] = lbvh::query_device(S2(h).bvh_dev, lbvh::ray_intersect<true>(r, max_dist[i]), lbvh::scene<2>::intersect_test()) ? 1 : 0;
            }
            else
            {
                lbvh::ray<float, 3> r(make_float3(org[3 * i], org[3 * i + 1], org[3 * i + 2]), make_float3(dir[3 * i], dir[3 * i + 1], dir[3 * i + 2]));
                hit[i] = lbvh::query_device(S3(h).bvh_dev, lbvh::ray_intersect<true>(r, max_dist[i]), lbvh::scene<3>::intersect_test()) ? 1 : 0;
            }
        });
    }

    // Green-weighted single-path descent (core/sample.cuh:23-92): obj[n] (-1 = none), pdf[n]
    void ref_sample_in_sphere(void *h, const float *ctr, const float *radius, const float *u, int n, int *obj, float *pdf)
    {
        AnyHandle *a = static_cast<AnyHandle *>(h);
        for (int i = 0; i < n; ++i)
        {
            if (a->dim == 2)
            {
                auto r = lbvh::sample_object_in_sphere(S2(h).bvh_dev,
                                                       lbvh::sphere_intersect(lbvh::sphere<float, 2>(make_float2(ctr[2 * i], ctr[2 * i + 1]), radius[i])),
                                                       lbvh::scene<2>::intersect_sphere(), lbvh::scene<2>::measurement_getter(),
                                                       lbvh::scene<2>::green_weight(), u[i]);
                obj[i] = r.first;
                pdf[i] = r.first >= 0 ? r.second : 0.0f;
            }
            else
            {
                auto r = lbvh::sample_object_in_sphere(S3(h).bvh_dev,
                                                       lbvh::sphere_intersect(lbvh::sphere<float, 3>(make_float3(ctr[3 * i], ctr[3 * i + 1], ctr[3 * i + 2]), radius[i])),
                                                       lbvh::scene<3>::intersect_sphere(), lbvh::scene<3>::measurement_getter(),
                                                       lbvh::scene<3>::green_weight(), u[i]);
                obj[i] = r.first;
                pdf[i] = r.first >= 0 ? r.second : 0.0f;
            }
        }
    }
}
