// Host-side scene object behind the opaque `wost_scene` handle.
#pragma once
#include <algorithm>
#include <cmath>
#include "common.cuh"
#include "lbvh.cuh"

struct wost_scene
{
    int dim = 0;
    int device = 0;
    int nverts = 0, nprims = 0, nsil = 0;
    bool used_morton64 = false;
    std::vector<float> h_verts; // nverts*dim
    std::vector<int> h_prims;   // nprims*dim
    // inputs on device
    wost::DevBuf<float> d_verts;
    wost::DevBuf<int> d_prims, d_sil_rec, d_prim_sil;
    // canonical LBVH (reference layout) -- parity probes + source of the traversal layout
    wost::DevBuf<wost::Node> d_nodes;
    wost::DevBuf<float> d_boxes; // num_nodes * 2*dim (up, lo)
    wost::DevBuf<float> d_cones; // num_nodes * (dim+2)
    wost::DevBuf<uint32_t> d_morton, d_order;
    // traversal layout
    wost::DevBuf<char> d_tnodes, d_cnodes, d_leaf_prim, d_leaf_sil;
    wost::DevBuf<uint32_t> d_leaf_obj, d_leaf_rank;
    wost::DevBuf<int> d_overflow;
    float root_box[6] = {0, 0, 0, 0, 0, 0};
    float root_cone[5] = {0, 0, 0, 0, 0};
    float root_ext[6] = {0, 0, 0, 0, 0, 0}; // exact (unpadded) extents of the whole mesh
    int top_nodes = 0, top_depth = 0;       // the first top_nodes fused records are the levels [0, top_depth) in breadth-first order
    size_t tnode_bytes() const { return (size_t)std::max(0, nprims - 1) * (dim == 2 ? sizeof(wost::TNode<2>) : sizeof(wost::TNode<3>)); }
    size_t top_bytes() const { return (size_t)top_nodes * (dim == 2 ? sizeof(wost::TNode<2>) : sizeof(wost::TNode<3>)); }

    template <int D>
    wost::DevBVH<D> view() const
    {
        wost::DevBVH<D> b;
        b.nprims = nprims;
        b.tnodes = reinterpret_cast<const wost::TNode<D> *>(d_tnodes.ptr);
        b.cnodes = reinterpret_cast<const wost::CNode<D> *>(d_cnodes.ptr);
        b.leaf_prim = reinterpret_cast<const wost::LeafPrim<D> *>(d_leaf_prim.ptr);
        b.leaf_obj = d_leaf_obj.ptr;
        b.leaf_rank = d_leaf_rank.ptr;
        b.leaf_sil = reinterpret_cast<const wost::LeafSil<D> *>(d_leaf_sil.ptr);
        for (int k = 0; k < 2 * D; ++k)
            b.root_box[k] = root_box[k];
        for (int k = 0; k < D + 2; ++k)
            b.root_cone[k] = root_cone[k];
        float m = 0.0f;
        for (int k = 0; k < 2 * D; ++k)
            m = std::max(m, std::fabs(root_ext[k]));
        b.max_abs_coord = m;
        return b;
    }
};

namespace wost
{
    constexpr size_t kL2WindowBudget = 32u << 20; // bytes of top-level records kept contiguous (and pinned in L2 for trees beyond L2)
    template <int D>
    inline DevBVH<D> empty_bvh()
    {
        DevBVH<D> b;
        b.nprims = 0;
        b.tnodes = nullptr;
        b.cnodes = nullptr;
        b.leaf_prim = nullptr;
        b.leaf_obj = nullptr;
        b.leaf_rank = nullptr;
        b.max_abs_coord = 0.0f;
        b.leaf_sil = nullptr;
        for (int k = 0; k < 2 * D; ++k)
            b.root_box[k] = 0.0f;
        for (int k = 0; k < D + 2; ++k)
            b.root_cone[k] = 0.0f;
        return b;
    }
}
