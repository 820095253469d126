// Host-side problem object behind `wost_problem` + the POD the walk kernels receive by value.
#pragma once
#include "scene.cuh"

namespace wost
{
    // Everything a walk step needs, passed to kernels by value (lives in constant/param space).
    template <int D>
    struct WalkCtx
    {
        DevBVH<D> dir, neu;         // nprims == 0 when disabled
        const int *dir_prims;       // object -> vertex indices (colour lookup)
        const int *neu_prims;
        const float *color_dir;     // nverts*6 (left rgb, right rgb) or nullptr = zeros
        const float *color_neu;
        float dirichlet_intensity, neumann_intensity, source_intensity;
        int neumann_flux_enabled;   // 0: every Neumann colour is zero -> the boundary-sample stage cannot change the result
        const float *source;        // dense Vec3f grid or nullptr
        int source_dims[3];
        float source_origin[3];
        float source_inv_voxel;
        float grid_scale;
        float grid_pos[3], grid_up[3], grid_right[3];
        const uint8_t *mask;        // frame_w*frame_h or nullptr
        int frame_w, frame_h;
        int max_depth;
        float eps;
        uint32_t seed;
        float aabb_min[3], aabb_max[3];
    };
}

struct wost_problem
{
    int dim = 0;
    int device = 0;
    wost_scene *dirichlet = nullptr; // not owned
    wost_scene *neumann = nullptr;   // not owned
    wost::DevBuf<float> d_color_dir, d_color_neu, d_source;
    wost::DevBuf<uint8_t> d_mask;
    int mask_w = 0, mask_h = 0;
    bool neumann_flux_enabled = false;
    wost_problem_desc desc; // host copy (pointers inside are NOT valid after create)

    template <int D>
    void fill_ctx(wost::WalkCtx<D> &c) const
    {
        c.dir = dirichlet ? dirichlet->view<D>() : wost::empty_bvh<D>();
        c.neu = neumann ? neumann->view<D>() : wost::empty_bvh<D>();
        c.dir_prims = dirichlet ? dirichlet->d_prims.ptr : nullptr;
        c.neu_prims = neumann ? neumann->d_prims.ptr : nullptr;
        c.color_dir = d_color_dir.ptr;
        c.color_neu = d_color_neu.ptr;
        c.dirichlet_intensity = desc.dirichlet_intensity;
        c.neumann_intensity = desc.neumann_intensity;
        c.source_intensity = desc.source_intensity;
        c.neumann_flux_enabled = neumann_flux_enabled ? 1 : 0;
        c.source = d_source.ptr;
        for (int k = 0; k < 3; ++k)
        {
            c.source_dims[k] = desc.source_dims[k];
            c.source_origin[k] = desc.source_origin[k];
            c.grid_pos[k] = desc.grid_pos[k];
            c.grid_up[k] = desc.grid_up[k];
            c.grid_right[k] = desc.grid_right[k];
            c.aabb_min[k] = desc.aabb_min[k];
            c.aabb_max[k] = desc.aabb_max[k];
        }
        c.source_inv_voxel = d_source.ptr ? 1.0f / desc.source_voxel : 0.0f;
        c.grid_scale = desc.grid_scale;
        c.mask = d_mask.ptr;
    }
};
