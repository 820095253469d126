/* wost.h -- C ABI of the B200-native walk-on-stars (WoSt) walker.
 *
 * Drop-in boundary for the walker hot path of tyanyuy3125/elaina (paths below are relative to the
 * reference tree). Elaina has no FFI of its own: its integrators are C++ class templates selected by a
 * std::variant in exec.cu:77-78,145-215. Each entry point here names the reference interface it replaces.
 * Plain pointers and sizes only; no C++/torch types cross this boundary; nothing throws or exits.
 * All "host" buffers are caller-owned; the library owns every device allocation behind the handles.
 * One handle set per GPU (wost_set_device selects the CUDA device of the calling thread's next creations).
 */
#ifndef WOST_H
#define WOST_H
#include <stdint.h>
#ifdef __cplusplus
extern "C"
{
#endif

    typedef enum wost_status
    {
        WOST_OK = 0,
        WOST_ERR_INVALID = 1,     /* bad argument */
        WOST_ERR_CUDA = 2,        /* CUDA runtime failure (message in wost_last_error) */
        WOST_ERR_IO = 3,          /* file not found / unreadable (reference: std::runtime_error, scene_loader.cuh:52-55) */
        WOST_ERR_CONFIG = 4,      /* missing JSON key (reference: json_get_or_throw, core/common.h:152-171) */
        WOST_ERR_UNSUPPORTED = 5, /* unknown integrator / dimensionality (reference: exit(1), exec.cu:96-100,125-129) */
        WOST_ERR_OVERFLOW = 6     /* traversal stack overflow (reference: silent memory corruption past stack[64]) */
    } wost_status;

    const char *wost_last_error(void);
    const char *wost_version(void);
    /* replaces Context::initialize (core/device/context.cpp:12-39, device 0 hard-coded there) */
    wost_status wost_set_device(int device);
    int wost_device_count(void);

    /* ------------------------------------------------------------------ geometry: one mesh + its SNCH-LBVH
     * replaces lbvh::scene<DIM>(verts, idx) + compute_silhouettes() + build_bvh()
     * (ext/lbvh/include/snch_lbvh/scene.cuh:624-676, 1128-1229; core/bvh.cuh:380-613), called from
     * core/problem.cu:21-61. verts: nverts*dim floats; prims: nprims*dim 0-based vertex indices. */
    typedef struct wost_scene wost_scene;
    wost_status wost_scene_create(int dim, const float *verts, int nverts, const int *prims, int nprims, wost_scene **out);
    void wost_scene_destroy(wost_scene *scene);
    wost_status wost_scene_info(const wost_scene *scene, int *dim, int *num_prims, int *num_nodes, int *used_morton64);
    /* parity probes: the canonical LBVH arrays in the reference's own layout (core/bvh.cuh:27-33,48-54):
     * nodes num_nodes*4 u32 (parent,left,right,object), aabbs num_nodes*2*dim (upper,lower), cones num_nodes*(dim+2) */
    wost_status wost_scene_dump_bvh(const wost_scene *scene, uint32_t *nodes, float *aabbs, float *cones);
    wost_status wost_scene_dump_sorted(const wost_scene *scene, uint32_t *morton, uint32_t *object_idx);

    /* lbvh::query_device(bvh, nearest(q), distance_calculator) -- core/query.cuh:238-318.
     * Host buffers; pts n*dim. idx = closest primitive, dist = distance (the star-radius input R_D). */
    wost_status wost_lbvh_nearest(const wost_scene *scene, const float *pts, int n, uint32_t *idx, float *dist);
    /* lbvh::query_device(bvh, nearest_silhouette(q, flip), silhouette_distance_calculator) -- core/query.cuh:325-423 */
    wost_status wost_lbvh_silhouette(const wost_scene *scene, const float *pts, int n, int flip, float *dist);
    /* lbvh::query_device(bvh, ray_intersect<false>(ray, max_dist), intersect_test) -- core/query.cuh:79-169.
     * uv: n*2 (2D: uv[2i] = segment parameter). */
    wost_status wost_lbvh_ray(const wost_scene *scene, const float *org, const float *dir, const float *max_dist, int n,
                              int *hit, float *t, float *uv, uint32_t *obj);
    /* ray_intersect<true>: any hit */
    wost_status wost_lbvh_ray_any(const wost_scene *scene, const float *org, const float *dir, const float *max_dist, int n, int *hit);
    /* lbvh::sample_object_in_sphere -- core/sample.cuh:23-92. obj = -1 when nothing was selected. */
    wost_status wost_lbvh_sample_sphere(const wost_scene *scene, const float *ctr, const float *radius, const float *u, int n,
                                        int *obj, float *pdf);
    /* Device-resident variants used by bench.py (all pointers are device pointers on the scene's GPU;
     * launched on `stream`, a cudaStream_t passed as void*; no synchronisation). */
    wost_status wost_lbvh_nearest_dev(const wost_scene *scene, const float *d_pts, int n, uint32_t *d_idx, float *d_dist, void *stream);
    wost_status wost_lbvh_silhouette_dev(const wost_scene *scene, const float *d_pts, int n, int flip, float *d_dist, void *stream);
    wost_status wost_lbvh_ray_dev(const wost_scene *scene, const float *d_org, const float *d_dir, const float *d_max_dist, int n,
                                  int *d_hit, float *d_t, float *d_uv, uint32_t *d_obj, void *stream);

    /* ------------------------------------------------------------------ problem: replaces Problem<DIM>::loadConfig
     * (core/problem.cu:151-213): meshes, per-vertex left/right colour pairs (core/problem.cu:63-133), dense Vec3f
     * source grid (stands in for the NanoVDB grid of core/problem.cu:135-149), mask (:215-249), intensities,
     * AABB and evaluation grid (core/evaluation_grid.h). */
    typedef struct wost_problem wost_problem;
    typedef struct wost_problem_desc
    {
        int dim;
        wost_scene *dirichlet; /* may be NULL */
        wost_scene *neumann;   /* may be NULL */
        const float *color_dirichlet; /* nverts_d*6 (left rgb, right rgb); NULL = zeros (core/problem.cu:113-117) */
        int n_color_dirichlet;
        const float *color_neumann;
        int n_color_neumann;
        float dirichlet_intensity, neumann_intensity, source_intensity;
        const float *source; /* nx*ny*nz*3 floats, x fastest; NULL = no source term */
        int source_dims[3];
        float source_origin[3];
        float source_voxel;
        float grid_scale; /* EvaluationGrid::ProbeData (core/evaluation_grid.h:17-24,49-57) */
        float grid_pos[3];
        float grid_up[3];
        float grid_right[3];
        const uint8_t *mask; /* mask_w*mask_h, nonzero = evaluate; NULL = all pixels */
        int mask_w, mask_h;
        float aabb_min[3], aabb_max[3];
    } wost_problem_desc;
    wost_status wost_problem_create(const wost_problem_desc *desc, wost_problem **out);
    void wost_problem_destroy(wost_problem *problem);

    /* ------------------------------------------------------------------ integrators
     * replaces UniformIntegrator<DIM> (integrator/uniform/integrator.h:54-131) and GuidedIntegrator<DIM>
     * (integrator/guided/integrator.h:95-253): ctor(Problem&, settings, basePath), solve(), render*SDF(),
     * renderSource(), exportImage(), resetNetwork(), queryNetwork(). */
    typedef enum wost_integrator_kind
    {
        WOST_UNIFORM = 0,
        WOST_GUIDED = 1
    } wost_integrator_kind;
    typedef enum wost_channel /* ExportImageChannel (util/film.h / exec.cu:131-143) */
    {
        WOST_CH_SOLUTION = 0,
        WOST_CH_DIRICHLET_SDF = 1,
        WOST_CH_NEUMANN_SDF = 2,
        WOST_CH_SOURCE = 3,
        WOST_CH_COUNT = 4
    } wost_channel;
    typedef enum wost_walk_mode
    {
        WOST_MODE_DEFAULT = 0,        /* = WOST_MODE_PERSISTENT */
        WOST_MODE_PERSISTENT = 1,     /* register-resident walks: persistent lanes pull (pixel, sample) jobs, one launch per spp batch */
        WOST_MODE_REFERENCE_ORDER = 2, /* wavefront kernel, closest-point queries in the reference's visit order (core/query.cuh:238-318):
                                          identical results; the on-device stand-in for the reference's traversal cost */
        WOST_MODE_WAVEFRONT = 3,      /* one fused kernel per bounce over compacted SoA queues */
        WOST_MODE_PERSISTENT_REFERENCE_ORDER = 4 /* persistent kernel with the reference's visit order */
    } wost_walk_mode;

    typedef struct wost_settings /* IntegratorSettings: integrator/uniform/integrator.h:28-49, guided/integrator.h:56-90 */
    {
        int frame_w, frame_h;
        int max_walking_depth;
        int samples_per_pixel;
        float epsilon_shell;
        int debug_pixel;
        int save_spp_metrics_duration, save_spp_metrics_until, save_time_metrics_duration;
        /* guided only */
        int train_spp_count;
        int max_guided_depth_training, max_guided_depth_guiding;
        float uniform_fraction_training, uniform_fraction_guiding;
        /* extensions (0 = defaults) */
        uint32_t seed;
        int walk_mode;       /* wost_walk_mode */
        int spp_batch;       /* samples of every pixel in flight per wavefront (0 = auto) */
        int pixel_begin, pixel_end; /* shard of the frame owned by this handle; pixel_end = 0 -> whole frame */
        int pixel_stride_rows;      /* >0: interleaved row tiles: rows r with (r / stride_rows) % num_shards == shard */
        int num_shards, shard;
    } wost_settings;

    typedef struct wost_integrator wost_integrator;
    wost_status wost_integrator_create(wost_problem *problem, const wost_settings *settings, int kind, wost_integrator **out);
    void wost_integrator_destroy(wost_integrator *it);
    /* solve(): returns wall-clock ms like the reference (integrator/uniform/integrator.cu:666-672) plus the number of
     * walk-steps (items consumed by the separate stage, integrator.cu:128) and the device time in ms (CUDA events). */
    wost_status wost_solve(wost_integrator *it, uint64_t *ms, uint64_t *walk_steps, float *device_ms);
    /* renderDirichletSDF / renderSilhouetteSDF / renderSource (integrator/common.h:51-163) */
    wost_status wost_render_channel(wost_integrator *it, int channel);
    /* Film (util/film.h:16-155): per pixel {r,g,b,a, weight}, index x + y*W, NOT flipped (files are flipped on save only) */
    wost_status wost_film_download(wost_integrator *it, int channel, float *out_rgbaw);
    /* device pointer of the film of `channel` (W*H*5 floats) for NCCL gathers issued by the host language */
    wost_status wost_film_device_ptr(wost_integrator *it, int channel, void **d_ptr);
    /* counters of the last solve */
    wost_status wost_last_launch_count(wost_integrator *it, uint64_t *kernel_launches);
    /* launches of the fused walk-step kernel in the last solve and their summed device time (CUDA events on the
     * integrator's stream around each spp batch's run of step launches) -- the roofline numerator of bench.py */
    wost_status wost_last_step_kernel_time(wost_integrator *it, uint64_t *step_launches, float *step_kernel_ms);

#ifdef __cplusplus
}
#endif
#endif /* WOST_H */
