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
    /* libwost.so parks released device blocks >= 1 MiB in a per-device free list (the reference never frees anything,
     * integrator/uniform/integrator.cu:25-60; here handles can be destroyed and re-created at no allocation cost).
     * This call returns every parked block to the driver. WOST_NO_ALLOC_CACHE=1 in the environment disables the cache. */
    wost_status wost_trim_allocation_cache(void);
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
    /* diagnostics of the B200 traversal layout (no reference counterpart): the fused internal-node records
     * (num_prims-1 records of *record_floats floats: both children's exact extents, child references and oriented boxes,
     * csrc/lbvh.cuh TNode) and the Morton-ordered leaf vertices (num_prims * dim * dim floats). Either pointer may be NULL. */
    wost_status wost_scene_dump_traversal_layout(const wost_scene *scene, float *records, int *record_floats, float *leaf_verts);
    /* Closest-point traversal statistics of the parity probes on this scene since the last reset: queries that ran a FULL
     * traversal in the reference's visit order (3D far field only) and queries that finished with the restricted
     * reference-order second pass (near-ties in 3D, and in 2D beyond |q| ~ 2000 where the reference's 1e-3 box padding no
     * longer dominates rounding; core/query.cuh:238-318). Everything else ran the B200 order alone. */
    wost_status wost_scene_traversal_stats(const wost_scene *scene, uint64_t *full_reference_order, uint64_t *restricted_second_pass, int reset);

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
    /* Closest-point query for hosts that must not allocate GPU memory. No counterpart in the reference (Elaina allocates
     * freely and never frees, integrator/uniform/integrator.cu:25-60); it exists for the embedders of this library that
     * cannot afford wost_lbvh_nearest's three staging buffers per call:
     *   - applications that hand the whole card to their own pool at start-up (renderers, DCC plug-ins, inference
     *     servers), where a later cudaMalloc from a plug-in fails or eats into a budget someone else planned;
     *   - frame loops and other latency-bound threads: cudaMalloc / cudaFree serialise against all work on the device, so
     *     one call stalls every stream of the process;
     *   - code running after set-up in a sandbox whose policy is "no allocation after initialisation".
     * Same arguments, same results and same error codes as wost_lbvh_nearest (host buffers in, host buffers out, idx may
     * be NULL), but the points and results are staged through a fixed array compiled into libwost.so instead of freshly
     * allocated blocks. Consequences of that design:
     *   - n is limited to WOST_NEAREST_STATIC_MAX_POINTS per call; a larger n is refused with WOST_ERR_INVALID and a
     *     message naming the limit (split the batch, or use wost_lbvh_nearest). The array costs about 1.25 MiB of device
     *     memory on every GPU the process loads the library on, for the lifetime of the process;
     *   - the scene's device must be the current one (wost_set_device); otherwise WOST_ERR_INVALID;
     *   - the first call on each device also looks up the array's device address (cudaGetSymbolAddress) and makes the
     *     CUDA runtime load the kernel; make that first call during set-up, before any stream capture begins;
     *   - the call itself invokes no cudaMalloc, cudaFree or allocating library routine. It runs on the default stream and
     *     returns after the results are in `idx` / `dist`;
     *   - all calls on one device share the one array, which keeps the previous call's contents (the control words are
     *     zeroed at the start of each call). Two calls in flight on the same GPU at once (two host threads, whether on
     *     one scene or on different scenes) would overwrite each other's points and results: serialise them;
     *   - its traversals are not added to wost_scene_traversal_stats (those counters live in the scene's own block). */
#define WOST_NEAREST_STATIC_MAX_POINTS 65536
    wost_status wost_lbvh_nearest_static(const wost_scene *scene, const float *pts, int n, uint32_t *idx, float *dist);
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
        /* guided extensions (0 = the reference's constants) */
        int mlp_impl;                /* 0: fused tcgen05 inference inside the bounce kernel; 1: validation mode, SIMT MLP kernel
                                        writing the raw outputs to HBM like the reference (guided/integrator.cu:528-563) */
        int max_train_depth;         /* Guidance::maxTrainDepth, default 3 (guided/integrator.h:222), <= 4 */
        int train_batch_size;        /* TRAIN_BATCH_SIZE 524288 (guided/parameters.h:12) */
        int train_min_batch;         /* MIN_TRAIN_BATCH_SIZE 65536 (guided/parameters.h:13) */
        int train_batches_per_frame; /* Guidance::batchPerFrame 5 (guided/integrator.h:223) */
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
    /* Standard error of the solution film of the last solve, per pixel and colour channel (W*H*3 floats, 0 for pixels this
     * shard does not evaluate). Measured, not modelled: the accumulators keep one running sum per (slot, pixel) where a slot
     * receives every spp_batch-th sample, so se^2 = sum_s n_s (m_s - m)^2 / ((S - 1) spp) over the slot means m_s; with
     * spp_batch == samples_per_pixel every slot holds ONE sample and this is the exact sample variance / spp. The reference
     * has no such output (its PixelState keeps only the running sum, integrator/uniform/workqueue.h:15-30); the statistical
     * parity tests (|delta| <= 3 standard errors per pixel) need it. */
    wost_status wost_film_stderr(wost_integrator *it, float *out_se_rgb);
    /* device pointer of the film of `channel` (W*H*5 floats) for NCCL gathers issued by the host language */
    wost_status wost_film_device_ptr(wost_integrator *it, int channel, void **d_ptr);
    /* the same statistics for the closest-point queries of this integrator's solves and channel renders */
    wost_status wost_traversal_stats(wost_integrator *it, uint64_t *full_reference_order, uint64_t *restricted_second_pass, int reset);
    /* diagnostics, meaningful only when the library was built with -DWOST_TRAV_DEBUG. The caller provides 48 words:
     * closest-point trips [0..7] and queries [8..15] of this integrator per decade of the query's max |coordinate|
     * (<600, <2e3, <1e4, ..., >=1e8), then [16..47] = 16 little-endian 64-bit sums of warp cycles (8 decades of trip iterations,
     * 8 of tail iterations); read and reset */
    wost_status wost_debug_counters(wost_integrator *it, unsigned int *out48);
    /* bytes of top-level LBVH records this integrator pinned in L2 (cudaAccessPolicyWindow on its stream; 0 when the tree
     * fits the L2 or WOST_L2_PERSIST=0) -- north_star (2), no reference counterpart (the reference leaves L2 to the hardware) */
    wost_status wost_l2_window_bytes(wost_integrator *it, uint64_t *bytes);
    /* counters of the last solve */
    wost_status wost_last_launch_count(wost_integrator *it, uint64_t *kernel_launches);
    /* launches of the fused walk-step kernel in the last solve and their summed device time (CUDA events on the
     * integrator's stream around each spp batch's run of step launches) -- the roofline numerator of bench.py */
    wost_status wost_last_step_kernel_time(wost_integrator *it, uint64_t *step_launches, float *step_kernel_ms);

    /* ------------------------------------------------------------------ configuration runner
     * run_expr(conf_path) (exec.cu:39-221): parse an elaina-exec configuration file, build problem and integrator, run
     * the requested channels, export images, write <base_path>/<exp_name>/{conf.json,result.json,...}. */
    wost_status wost_run_config(const char *json_path);

    /* ------------------------------------------------------------------ guided integrator: the guiding network
     * resetNetwork(json) (integrator/guided/integrator.cu:1096-1137): `tcnn_json` is the "network" block of the
     * configuration file ({encoding{otype DenseGrid,n_levels,n_features_per_level 4,base_resolution,per_level_scale},
     * network{FullyFusedMLP, 64 neurons, 3 hidden layers, ReLU}, optimizer{Ema{decay, nested Adam{...}}}}). */
    wost_status wost_reset_network(wost_integrator *it, const char *tcnn_json);
    /* resetTraining() (integrator.cu:1139-1146): re-initialise parameters and optimiser state */
    wost_status wost_reset_training(wost_integrator *it);
    /* queryNetwork(p) (integrator.cu:565-615): raw network output at a world-space point, 33 (2D) / 41 (3D) floats */
    wost_status wost_query_network(wost_integrator *it, const float *p, float *out);
    wost_status wost_guided_network_info(wost_integrator *it, int *n_params, int *n_mlp_params, int *encoded_width);
    /* parameter slab [W0|W1|W2|W3|grid] in fp32; which = 0 training weights, 1 EMA ("inference") weights */
    wost_status wost_guided_get_params(wost_integrator *it, int which, float *params);
    wost_status wost_guided_set_params(wost_integrator *it, const float *params);
    /* parity probes. infer: raw outputs (n*48, padded) at world-space points; impl 0 = tcgen05 tile kernel, 1 = SIMT.
     * train_batch: ONE optimiser step on the given samples (n rounded down to a multiple of 128) = forward, KL gradient
     * (compute_dL_doutput_divergence, guided/train.h:491-553), backward, Adam + EMA; grad_out (n_params floats or NULL)
     * receives the loss-scaled gradient slab of the step, loss the mean per-sample loss. */
    wost_status wost_guided_infer(wost_integrator *it, const float *pts, int n, int impl, int which_weights, float *out);
    wost_status wost_guided_train_batch(wost_integrator *it, const float *pos, const float *dir, const float *nrm, const float *dir_pdf,
                                        const float *li, const uint8_t *on_neumann, int n, float *loss, float *grad_out);
    /* VMM<DIM,8> probe (guided/distribution.h:136-444) evaluated on the device: pdf (n) of the mixture decoded from raw
     * (n * (8*(dim+2)+1)) at dir (+ reflected dir when on_neumann), grad (n * 8*(dim+2)) of that pdf w.r.t. the
     * activated lobe parameters, and one sampled direction (n*dim) drawn with Philox key (seed, i). */
    wost_status wost_vmm_eval(int dim, const float *raw, const float *dir, const float *nrm, const uint8_t *on_neumann, int n, uint32_t seed,
                              float *pdf, float *grad, float *sampled_dir);
    /* Multi-GPU training: called on the integrator's stream between the backward pass and the optimiser step of every
     * batch with the device gradient slab (n_floats floats, last word = number of ranks whose batch was valid). The
     * callback must enqueue a SUM all-reduce on `stream` (a cudaStream_t) and return 0. */
    typedef int (*wost_allreduce_fn)(void *user, void *d_buf, uint64_t n_floats, void *stream);
    wost_status wost_set_allreduce(wost_integrator *it, wost_allreduce_fn fn, void *user);

    /* ------------------------------------------------------------------ multi-GPU: NCCL inside the library
     * The reference is single-device (core/device/context.cpp:23-24 hard-codes device 0), so these entry points have no
     * counterpart to mirror; they are what a multi-GPU exec.cu needs (north_star: evaluation points shard across the GPUs
     * of one box, NCCL only for the gradient all-reduce of the replicated guiding network and the final grid gather).
     * One process per GPU. libnccl.so.2 is resolved at run time (dlopen; WOST_NCCL_LIB overrides the search). */
    typedef struct wost_comm wost_comm;
    /* rank 0: 128-byte NCCL unique id, to be handed to the other ranks by the launcher (torch.distributed broadcast, MPI, file) */
    wost_status wost_comm_unique_id(void *id128);
    /* collective over `world` processes; uses the calling thread's current CUDA device */
    wost_status wost_comm_create(int rank, int world, const void *id128, wost_comm **out);
    /* torchrun / mpirun environment (RANK, WORLD_SIZE, LOCAL_RANK, MASTER_PORT | OMPI_COMM_WORLD_*): selects device LOCAL_RANK and
     * exchanges the id through a file in $TMPDIR named after the launcher. *out = NULL (and WOST_OK) for a single process. */
    wost_status wost_comm_create_from_env(wost_comm **out);
    void wost_comm_destroy(wost_comm *comm);
    wost_status wost_comm_info(const wost_comm *comm, int *rank, int *world, int *nccl_version);
    /* fills pixel_stride_rows / num_shards / shard of `settings` for this rank (interleaved tiles of stride_rows rows) */
    wost_status wost_comm_shard_settings(const wost_comm *comm, wost_settings *settings, int stride_rows);
    /* guided integrator: ncclAllReduce(sum) of the gradient slab on the integrator's stream before every optimiser step */
    wost_status wost_integrator_set_comm(wost_integrator *it, wost_comm *comm);
    /* final grid gather: the row tiles every rank evaluated are collected into the device film of rank `root` (root < 0: of
     * every rank) -- (N-1)/N of one film over NVLink, packed rows, grouped send/recv or one all-gather */
    wost_status wost_film_gather(wost_integrator *it, wost_comm *comm, int channel, int root);
    /* host scalars across ranks: sum (op_max = 0) or max (op_max = 1) */
    wost_status wost_comm_allreduce_f64(wost_comm *comm, double *vals, int n, int op_max);

    /* NanoVDB source grids (core/problem.cu:135-149 reads them with nanovdb::io::readGrid; the walker samples them with
     * worldToIndex + SampleFromVoxels<Vec3fTree, 1, false>, integrator/uniform/integrator.cu:309-311). Host-side: flattens
     * the first grid of an uncompressed .nvdb file (Vec3f, axis-aligned uniform map, zero background) over its index
     * bounding box into the dense grid wost_problem_desc takes. Call with dense == NULL to get dims / origin / voxel, then
     * with a buffer of dims[0]*dims[1]*dims[2]*3 floats ([z][y][x][rgb]). */
    wost_status wost_nvdb_read_vec3f(const char *path, int dims[3], float origin[3], float *voxel, float *dense, uint64_t capacity_floats);

    /* Progressive metrics (integrator/uniform/integrator.cu:577-609, "frames/<sampleId>" and "frames_time/<ms>"): when
     * settings.save_spp_metrics_duration > 0 or save_time_metrics_duration > 0, wost_solve() hands the running estimate
     * solution / (sample_id + 1) to this callback after every sample index the reference would dump
     * (sample_id % duration == 0 [and sample_id < save_spp_metrics_until for kind 0]). kind: 0 = spp metric, 1 = time
     * metric. film: frame_w*frame_h*5 floats {r,g,b,a,weight}, host memory owned by the library, valid during the call.
     * The solver cuts its sample batches at those indices, so a dump costs one stream synchronisation, as in the reference. */
    typedef void (*wost_frame_fn)(void *user, int kind, int sample_id, uint64_t elapsed_ms, const float *film, int frame_w, int frame_h);
    wost_status wost_set_frame_callback(wost_integrator *it, wost_frame_fn fn, void *user);
    /* with WOST_PROFILE=1 in the environment: device ms of the last guided solve by kernel class
     * [0] arrive (closest point + contributions), [1] bounce (MLP + sampling + move), [2] finish (uniform tail),
     * [3] gaps between classes, [4] training frames (samples + forward/backward + optimiser), [5..7] unused */
    wost_status wost_guided_last_profile(wost_integrator *it, float *ms8);
    /* parity probe: the compacted training set of the last training frame (generate_training_data, integrator/guided/train.h:422-471),
     * one row per sample: pos[dim], dir[dim], nN[dim], dirPdf, Li, onN -> 3 * dim + 3 floats. rows == NULL: only *n_rows. */
    wost_status wost_guided_last_samples(wost_integrator *it, float *rows, uint64_t capacity_rows, uint64_t *n_rows);
    /* counters of the last guided solve: training samples consumed, optimiser steps, mean loss of the last batch */
    wost_status wost_guided_last_stats(wost_integrator *it, uint64_t *train_samples, uint64_t *train_steps, float *last_loss);

#ifdef __cplusplus
}
#endif
#endif /* WOST_H */
