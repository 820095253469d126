/* TEST INFRASTRUCTURE ONLY -- the scalar CPU oracle. Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load this library; the product never does.
 *
 * A restatement (not a copy) of the reference algorithm for the walker hot path:
 *   - SNCH-LBVH build            ext/lbvh/include/snch_lbvh/core/bvh.cuh:380-613, scene.cuh:420-502,631-676,870-961,1135-1229
 *   - closest point / silhouette / ray / sphere-sample queries
 *                                core/query.cuh:79-169,238-423, core/sample.cuh:23-92
 *   - uniform WoSt integrator    integrator/uniform/integrator.cu:64-623
 *   - guided sampling math       integrator/guided/distribution.h, util/vonmises.h, util/vmf.h
 * Parity is PINNED: in math mode 1 (libm) the oracle reproduces the host-compiled reference LBVH
 * (oracle/_ref/libref_lbvh.so) bit for bit on nodes, boxes, cones and every query, and the golden
 * vectors of SURVEY.md section 8(c) (tests/golden/lbvh_golden.json) hold in both modes.
 * In math mode 0 (detmath, default) transcendental functions come from elaina_b200/csrc/detmath.h,
 * the arithmetic the CUDA product uses, so product == oracle bit for bit.
 */
#ifndef WOST_ORACLE_H
#define WOST_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C"
{
#endif

    void orc_set_math_mode(int mode); /* 0 = detmath (default), 1 = libm (matches oracle/_ref) */
    /* uniform walker draws: 0 = Philox stream of the product (default); 1 = the reference's per-pixel PCG32 in its
     * consumption order (core/sampler.h:10-103, integrator/uniform/integrator.cu:72-77), to compare orc_solve_uniform
     * field-for-field with the reference's own CUDA build (tests/golden/b1_*.npz) */
    void orc_set_rng_mode(int mode);
    void orc_set_pixel_steps(uint32_t *buf); /* diagnostics: per-pixel walk-steps of orc_solve_uniform (W*H uint32), NULL = off */
    void orc_set_trace(float *buf); /* diagnostics: walk positions [depth][dim] of a single-pixel, 1-spp, 1-thread solve */
    void orc_pcg32_probe(uint64_t initstate, uint64_t initseq, int64_t delta, int n, uint32_t *out); /* PCG32 known answers */

    /* ---- geometry ---- */
    void *orc_scene_create(int dim, const float *verts, int nverts, const int *idx, int nprims);
    void orc_scene_destroy(void *scene);
    int orc_num_nodes(void *scene);
    int orc_used_morton64(void *scene);
    /* nodes: num_nodes*4 u32 (parent,left,right,object); aabbs: num_nodes*2*dim (upper, lower); cones: num_nodes*(dim+2) */
    void orc_dump_bvh(void *scene, uint32_t *nodes, float *aabbs, float *cones);
    /* per leaf (sorted order): morton code and original object index */
    void orc_dump_sorted(void *scene, uint32_t *morton, uint32_t *object_idx);
    void orc_dump_silhouette_assignment(void *scene, int *sil);
    int orc_num_silhouettes(void *scene);
    void orc_dump_silhouettes(void *scene, int *rec);

    void orc_query_nearest(void *scene, const float *pts, int n, uint32_t *idx, float *dist);
    void orc_query_silhouette(void *scene, const float *pts, int n, int flip, float *dist);
    void orc_query_ray(void *scene, const float *org, const float *dir, const float *max_dist, int n,
                       int *hit, float *t, float *uv, uint32_t *obj);
    void orc_query_ray_any(void *scene, const float *org, const float *dir, const float *max_dist, int n, int *hit);
    void orc_sample_in_sphere(void *scene, const float *ctr, const float *radius, const float *u, int n, int *obj, float *pdf);

    /* ---- problem + uniform walker ---- */
    typedef struct orc_problem_desc
    {
        int dim;
        void *dirichlet; /* orc scene or NULL */
        void *neumann;   /* orc scene or NULL */
        const float *color_dirichlet; /* nverts_d * 6 (left rgb, right rgb) or NULL (zeros) */
        const float *color_neumann;   /* nverts_n * 6 or NULL */
        float dirichlet_intensity, neumann_intensity, source_intensity;
        /* dense source grid (Vec3f), index = (world - origin) / voxel; NULL = no source */
        const float *source; /* nx*ny*nz*3, x fastest */
        int source_dims[3];
        float source_origin[3];
        float source_voxel;
        /* evaluation grid (core/evaluation_grid.h) */
        float grid_scale;
        float grid_pos[3];
        float grid_up[3];
        float grid_right[3];
        const uint8_t *mask; /* W*H or NULL (all true) */
        float aabb_min[3], aabb_max[3];
    } orc_problem_desc;

    typedef struct orc_settings
    {
        int frame_w, frame_h;
        int max_walking_depth;
        int samples_per_pixel;
        float epsilon_shell;
        uint32_t seed;
    } orc_settings;

    /* Uniform WoSt over pixels [pixel_begin, pixel_end) with the Philox stream of elaina_b200/csrc/philox.h.
     * solution: W*H*3 floats (sum over spp divided by spp); steps: walk-steps consumed (items through "separate"). */
    void orc_solve_uniform(const orc_problem_desc *p, const orc_settings *s, int pixel_begin, int pixel_end,
                           float *solution, uint64_t *steps, int nthreads);

    /* ---- guided integrator (wost_oracle_guided.inl) ---- */
    typedef struct orc_guided_settings
    {
        int max_guided_depth;   /* maxGuidedDepthIn{Training,Guiding}Phase */
        float uniform_fraction; /* uniformFractionIn{Training,Guiding}Phase */
        int max_train_depth;    /* Guidance::maxTrainDepth (3) */
    } orc_guided_settings;
    void *orc_net_create(int dim, int n_levels, int base_res, float per_level_scale, const float *aabb_min, const float *aabb_max);
    void orc_net_destroy(void *net);
    int orc_net_num_params(void *net, int *n_mlp, int *k0);
    void orc_net_set_params(void *net, const float *params);
    void orc_net_forward(void *net, const float *pts, int n, float *out, float *acts, uint32_t *cell_idx, float *cell_w);
    void orc_vmm_eval(int dim, const float *raw, const float *dir, const float *nrm, const uint8_t *onN, int n, uint32_t seed, float *pdf, float *grad,
                      float *sampled);
    void orc_kl_grad(int dim, const float *raw48, const float *dir, const float *nrm, const uint8_t *onN, const float *dirPdf, const float *li, int n,
                     float scale, float *dl, float *loss);
    float orc_eval_poly_large0(float y);
    float orc_log_bessel(float x, int order);
    float orc_vonmises_log_pdf(float kappa, float cos_theta);
    float orc_vonmises_pdf(float kappa, float cos_theta);
    float orc_vonmises_dlog_dkappa(float kappa, float cos_theta);
    float orc_vmf_pdf(float kappa, float cos_theta);
    int orc_solve_guided(const orc_problem_desc *p, const orc_settings *s, const orc_guided_settings *g, void *net, int pixel_begin, int pixel_end,
                         int sample0, int nsamples, float *solution, uint64_t *steps, float *samples_out, uint8_t *samples_on, int cap, int nthreads);

    /* evaluation point of a pixel (dim floats) */
    void orc_eval_point(const orc_problem_desc *p, const orc_settings *s, int pixel, float *out);

    /* ---- detmath / philox probes for tests ---- */
    void orc_detmath_eval(int fn, const float *x, const float *y, int n, float *out);
    void orc_philox(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1, uint32_t *out4);

#ifdef __cplusplus
}
#endif
#endif
