// Host-side state of the guided integrator: the guiding network (parameters, optimiser state, fp16 working copies),
// per-pixel walk records and the training-sample buffer. Owned by wost_integrator (integrator.cuh).
#pragma once
#include <memory>
#include "guiding.cuh"
#include "problem.cuh"

namespace wost
{
    constexpr int kMaxTrainDepth = 4; // MAX_TRAIN_DEPTH (parameters.h:6); Guidance::maxTrainDepth defaults to 3 (integrator.h:222)

    struct AdamConfig // tcnn Adam nested in Ema (data/ladybug/n.json:68-80; optimizers/adam.h, ema.h)
    {
        float learning_rate = 8e-3f, beta1 = 0.9f, beta2 = 0.99f, epsilon = 1e-15f, l2_reg = 1e-6f;
        float ema_decay = 0.95f;
        float relative_decay = 0.0f, absolute_decay = 0.0f;
    };

    // device-side control block of one training frame: lets the train / optimiser kernels decide by themselves
    // whether their batch exists, so that the host never reads a counter back (the reference syncs >= 6 times per batch)
    struct TrainCtl
    {
        int n_samples;       // written by the sample scatter kernel
        int opt_steps[2];    // [0] optimiser steps taken so far (EMA debiasing); [1] scratch of the optimiser kernel
        int batch_valid[8];  // per batch of the frame
        int batch_rows[8];
        float loss[8];       // sum of per-sample losses per batch
        int pad;
    };

    struct GuideNet
    {
        int dim = 0;
        int n_levels = 0, k0 = 0, base_res = 0;
        float per_level_scale = 1.0f;
        int n_mlp = 0, n_grid = 0, n_params = 0;
        AdamConfig opt;
        uint32_t seed = 1337u;
        GuideNetView view_train, view_infer; // same tables, different weight pointers
        DevBuf<float> master, adam_m, adam_v, ema, grad; // grad: n_params + 4 (last words: valid-rank count)
        DevBuf<uint32_t> steps;
        DevBuf<__half> w16_train, w16_infer; // [canonical MLP | grid]
        DevBuf<TrainCtl> ctl;
    };

    // SoA records of the first <= maxTrainDepth vertices of every walk (GuidedPixelStateBuffer, guided.h:12-69) and
    // the compacted training set (TrainBuffer, train.h:215-286)
    struct GuidedBuffers
    {
        DevBuf<float> rec;       // [(3*D + 5)][kMaxTrainDepth][npix]: pos, dir, nN, pdf, thp, base rgb
        DevBuf<uint8_t> rec_onN; // [kMaxTrainDepth][npix]
        DevBuf<int> rec_count;   // [npix]
        DevBuf<float> total;     // [3][npix] final contribution sum of the pixel's walk
        DevBuf<int> sample_off;  // [n_active + 1] exclusive scan of valid records per active pixel
        DevBuf<float> smp;       // [(3*D + 2)][cap]: pos(normalised by the kernel), dir, nN, pdf, Li
        DevBuf<uint8_t> smp_onN; // [cap]
        DevBuf<char> scan_tmp;
        size_t scan_tmp_bytes = 0;
        int cap = 0;
        DevBuf<float> infer_out; // validation mode (mlp_impl = 1): [queue_capacity][48] raw outputs written by the SIMT kernel
    };
} // namespace wost

struct wost_integrator;
namespace wost
{
    wost_status guided_alloc(wost_integrator *it);                       // guided.cu
    wost_status solve_guided(wost_integrator *it, uint64_t *walk_steps); // guided.cu
    void launch_resolve(wost_integrator *it, int nslots, float spp);     // integrator.cu
}
