// Host-side integrator object behind `wost_integrator`.
#pragma once
#include "guided.cuh"
#include "problem.cuh"

#include <chrono>

namespace wost
{
    // Work queue of walk states: "SoA of 128-bit lanes". Replaces WorkQueue<T>/SOA<T>
    // (integrator/uniform/workqueue.h:32-69, util/soac.cpp:302-508): one float4 column per 16 bytes of state, so a warp
    // moves 512 contiguous bytes per instruction; slots are claimed with ONE atomic per warp (ballot + popc), not one
    // per item, and the queue of bounce d+1 is written compacted, so later bounces only touch live walks.
    //   2D: a = {x, y, bits(ai), bits(sample | onN << 31)}   b = {nx, ny, thp, 0}
    //   3D: a = {x, y, z, thp}   b = {nx, ny, nz, bits(ai)}   c = sample | onN << 31
    // ai = rank of the pixel among the shard's active pixels (pixelId = active[ai])
    struct QueueView
    {
        float4 *a;
        float4 *b;
        uint32_t *c;
    };
    // L2 persistence bookkeeping (integrator.cu): the persisting-L2 carve-out is a per-DEVICE limit shared by all integrators
    void release_l2_window(int device);
}

struct wost_integrator
{
    wost_problem *problem = nullptr; // not owned
    wost_settings s;
    int kind = 0, dim = 0, device = 0;
    int npix = 0;
    int n_active = 0;  // pixels of this shard that pass the mask
    int n_active_frame = 0; // the same count over ALL shards (identical on every rank: sizes the training batches)
    int spp_batch = 1; // S: samples of every pixel in flight per wavefront
    size_t queue_capacity = 0;
    wost::DevBuf<uint32_t> d_active;
    wost::DevBuf<uint2> d_first; // per active pixel: closest-point result of the evaluation point {leaf position, re-squared distance}
    wost::DevBuf<float4> qa[2], qb[2];
    wost::DevBuf<uint32_t> qc[2];
    wost::DevBuf<int> d_counts; // [0..max_depth]: items of bounce d; [max_depth+1 ..]: claim cursors
    wost::DevBuf<unsigned long long> d_steps;
    wost::DevBuf<float4> d_acc; // [slot][active rank] rgb accumulators (shard-sized: S * n_active)
    std::vector<int> slot_n;    // samples of every pixel that went to accumulator slot s in the last solve (wost_film_stderr)
    wost::DevBuf<int> d_slot_n;
    wost::DevBuf<float> d_se;   // [pixel][rgb] standard error of the pixel mean, filled on request
    wost::DevBuf<float> film[WOST_CH_COUNT];
    wost::DevBuf<int> d_overflow;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    uint64_t launches = 0;
    uint64_t step_launches = 0;  // launches of the fused walk-step kernel in the last solve
    float step_kernel_ms = 0.0f; // device time between the first and last walk-step launch of every spp batch, summed
    std::vector<cudaEvent_t> batch_ev; // 2 events per spp batch
    // guided integrator
    std::unique_ptr<wost::GuideNet> net;
    wost::GuidedBuffers gb;
    wost::DevBuf<float4> gq[2][3]; // guided out-shell queues: {x,thp} {n,R_B} {pixel,sample|onN,Tr,Tg}
    wost::DevBuf<float> gq3[2];    // Tb
    wost::DevBuf<unsigned long long> d_cursor64;
    wost_allreduce_fn allreduce = nullptr;
    void *allreduce_user = nullptr;
    wost_frame_fn frame_cb = nullptr; // progressive metrics (frames/, frames_time/)
    void *frame_user = nullptr;
    std::vector<float> frame_host;
    std::chrono::high_resolution_clock::time_point solve_t0;
    uint64_t train_samples = 0, train_steps = 0;
    float last_loss = 0.0f;
    std::vector<cudaEvent_t> prof_ev; // WOST_PROFILE=1
    float prof_ms[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    size_t l2_window_bytes = 0; // top-level LBVH records pinned in L2 through the stream's access-policy window (0: tree fits L2)

    // progressive metrics: is sample index `sid` one the reference dumps (uniform/integrator.cu:577-609)? bit 0 spp, bit 1 time
    int frame_kinds(int sid) const
    {
        if (!frame_cb)
            return 0;
        int k = 0;
        if (s.save_spp_metrics_duration > 0 && sid % s.save_spp_metrics_duration == 0 && sid < s.save_spp_metrics_until)
            k |= 1;
        if (s.save_time_metrics_duration > 0 && sid % s.save_time_metrics_duration == 0)
            k |= 2;
        return k;
    }
    // largest batch [s0, s0 + n) with n <= want that ends right after the next dumped sample index (or want if none)
    int frame_batch(int s0, int want) const
    {
        if (!frame_cb)
            return want;
        for (int i = 0; i < want; ++i)
            if (frame_kinds(s0 + i))
                return i + 1;
        return want;
    }

    ~wost_integrator()
    {
        for (cudaEvent_t e : batch_ev)
            cudaEventDestroy(e);
        for (cudaEvent_t e : prof_ev)
            cudaEventDestroy(e);
        if (ev0)
            cudaEventDestroy(ev0);
        if (ev1)
            cudaEventDestroy(ev1);
        if (stream)
            cudaStreamDestroy(stream);
        if (l2_window_bytes)
            wost::release_l2_window(device);
    }
};

namespace wost
{
    // progressive metrics: resolve solution / (sample_id + 1), download, hand to the frame callback (no-op when sample_id is not dumped)
    wost_status emit_frames(wost_integrator *it, int sample_id);
}
