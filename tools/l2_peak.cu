// L2 bandwidth microbenchmark: the denominator for the L2-resident traversals (C1-C3, c4-small), whose DRAM
// traffic is ~3 B per walk-step and for which the HBM roofline is only nominal.
//   random16 : every lane loads one float4 (16 B) from an independent random 16-byte slot of the buffer -- the access shape of the
//              per-lane node fetches of csrc/lbvh.cuh (each lane is in a different node). Bytes counted: the 16 B asked for
//              ("useful") and the 32-byte sectors the L2 actually delivers.
//   random64 : every lane loads 4 consecutive float4 (one 64-byte half record) at a random 64-byte-aligned offset -- the shape
//              of one child's extents + oriented box in a TNode.
//   warp512  : a warp loads 512 contiguous bytes at a random 512-byte-aligned offset (fully coalesced, upper bound).
// Buffer: 32 MiB by default (argv[1] in MiB), far below the 126 MB L2 and beyond the 148 x 256 KB of L1. Timed with CUDA events,
// best and median of 7 launches after 2 warm-ups. Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/l2_peak tools/l2_peak.cu
#include <cuda_runtime.h>
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x)                                                                                   \
    do                                                                                          \
    {                                                                                           \
        cudaError_t e_ = (x);                                                                   \
        if (e_ != cudaSuccess)                                                                  \
        {                                                                                       \
            fprintf(stderr, "%s:%d %s: %s\n", __FILE__, __LINE__, #x, cudaGetErrorString(e_)); \
            exit(1);                                                                            \
        }                                                                                       \
    } while (0)

__device__ __forceinline__ unsigned lcg(unsigned &s)
{
    s = s * 1664525u + 1013904223u;
    return s;
}
__device__ __forceinline__ float4 ld_cg(const float4 *p)
{
    float4 v;
    asm volatile("ld.global.cg.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p)); // L2 only: bypass L1
    return v;
}

// MODE 0: random 16 B per lane, 1: random 64 B per lane, 2: 512 B per warp
template <int MODE>
__global__ void __launch_bounds__(256) k_l2(const float4 *buf, unsigned slots16, int iters, float *sink)
{
    unsigned s = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 12345u;
    const unsigned lane = threadIdx.x & 31u;
    unsigned ws = (blockIdx.x * blockDim.x + threadIdx.x) / 32u * 2246822519u + 777u; // warp-uniform stream for MODE 2
    float acc = 0.0f;
    for (int i = 0; i < iters; ++i)
    {
        float4 v[4];
        if (MODE == 0)
        {
#pragma unroll
            for (int k = 0; k < 4; ++k)
                v[k] = ld_cg(buf + (lcg(s) >> 4) % slots16);
        }
        else if (MODE == 1)
        {
            const unsigned base = ((lcg(s) >> 4) % (slots16 / 4)) * 4;
#pragma unroll
            for (int k = 0; k < 4; ++k)
                v[k] = ld_cg(buf + base + k);
        }
        else
        {
#pragma unroll
            for (int k = 0; k < 4; ++k)
                v[k] = ld_cg(buf + ((lcg(ws) >> 4) % (slots16 / 32)) * 32 + lane);
        }
#pragma unroll
        for (int k = 0; k < 4; ++k)
            acc += v[k].x + v[k].w;
    }
    if (acc == 123.456f)
        *sink = acc;
}

template <int MODE>
static void run(const char *name, const float4 *buf, unsigned slots16, int sms, float *sink, double sector_factor, bool last)
{
    const int blocks = sms * 8, threads = 256, iters = 2048;
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    std::vector<float> ms;
    for (int r = 0; r < 9; ++r)
    {
        CK(cudaEventRecord(e0));
        k_l2<MODE><<<blocks, threads>>>(buf, slots16, iters, sink);
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float t;
        CK(cudaEventElapsedTime(&t, e0, e1));
        if (r >= 2)
            ms.push_back(t);
    }
    std::sort(ms.begin(), ms.end());
    const double bytes = (double)blocks * threads * iters * 4 * 16;
    printf("  \"%s\": {\"useful_gbs_best\": %.1f, \"useful_gbs_median\": %.1f, \"sector_gbs_best\": %.1f, \"ms_best\": %.4f, \"bytes_per_launch\": %.0f}%s\n", name,
           bytes / ms.front() * 1e-6, bytes / ms[ms.size() / 2] * 1e-6, bytes * sector_factor / ms.front() * 1e-6, ms.front(), bytes, last ? "" : ",");
}

int main(int argc, char **argv)
{
    const size_t mib = argc > 1 ? (size_t)atoi(argv[1]) : 32;
    const size_t bytes = mib << 20;
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, 0));
    float4 *buf;
    float *sink;
    CK(cudaMalloc(&buf, bytes));
    CK(cudaMalloc(&sink, 4));
    CK(cudaMemset(buf, 0, bytes));
    const unsigned slots16 = (unsigned)(bytes / 16);
    printf("{\n  \"device\": \"%s\", \"sms\": %d, \"l2_bytes\": %d, \"buffer_mib\": %zu, \"loads\": \"ld.global.cg.v4.f32 (L1 bypassed)\",\n", prop.name, prop.multiProcessorCount,
           prop.l2CacheSize, mib);
    run<0>("random16", buf, slots16, prop.multiProcessorCount, sink, 2.0, false); // a 16 B request moves one 32 B sector
    run<1>("random64", buf, slots16, prop.multiProcessorCount, sink, 1.0, false);
    run<2>("warp512", buf, slots16, prop.multiProcessorCount, sink, 1.0, true);
    printf("}\n");
    return 0;
}
