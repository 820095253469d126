// Host-side plumbing shared by the translation units of libwost.so: error state, CUDA checks, RAII device buffers.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string>
#include <vector>
#include "../../include/wost.h"

namespace wost
{
    void set_last_error(const std::string &msg);
    // Device allocations of libwost.so. Blocks are not returned to the driver when released but parked in a
    // per-device free list keyed by exact size (scene.cu): re-creating a scene / problem / integrator of the same shape --
    // what a host does between solves -- then costs neither cudaMalloc nor cudaFree (the latter took up to 60 ms for the
    // accumulator and queue buffers). wost_trim_allocation_cache() hands everything back.
    cudaError_t dev_alloc(void **ptr, size_t bytes);
    void dev_free(void *ptr, size_t bytes);

    struct CudaError
    {
        cudaError_t code;
        const char *file;
        int line;
    };

#define WOST_CUDA_TRY(expr)                                                                                       \
    do                                                                                                            \
    {                                                                                                             \
        cudaError_t _e = (expr);                                                                                  \
        if (_e != cudaSuccess)                                                                                    \
        {                                                                                                         \
            char _buf[512];                                                                                       \
            snprintf(_buf, sizeof(_buf), "%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
            ::wost::set_last_error(_buf);                                                                         \
            return WOST_ERR_CUDA;                                                                                 \
        }                                                                                                         \
    } while (0)

    template <typename T>
    struct DevBuf
    {
        T *ptr = nullptr;
        size_t count = 0;
        DevBuf() = default;
        DevBuf(const DevBuf &) = delete;
        DevBuf &operator=(const DevBuf &) = delete;
        ~DevBuf() { release(); }
        void release()
        {
            if (ptr)
                dev_free(ptr, count * sizeof(T));
            ptr = nullptr;
            count = 0;
        }
        cudaError_t alloc(size_t n)
        {
            release();
            count = n;
            if (n == 0)
                return cudaSuccess;
            return dev_alloc(reinterpret_cast<void **>(&ptr), n * sizeof(T));
        }
        cudaError_t upload(const T *host, size_t n, cudaStream_t s = 0)
        {
            cudaError_t e = alloc(n);
            if (e != cudaSuccess || n == 0)
                return e;
            return cudaMemcpyAsync(ptr, host, n * sizeof(T), cudaMemcpyHostToDevice, s);
        }
        cudaError_t download(T *host, size_t n, cudaStream_t s = 0) const
        {
            if (n == 0)
                return cudaSuccess;
            return cudaMemcpyAsync(host, ptr, n * sizeof(T), cudaMemcpyDeviceToHost, s);
        }
    };

    inline int div_up(long long a, long long b) { return (int)((a + b - 1) / b); }
    int sm_count();
} // namespace wost
