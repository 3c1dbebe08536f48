// common.cuh — shared host/device helpers for the cpprl_b200 kernels (sm_100a only).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <string>

#include "cpprl_b200.h"

namespace cpprl_b200
{
// ---- host side ---------------------------------------------------------------------------
void set_error(const std::string &msg);
int fail(int code, const std::string &msg);
int check_cuda(cudaError_t e, const char *what);
extern std::atomic<uint64_t> g_launches;

// Scoped kernel timer: records CUDA events on `stream` around the launches of one family when
// timing is enabled (cpprl_b200_kernel_time_enable).
struct KernelTimer
{
    KernelTimer(const char *family, cudaStream_t stream);
    ~KernelTimer();
    const char *family;
    cudaStream_t stream;
    int slot;
};

inline void count_launch(int n = 1) { g_launches.fetch_add(n, std::memory_order_relaxed); }

#define CPPRL_CUDA_TRY(expr)                                          \
    do                                                                \
    {                                                                 \
        int _rc = ::cpprl_b200::check_cuda((expr), #expr);            \
        if (_rc != 0)                                                 \
            return _rc;                                               \
    } while (0)

#define CPPRL_LAUNCH_CHECK(name)                                                       \
    do                                                                                 \
    {                                                                                  \
        ::cpprl_b200::count_launch();                                                  \
        int _rc = ::cpprl_b200::check_cuda(cudaGetLastError(), "launch of " name);     \
        if (_rc != 0)                                                                  \
            return _rc;                                                                \
    } while (0)

int require_device(); // 0 if current device is sm_10x, else ECUDA with message
int sm_count();

// Function attributes (dynamic shared-memory limit, carve-out) are per DEVICE: a launcher keeps one bit per device
// ordinal and configures its kernel the first time it runs on each.
inline bool configured_on_device(const std::atomic<uint64_t> &mask)
{
    int dev = 0;
    cudaGetDevice(&dev);
    return (mask.load(std::memory_order_acquire) >> (dev & 63)) & 1u;
}
inline void mark_configured_on_device(std::atomic<uint64_t> &mask)
{
    int dev = 0;
    cudaGetDevice(&dev);
    mask.fetch_or(1ull << (dev & 63), std::memory_order_release);
}

// ---- device side ---------------------------------------------------------------------------
#ifdef __CUDACC__
__device__ __forceinline__ float warp_sum(float v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
        v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__device__ __forceinline__ double warp_sum(double v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
        v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Streaming (read-once) 16-byte load / store that do not allocate in L1.
__device__ __forceinline__ float4 ld_stream4(const float *p)
{
    float4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
                 : "l"(p));
    return r;
}
__device__ __forceinline__ float ld_stream1(const float *p)
{
    float r;
    asm volatile("ld.global.nc.L1::no_allocate.f32 %0, [%1];" : "=f"(r) : "l"(p));
    return r;
}
__device__ __forceinline__ void st_stream4(float *p, float4 v)
{
    asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z),
                 "f"(v.w)
                 : "memory");
}

// Grid-wide barrier for cooperative (co-resident) launches.  `bar` points at two zero-initialised
// unsigned ints {arrive count, generation}; every CTA calls with the same `nblocks`.
__device__ __forceinline__ void grid_barrier(unsigned int *bar, unsigned int nblocks)
{
    __syncthreads();
    if (threadIdx.x == 0)
    {
        volatile unsigned int *gen_ptr = bar + 1;
        unsigned int gen = *gen_ptr;
        __threadfence();
        unsigned int ticket = atomicAdd(bar, 1u);
        if (ticket == nblocks - 1)
        {
            bar[0] = 0u;
            __threadfence();
            atomicAdd(bar + 1, 1u);
        }
        else
        {
            while (*gen_ptr == gen)
            {
            }
        }
        __threadfence();
    }
    __syncthreads();
}
#endif
} // namespace cpprl_b200
