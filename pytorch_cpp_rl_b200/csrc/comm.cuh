// comm.cuh — peer-memory exchange area shared by the kernels that reduce across GPUs (env-sharded updates).
//
// Every rank (one process per GPU) owns ONE exchange area in its own HBM, exported to the other ranks of the box
// with CUDA IPC and mapped over NVLink / NVSwitch.  A rank only ever WRITES its own area and READS its peers';
// readiness is signalled with monotonically increasing sequence numbers (st.release.sys / ld.acquire.sys), so no
// area is ever reset while a peer may still look at it.
#pragma once
#include <stdint.h>

namespace cpprl_b200
{
constexpr int kMaxRanks = 8;
constexpr int kMaxCtas = 148;
constexpr int kXMaxDim = 256; // observation columns the normaliser exchange can carry

struct ExchangeArea
{
    // gradient exchange: one flag per CTA index, two record buffers selected by the parity of the sequence number
    unsigned int step_flag[kMaxCtas + 12];
    // advantage moments {sum, sum of squares} of the local shard + flag (sequence = update count)
    double adv[4];
    unsigned int adv_flag, rms_flag, pad0, pad1;
    // observation-normaliser exchange: column sums, column sums of squares, row count
    double rms[2 * kXMaxDim + 2];
    // followed by float xbuf[2][x_stride]
};

struct CommView // passed by value to kernels
{
    int world, rank;
    int x_stride;                 // floats per parity buffer of the gradient exchange
    ExchangeArea *area[kMaxRanks]; // area[rank] is local memory, the others are peer mappings
};

#ifdef __CUDACC__
__device__ __forceinline__ float *exchange_xbuf(ExchangeArea *a) { return reinterpret_cast<float *>(a + 1); }

__device__ __forceinline__ void flag_publish(unsigned int *flag, unsigned int seq)
{
    __threadfence_system();
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(flag), "r"(seq) : "memory");
}

// Bounded wait (~ seconds): returns false when the peer never showed up, so a lost rank cannot hang the GPU.
__device__ __forceinline__ bool flag_wait(const unsigned int *flag, unsigned int seq)
{
    for (long long spins = 0; spins < (1ll << 24); ++spins)
    {
        unsigned int v;
        asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(flag) : "memory");
        if (static_cast<int>(v - seq) >= 0)
            return true;
        __nanosleep(100);
    }
    return false;
}

__device__ __forceinline__ float peer_load(const float *p)
{
    float v;
    asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ double peer_load(const double *p)
{
    double v;
    asm volatile("ld.relaxed.sys.global.f64 %0, [%1];" : "=d"(v) : "l"(p) : "memory");
    return v;
}
#endif
} // namespace cpprl_b200
