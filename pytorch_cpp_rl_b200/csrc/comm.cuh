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
    unsigned int adv_flag, rms_flag;
    int rms_error; // a peer never published its normaliser sums (bounded wait expired): the merge was skipped
    unsigned int pad1;
    // observation-normaliser exchange: column sums, column sums of squares, row count; two buffers selected by the
    // parity of the sequence number, so back-to-back updates never overwrite sums a slower peer is still reading
    // (a rank enters sequence n+2 only after every peer published n+1, i.e. finished reading n)
    double rms[2][2 * kXMaxDim + 2];
    // followed by the gradient exchange: LLWord xbuf[2 (parity)][world (sender)][x_stride]
};

// Low-latency exchange word: value and sequence number travel in ONE 8-byte store, so a reader that sees the
// expected sequence number also sees the value — no fence, no separate flag (the NCCL "LL" idea).
struct LLWord
{
    float value;
    unsigned int seq;
};

struct CommView // passed by value to kernels
{
    int world, rank;
    int x_stride;                 // floats per parity buffer of the gradient exchange
    ExchangeArea *area[kMaxRanks]; // area[rank] is local memory, the others are peer mappings
};

#ifdef __CUDACC__
__device__ __forceinline__ LLWord *exchange_xbuf(ExchangeArea *a) { return reinterpret_cast<LLWord *>(a + 1); }

// PUSH a word into (possibly peer) memory / poll a word of LOCAL memory a peer pushes into.
__device__ __forceinline__ void ll_store(LLWord *dst, float value, unsigned int seq)
{
    asm volatile("st.relaxed.sys.global.v2.b32 [%0], {%1, %2};" ::"l"(dst), "r"(__float_as_uint(value)), "r"(seq) : "memory");
}
__device__ __forceinline__ bool ll_load(const LLWord *src, unsigned int seq, float &value)
{
    for (long long spins = 0; spins < (1ll << 24); ++spins)
    {
        unsigned int v, s;
        asm volatile("ld.relaxed.sys.global.v2.b32 {%0, %1}, [%2];" : "=r"(v), "=r"(s) : "l"(src) : "memory");
        if (s == seq)
        {
            value = __uint_as_float(v);
            return true;
        }
        if (spins > 8192)
            __nanosleep(200);
    }
    value = 0.f;
    return false;
}

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
        if (spins > 4096)
            __nanosleep(200);
    }
    return false;
}

__device__ __forceinline__ float peer_load(const float *p)
{
    float v;
    asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ float4 peer_load4(const float *p)
{
    float4 v;
    asm volatile("ld.relaxed.sys.global.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
    return v;
}
// The gradient exchange seen as plain floats (bulk records published under the per-CTA step flags: the CnnBase
// all-reduce): buffer `parity` of area `a` holds 2 * x_stride floats per sender slot; a rank uses slot 0 of its own area.
__device__ __forceinline__ float *exchange_fbuf(ExchangeArea *a, int parity, int world, int x_stride)
{
    return reinterpret_cast<float *>(exchange_xbuf(a)) + static_cast<size_t>(parity) * world * 2 * x_stride;
}
__device__ __forceinline__ double peer_load(const double *p)
{
    double v;
    asm volatile("ld.relaxed.sys.global.f64 %0, [%1];" : "=d"(v) : "l"(p) : "memory");
    return v;
}
#endif
} // namespace cpprl_b200
