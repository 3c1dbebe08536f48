// mlp_tc.cuh — tensor-core building blocks of the fused MLP update (large minibatches; BASELINE north_star (f):
// "tcgen05 tensor-core GEMMs ... for large-batch hidden-layer contractions").
//
// Orientation: D[feature][row].  The 128 TMEM lanes are the 2 x 64 hidden units of the two towers (lane F: tower F / 64,
// unit F % 64), the columns are the rows of the tile.  Every contraction of the towers is
//     D[F][n] (+)= sum_k A[F][k] * B[n][k]
// with A read from TENSOR MEMORY (lane F, column k: weights W1 / W2 / W2^T for the forward pass and the data
// gradient, the activation gradients dz2 / dz1 for the weight gradients) and B from shared memory in the K-major
// SWIZZLE_NONE layout (activations / observations, one copy per contraction they feed).  fp32-grade products come
// from the 3xTF32 split (hi * hi + hi * lo + lo * hi into one fp32 accumulator; csrc/cnn_gemm_tc.cuh has the error
// measurements), so every operand exists as a hi and a lo plane.
//
// Tower-specific B operands (h1 of the actor vs h1 of the critic) run as TWO M = 128 instructions into two column
// ranges: lanes 0..63 of the first and lanes 64..127 of the second are the results, the other halves are never read
// (M = 64 costs the tensor core the same cycles as M = 128).  Probed on B200 with tools/umma_probe*.cu: A from TMEM
// is exact over all 128 lanes, a B operand whose last 8-row group does not exist only pollutes its own columns.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace cpprl_b200
{
namespace mtc
{
// ---- tensor-memory column map (512 columns allocated, 1 CTA per SM) ----------------------------------------------
constexpr uint32_t kColW = 0;    // [0,128)   W2 hi | lo (forward)  ->  W2^T hi | lo (data gradient)  ->  dz1 hi | lo (dW1)
constexpr uint32_t kColX = 128;  // [128,192) W1 hi | lo (32 + 32), resident for the whole minibatch
constexpr uint32_t kColA = 192;  // [192,320) layer-1 / layer-2 accumulators  ->  dz2 hi | lo (dW2)
constexpr uint32_t kColG = 320;  // [320,448) dh1 accumulators (data gradient)  ->  dW2 accumulators of the tile
constexpr uint32_t kColD1 = 448; // [448,480) dW1 accumulator of the tile
constexpr uint32_t kTmemCols = 512;
constexpr int kMaxDp = 32;       // observation columns (padded to 8) this path takes

// ---- shared-memory operand geometry (bytes).  K-major SWIZZLE_NONE with the 8-row groups 128 bytes apart:
//      element (n, k) of an operand lives at 16 n + (k / 4) * LBO + (k % 4) * 4.
template <int R>
struct Geo
{
    static_assert(R % 8 == 0 && R >= 8 && R <= 64, "tile rows");
    static constexpr int CH = R / 2;                       // columns (tile rows) per thread in the epilogues
    static constexpr int kN = (R + 15) / 16 * 16;          // N of the instructions whose columns are the tile's rows (M = 128: N % 16 == 0)
    static constexpr int LD = R + 4;                       // row stride of the fp32 feature-major h2 tile (heads)
    static constexpr uint32_t kLboB = (R / 8) * 128 + 16;  // row-major operands [n = row][k = feature]: 16 B off a multiple of
                                                           // 128 -> the per-feature 4-byte stores of a warp are conflict-free
    static constexpr uint32_t kPlaneB = 16 * kLboB;        // one plane of a [R][64] operand (16 k chunks)
    static constexpr uint32_t kLboF = 8 * 128;             // feature-major operands [n = feature (64)][k = row]
    static constexpr uint32_t kPlaneF = (R / 4) * kLboF;
    static constexpr uint32_t kLboXF = 4 * 128 + 16;       // x feature-major [n = obs column (<= 32)][k = row]
    static constexpr uint32_t kPlaneXF = (R / 4) * kLboXF;
    __host__ __device__ static constexpr uint32_t plane_xb(int Dp) { return static_cast<uint32_t>(Dp / 4) * kLboB; }
};

__device__ __forceinline__ uint32_t smem_u32_tc(const void *p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

__device__ __forceinline__ uint64_t make_desc(uint32_t addr, uint32_t lbo_bytes)
{
    // UMMA shared-memory matrix descriptor, SWIZZLE_NONE, K-major: start address, leading byte offset (between the
    // 16-byte k chunks), stride byte offset 128 (between 8-row groups), all >> 4; descriptor version 1 in bits [46, 48)
    return static_cast<uint64_t>((addr >> 4) & 0x3fffu) | (static_cast<uint64_t>((lbo_bytes >> 4) & 0x3fffu) << 16) |
           (static_cast<uint64_t>(128u >> 4) << 32) | (1ull << 46);
}

// instruction descriptor: D = F32, A = B = TF32, both K-major, M = 128
__host__ __device__ constexpr uint32_t idesc_n(int N)
{
    return (1u << 4) | (2u << 7) | (2u << 10) | (static_cast<uint32_t>(N >> 3) << 17) | (static_cast<uint32_t>(128 >> 4) << 24);
}

// A from tensor memory, B through a shared-memory descriptor given as its two 32-bit words: the high word {stride byte
// offset 128, version 1} is the same for every operand, the low word {start address >> 4, leading byte offset >> 4} is a
// per-operand base plus a compile-time constant per k-step (shared-memory addresses stay below 2^18: the 14-bit
// start-address field never carries), so the single issuing thread spends one add per instruction on it.
__device__ __forceinline__ void mma_ts(uint32_t d_tmem, uint32_t a_tmem, uint32_t desc_lo, uint32_t idesc, uint32_t accumulate)
{
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        ".reg .b64 db;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "mov.b64 db, {%2, %5};\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], db, %3, {%6, %6, %6, %6}, p;\n\t"
        "}\n" ::"r"(d_tmem),
        "r"(a_tmem), "r"(desc_lo), "r"(idesc), "r"(accumulate), "r"((128u >> 4) | (1u << 14)), "r"(0u)
        : "memory");
}
__device__ __forceinline__ uint32_t desc_lo_of(uint32_t smem_addr, uint32_t lbo_bytes) { return ((smem_addr >> 4) & 0x3fffu) | ((lbo_bytes >> 4) << 16); }

// One contraction D = A * B^T as 3xTF32 over the k-steps [K0, K1) of 8: A hi / lo at TMEM columns a_hi / a_lo (advancing
// by 8 per k-step), B hi / lo planes at shared addresses b_hi / b_lo (advancing by two k chunks of LBO bytes per
// k-step).  Fully unrolled: the descriptor of every instruction is base + immediate.
// The tensor core adds into its fp32 accumulator with TRUNCATION: every instruction of a chain costs up to one ulp of
// the running sum, whatever the size of its addend.  The cross terms (hi * lo, lo * hi: 2^-11 of the result) therefore
// go FIRST, while the accumulator is still small and its ulps negligible, and only the hi * hi instructions work on
// the full-size sum: 8 ulp-sized truncations for a 64-long contraction instead of 24 (measured: 1.75e-7 of sum |a b|
// against 4.3e-7 for the interleaved order, tools/umma_probe2.cu).
template <int K0, int K1, uint32_t LBO>
__device__ __forceinline__ void issue_3xtf32(uint32_t d_tmem, uint32_t a_hi, uint32_t a_lo, uint32_t b_hi, uint32_t b_lo, uint32_t idesc)
{
    const uint32_t dh = desc_lo_of(b_hi, LBO), dl = desc_lo_of(b_lo, LBO);
#pragma unroll
    for (int ks = K0; ks < K1; ++ks)
    {
        mma_ts(d_tmem, a_hi + 8 * ks, dl + ((ks * 2 * LBO) >> 4), idesc, ks > K0 ? 1u : 0u);
        mma_ts(d_tmem, a_lo + 8 * ks, dh + ((ks * 2 * LBO) >> 4), idesc, 1u);
    }
#pragma unroll
    for (int ks = K0; ks < K1; ++ks)
        mma_ts(d_tmem, a_hi + 8 * ks, dh + ((ks * 2 * LBO) >> 4), idesc, 1u);
}

__device__ __forceinline__ void commit(void *mbar)
{
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32_tc(mbar)) : "memory");
}
__device__ __forceinline__ void fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// four consecutive columns of this thread's lane (warp-collective: 32 lanes x 4 columns).  The wait is part of the same
// asm statement: the destination registers must not be read before tcgen05.wait::ld, and nothing else would stop the
// compiler from scheduling their first use ahead of a separate wait statement.
__device__ __forceinline__ void ld4(uint32_t addr, float (&v)[4])
{
    uint32_t r0, r1, r2, r3;
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];\n\t"
                 "tcgen05.wait::ld.sync.aligned;"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
                 : "r"(addr)
                 : "memory");
    v[0] = __uint_as_float(r0), v[1] = __uint_as_float(r1), v[2] = __uint_as_float(r2), v[3] = __uint_as_float(r3);
}
// N consecutive columns of this thread's lane WITHOUT the wait: the caller issues all the loads of an epilogue, then one
// wait_ld(), then reads the registers (the pattern of CUTLASS's fence_view_async_tmem_load)
__device__ __forceinline__ void ld4_nowait(uint32_t addr, float *v)
{
    uint32_t r[4];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
#pragma unroll
    for (int i = 0; i < 4; ++i)
        v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ void ld8_nowait(uint32_t addr, float *v)
{
    uint32_t r[8];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "r"(addr));
#pragma unroll
    for (int i = 0; i < 8; ++i)
        v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ void ld16_nowait(uint32_t addr, float *v)
{
    uint32_t r[16];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
                   "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                 : "r"(addr));
#pragma unroll
    for (int i = 0; i < 16; ++i)
        v[i] = __uint_as_float(r[i]);
}
// CH (a multiple of 4, at most 32: 28 and 32 in the kernel) columns starting at addr, one wait
template <int CH>
__device__ __forceinline__ void ld_cols(uint32_t addr, float (&v)[CH])
{
    static_assert(CH % 4 == 0 && CH <= 32, "columns per thread");
    int c = 0;
#pragma unroll
    for (; c + 16 <= CH; c += 16)
        ld16_nowait(addr + c, v + c);
    if (c + 8 <= CH)
    {
        ld8_nowait(addr + c, v + c);
        c += 8;
    }
    if (c + 4 <= CH)
        ld4_nowait(addr + c, v + c);
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

__device__ __forceinline__ void st4(uint32_t addr, float a, float b, float c, float d)
{
    asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(__float_as_uint(a)), "r"(__float_as_uint(b)),
                 "r"(__float_as_uint(c)), "r"(__float_as_uint(d))
                 : "memory");
}

// sixteen consecutive columns of this thread's lane
__device__ __forceinline__ void st16(uint32_t addr, const float (&v)[16])
{
    asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(addr),
                 "r"(__float_as_uint(v[0])), "r"(__float_as_uint(v[1])), "r"(__float_as_uint(v[2])), "r"(__float_as_uint(v[3])),
                 "r"(__float_as_uint(v[4])), "r"(__float_as_uint(v[5])), "r"(__float_as_uint(v[6])), "r"(__float_as_uint(v[7])),
                 "r"(__float_as_uint(v[8])), "r"(__float_as_uint(v[9])), "r"(__float_as_uint(v[10])), "r"(__float_as_uint(v[11])),
                 "r"(__float_as_uint(v[12])), "r"(__float_as_uint(v[13])), "r"(__float_as_uint(v[14])), "r"(__float_as_uint(v[15]))
                 : "memory");
}
// sixteen floats of global memory (four 16-byte loads; the caller keeps them in flight) -> hi plane at `col`, lo plane
// at `col + lo_off` of this thread's tensor-memory lane
__device__ __forceinline__ void store_split16(uint32_t col, uint32_t lo_off, const float4 (&w)[4])
{
    float hi[16], lo[16];
#pragma unroll
    for (int i = 0; i < 4; ++i)
    {
        const float x[4] = {w[i].x, w[i].y, w[i].z, w[i].w};
#pragma unroll
        for (int j = 0; j < 4; ++j)
        {
            hi[4 * i + j] = __uint_as_float((__float_as_uint(x[j]) + 0x1000u) & 0xffffe000u);
            lo[4 * i + j] = x[j] - hi[4 * i + j];
        }
    }
    st16(col, hi);
    st16(col + lo_off, lo);
}

// tanh for the tensor-core epilogues, 1 - 2 / (exp(2x) + 1) on the special-function unit (ex2.approx + rcp.approx: six
// instructions against ~25 of tanhf; the epilogues are instruction-issue bound: 28 activations per thread per layer).
// Absolute error below 2e-7 -- the rounding level of an fp32 activation near 1 -- as in the recurrent kernel's gates
// (gru_update.cu); CPPRL_B200_MLP_TC=0 keeps the correctly rounded tanhf of the CUDA-core variant for A/B runs.
__device__ __forceinline__ float tanh_sfu(float x)
{
    const float e = __expf(2.f * fminf(fmaxf(x, -15.f), 15.f)); // tanh(+-15) = +-1 in fp32
    return 1.f - __fdividef(2.f, e + 1.f);
}

// x = hi + lo with hi on the TF32 grid (round to nearest, ties away) and lo = x - hi exact in fp32 (cnn_gemm_tc.cuh)
__device__ __forceinline__ float tf32_hi(float x) { return __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xffffe000u); }
} // namespace mtc
} // namespace cpprl_b200
