// cnn_gemm_tc.cuh — the CnnBase contractions on the 5th-generation tensor cores (tcgen05, accumulators in TMEM).
//
//   C[m][n] (+)= sum_k A(m, k) * B(k, n)       same loader / epilogue contract as cnn_gemm.cuh
//
// Parity needs fp32-grade products (post-update parameters to 1e-4 after hundreds of Adam steps), so every fp32
// operand element x is split into hi = tf32(x) (round to nearest) and lo = x - hi, and each k-step issues THREE
// kind::tf32 MMAs into the same fp32 TMEM accumulator: hi*hi + hi*lo + lo*hi (the dropped lo*lo term is 2^-22 of
// the product) — "3xTF32".
//
// CTA = 256 threads, tile 128 (UMMA M) x BN x 32.  All threads gather the operand tile from global memory (implicit
// im2col through the same loaders as the fp32 kernel), split it and store it into the canonical NO-SWIZZLE UMMA
// K-major shared-memory layout with 16-byte stores (operands whose vectors run along m / n are transposed in
// registers, 4x4 blocks; MN-major no-swizzle descriptors read back zeros for kind::tf32 on this part), then one
// elected thread issues the MMAs (descriptors advance by k-step) and commits them to an mbarrier; two
// shared-memory stages, the global loads of tile t+1 are in flight while tile t's MMAs run.  Epilogue: tcgen05.ld
// (32 lanes x 8 columns per warp instruction) -> registers -> epi(m, n, float4).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace cpprl_b200
{
namespace tc
{
constexpr int kThreads = 256;
constexpr int kBM = 128;
constexpr int kBK = 32;

__device__ __forceinline__ uint32_t smem_addr(const void *p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

__device__ __forceinline__ uint64_t make_desc(uint32_t addr, uint32_t lbo_bytes, uint32_t sbo_bytes)
{
    // UMMA shared-memory matrix descriptor, SWIZZLE_NONE: start address, leading / stride byte offsets (all >> 4),
    // descriptor version 1 (Blackwell) in bits [46, 48)
    return static_cast<uint64_t>((addr >> 4) & 0x3fffu) | (static_cast<uint64_t>((lbo_bytes >> 4) & 0x3fffu) << 16) |
           (static_cast<uint64_t>((sbo_bytes >> 4) & 0x3fffu) << 32) | (1ull << 46);
}

__device__ __forceinline__ void mma_tf32(uint32_t tmem_c, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate)
{
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n\t"
        "}\n" ::"r"(tmem_c),
        "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate), "r"(0u)
        : "memory");
}

// the same MMA with the descriptors given as 32-bit halves: the high word {stride byte offset, version} is common to
// all operands, the low word {start address >> 4, leading byte offset} is a loop-invariant base plus a constant
__device__ __forceinline__ void mma_tf32_halves(uint32_t tmem_c, uint32_t a_lo, uint32_t b_lo, uint32_t desc_hi, uint32_t idesc,
                                                uint32_t accumulate)
{
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        ".reg .b64 da, db;\n\t"
        "setp.ne.b32 p, %5, 0;\n\t"
        "mov.b64 da, {%1, %3};\n\t"
        "mov.b64 db, {%2, %3};\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], da, db, %4, {%6, %6, %6, %6}, p;\n\t"
        "}\n" ::"r"(tmem_c),
        "r"(a_lo), "r"(b_lo), "r"(desc_hi), "r"(idesc), "r"(accumulate), "r"(0u)
        : "memory");
}

__device__ __forceinline__ void mbar_init(void *bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_wait(void *bar, uint32_t parity)
{
    uint32_t done = 0;
    const uint32_t addr = smem_addr(bar);
    while (!done)
        asm volatile(
            "{\n\t"
            ".reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t"
            "}\n"
            : "=r"(done)
            : "r"(addr), "r"(parity)
            : "memory");
}

// x = hi + lo with hi on the TF32 grid (round to nearest, ties away: what cvt.rna.tf32.f32 does, written as integer
// arithmetic because the cvt expands to a four-instruction sequence with an Inf/NaN guard this path has no use for;
// Inf stays Inf, values within half a TF32 ulp of FLT_MAX round to Inf) and lo = x - hi exact in fp32.
__device__ __forceinline__ float tf32_hi(float x) { return __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xffffe000u); }
__device__ __forceinline__ void split_tf32(const float4 &x, float4 &hi, float4 &lo)
{
    hi = make_float4(tf32_hi(x.x), tf32_hi(x.y), tf32_hi(x.z), tf32_hi(x.w));
    lo = make_float4(x.x - hi.x, x.y - hi.y, x.z - hi.z, x.w - hi.w);
}

// Shared-memory operand layout (both operands K-major, SWIZZLE_NONE): [k chunk of 4 floats][8-row group][row % 8][4 floats]
// with the 8-row groups 144 bytes apart (128 + 16 padding: the descriptor's stride byte offset is free, and the
// padding makes the transposing 16-byte stores below conflict-free) and the k chunks (rows / 8) * 144 bytes apart.
constexpr int kGroupBytes = 144;
// The A operand's k chunks are a further 16 bytes apart (leading byte offset 16 * 144 + 16): eight lanes that store the
// eight k chunks of ONE row (the lanes-along-k gather below) then fall into eight different 16-byte bank groups.
__host__ __device__ constexpr int operand_bytes(int rows) { return (kBK / 4) * (rows / 8) * kGroupBytes; }
__host__ __device__ constexpr int a_lbo_bytes() { return (kBM / 8) * kGroupBytes + 16; }
__host__ __device__ constexpr int a_operand_bytes() { return (kBK / 4) * a_lbo_bytes(); }
__host__ __device__ constexpr size_t smem_bytes(int BN) { return static_cast<size_t>(2) * 2 * (a_operand_bytes() + operand_bytes(BN)) + 64; }

// A loader type may declare `static constexpr bool kTf32Exact = true`: every element it yields is ON the TF32 grid (raw
// 0..255 pixels are), so A_lo == 0 exactly, the A_lo x B_hi product contributes exactly nothing and is compiled out
// together with the lo plane's split and shared-memory stores: one MMA per k-step instead of two, same bits out.
// The host picks that loader type only after cnn_obs_scan_kernel has checked every observation of the update.
template <class L, class = void>
struct ExactOf
{
    static constexpr bool value = false;
};
template <class L>
struct ExactOf<L, decltype(void(L::kTf32Exact))>
{
    static constexpr bool value = L::kTf32Exact;
};

// An epilogue type may declare `static constexpr bool kColSum = true`: it then has store(m, n, v) returning what it wrote,
// a pointer `colsum` (may be null) and `colmod`; the kernel adds the column sums of what its CTA wrote to
// colsum[n % colmod] (one atomic per column and CTA).
template <class E, class = void>
struct ColSumOf
{
    static constexpr bool value = false;
};
template <class E>
struct ColSumOf<E, decltype(void(E::kColSum))>
{
    static constexpr bool value = E::kColSum;
};

// A_T: A(m, k) = X[k][m] (the loader's vectors run along m: transposed on the way into shared memory); else
// A(m, k) = X[m][k] (vectors along k).  B(k, n) = Y[k][n] always (vectors along n, transposed likewise).
template <int BN, bool A_T, class ALoad, class BLoad, class Epi>
__global__ void __launch_bounds__(kThreads, BN == 128 ? 1 : 2) gemm_tc_kernel(const ALoad al, const BLoad bl, const Epi epi, int M, int N, int K,
                                                           int k_slice, int z_is_k_split, const int *stop_flag)
{
    static_assert(BN == 32 || BN == 64 || BN == 128, "tile columns");
    extern __shared__ __align__(128) unsigned char tc_smem[];
    if (stop_flag && *stop_flag)
        return;
    constexpr int A_BYTES = a_operand_bytes(), B_BYTES = operand_bytes(BN), STAGE_BYTES = 2 * (A_BYTES + B_BYTES);
    // B holds its hi and lo planes as ONE operand of 2 BN rows (hi rows [0, BN), lo rows [BN, 2 BN)): A_hi x [B_hi; B_lo]
    // is a single N = 2 BN instruction whose two column halves are hi*hi and hi*lo, so a k-step is TWO instructions
    // (the second, A_lo x B_hi with N = BN, adds into the first half) and A — the 4 KB operand whose shared-memory
    // reads bound these small-N MMAs — is read twice instead of three times.
    constexpr uint32_t A_LBO = a_lbo_bytes(), B_LBO = (2 * BN / 8) * kGroupBytes;
    // K-major sources whose rows are contiguous in memory (activations, plain matrices) are gathered with the LANES
    // ALONG K: eight lanes read the 128 contiguous bytes one row contributes to a k tile, a warp request touches 4
    // lines instead of 32 (the L1 wavefront count had been the limiter of those kernels).  The thread then owns the
    // same k chunk of four rows (four row contexts).  conv1's NCHW gather keeps lanes along the output pixels.
    constexpr bool ALONG_K = !A_T && ALoad::kLanesAlongK;
    constexpr bool A_EXACT = ExactOf<ALoad>::value;
    uint64_t *mbar = reinterpret_cast<uint64_t *>(tc_smem + 2 * STAGE_BYTES);
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(mbar + 2);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int m0 = blockIdx.x * kBM, n0 = blockIdx.y * BN;
    const int k_begin = z_is_k_split ? blockIdx.z * k_slice : 0, k_end = min(K, k_begin + k_slice);

    // ---- TMEM allocation (one warp), barriers ----------------------------------------------------------------------
    constexpr uint32_t TMEM_COLS = 4 * BN; // two accumulators of 2 BN columns: see the chunked accumulation below
    if (warp == 0)
    {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_addr(tmem_slot)), "r"(TMEM_COLS)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    if (tid == 0)
    {
        mbar_init(&mbar[0], 1);
        mbar_init(&mbar[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_acc = *tmem_slot;

    // instruction descriptor: D = F32 (bits 4-5), A = B = TF32 (bits 7-9, 10-12), both K-major, N >> 3, M >> 4
    constexpr uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | (static_cast<uint32_t>(BN >> 3) << 17) |
                               (static_cast<uint32_t>(kBM >> 4) << 24);
    constexpr uint32_t idesc2 = (1u << 4) | (2u << 7) | (2u << 10) | (static_cast<uint32_t>((2 * BN) >> 3) << 17) |
                                (static_cast<uint32_t>(kBM >> 4) << 24);

    // byte offset of the 16-byte slot (k chunk k4, row r) inside an operand
    auto slot = [](int k4, int r, uint32_t lbo) { return k4 * lbo + (r >> 3) * kGroupBytes + (r & 7) * 16; };
    constexpr int B_BLOCKS = (BN / 4) * (kBK / 4); // 4x4 blocks of the B tile
    // two tiles of operands in flight per thread (the loads of tiles t+1 and t+2 overlap tile t's MMAs)
    float4 ra2[2][4], rb2[2][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
        ra2[0][i] = ra2[1][i] = rb2[0][i] = rb2[1][i] = make_float4(0.f, 0.f, 0.f, 0.f);
    // K-major A: every vector of this thread belongs to the SAME row m0 + (tid & 127): its context is formed once
    auto a_row = al.row(A_T ? 0 : m0 + (tid & (kBM - 1)));
    al.clip(a_row, k_end);
    decltype(a_row) a_rows[4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
    {
        a_rows[i] = al.row(ALONG_K ? m0 + (tid >> 3) + 32 * i : 0);
        al.clip(a_rows[i], k_end);
    }
    // every load is unconditional (out-of-range elements come from the loaders' zero page): no branches, no zeroing
    const int b_row_lim = bl.row_limit(n0 + 4 * (tid % (BN / 4)), k_end);
    // element offsets = per-thread column part (formed once) + per-tile part (uniform over the CTA)
    int a_col[4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
        a_col[i] = al.col_part(ALONG_K ? 4 * (tid & 7) : 4 * ((tid >> 7) + 2 * i));
    auto fetch = [&](int k0, float4 (&ra)[4], float4 (&rb)[4])
    {
        const int a_tile = al.tile_part(k0);
        (void)a_tile;
#pragma unroll
        for (int i = 0; i < 4; ++i)
        {
            if (A_T)
            {
                const int m4 = tid & 31, k4 = tid >> 5; // one 4 (k) x 4 (m) block per thread
                const int r = k0 + 4 * k4 + i;
                auto ctx = al.row(r);
                ctx.lim = r < k_end ? ctx.lim : 0;
                ra[i] = al.load(ctx, m0 + 4 * m4);
            }
            else if (ALONG_K)
                ra[i] = al.load_parts(a_rows[i], a_col[0] + a_tile, k0 + 4 * (tid & 7));
            else
                ra[i] = al.load_parts(a_row, a_col[i] + a_tile, k0 + 4 * ((tid >> 7) + 2 * i));
            if (tid < B_BLOCKS)
            {
                const int n4 = tid % (BN / 4), k4 = tid / (BN / 4);
                rb[i] = bl.at(k0 + 4 * k4 + i, n0 + 4 * n4, b_row_lim);
            }
        }
    };
    auto store_split = [](unsigned char *hi_base, unsigned char *lo_base, int off, const float4 &x)
    {
        float4 hi, lo;
        split_tf32(x, hi, lo);
        *reinterpret_cast<float4 *>(hi_base + off) = hi;
        *reinterpret_cast<float4 *>(lo_base + off) = lo;
    };
    auto stash = [&](int s, const float4 (&ra)[4], const float4 (&rb)[4])
    {
        unsigned char *a_hi = tc_smem + s * STAGE_BYTES, *a_lo = a_hi + A_BYTES, *b_hi = a_lo + A_BYTES, *b_lo = b_hi + (BN / 8) * kGroupBytes;
        // an exact A has no lo plane: the values go to shared memory as they are
        auto store_a = [&](int off, const float4 &x)
        {
            if (A_EXACT)
                *reinterpret_cast<float4 *>(a_hi + off) = x;
            else
                store_split(a_hi, a_lo, off, x);
        };
        if (A_T)
        {
            const int m4 = tid & 31, k4 = tid >> 5;
            store_a(slot(k4, 4 * m4 + 0, A_LBO), make_float4(ra[0].x, ra[1].x, ra[2].x, ra[3].x));
            store_a(slot(k4, 4 * m4 + 1, A_LBO), make_float4(ra[0].y, ra[1].y, ra[2].y, ra[3].y));
            store_a(slot(k4, 4 * m4 + 2, A_LBO), make_float4(ra[0].z, ra[1].z, ra[2].z, ra[3].z));
            store_a(slot(k4, 4 * m4 + 3, A_LBO), make_float4(ra[0].w, ra[1].w, ra[2].w, ra[3].w));
        }
        else
        {
#pragma unroll
            for (int i = 0; i < 4; ++i)
            {
                const int idx = tid + i * kThreads;
                const int mm = ALONG_K ? idx >> 3 : idx & (kBM - 1), k4 = ALONG_K ? idx & 7 : idx >> 7;
                store_a(slot(k4, mm, A_LBO), ra[i]);
            }
        }
        if (tid < B_BLOCKS)
        {
            const int n4 = tid % (BN / 4), k4 = tid / (BN / 4);
            store_split(b_hi, b_lo, slot(k4, 4 * n4 + 0, B_LBO), make_float4(rb[0].x, rb[1].x, rb[2].x, rb[3].x));
            store_split(b_hi, b_lo, slot(k4, 4 * n4 + 1, B_LBO), make_float4(rb[0].y, rb[1].y, rb[2].y, rb[3].y));
            store_split(b_hi, b_lo, slot(k4, 4 * n4 + 2, B_LBO), make_float4(rb[0].z, rb[1].z, rb[2].z, rb[3].z));
            store_split(b_hi, b_lo, slot(k4, 4 * n4 + 3, B_LBO), make_float4(rb[0].w, rb[1].w, rb[2].w, rb[3].w));
        }
    };
    // descriptor halves (make_desc): smem addresses stay below 2^18, so base + constant never carries out of the
    // 14-bit start-address field; s is a literal at both call sites, every offset below folds into one add
    const uint32_t smem0 = smem_addr(tc_smem) >> 4;
    const uint32_t a_desc_lo = smem0 | ((A_LBO >> 4) << 16), b_desc_lo = smem0 | ((B_LBO >> 4) << 16);
    constexpr uint32_t desc_hi = (kGroupBytes >> 4) | (1u << 14);
    auto issue = [&](int s, uint32_t acc, bool first_tile)
    {
#pragma unroll
        for (int j = 0; j < kBK / 8; ++j) // one MMA k-step = 8 tf32 = two 16-byte k chunks
        {
            const uint32_t a_off = (s * STAGE_BYTES + j * 2 * A_LBO) >> 4, b_off = (s * STAGE_BYTES + 2 * A_BYTES + j * 2 * B_LBO) >> 4;
            mma_tf32_halves(acc, a_desc_lo + a_off, b_desc_lo + b_off, desc_hi, idesc2, (first_tile && j == 0) ? 0u : 1u); // [hi*hi | hi*lo]
            if (!A_EXACT)
                mma_tf32_halves(acc, a_desc_lo + a_off + (A_BYTES >> 4), b_desc_lo + b_off, desc_hi, idesc, 1u); // lo*hi into the first half
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_addr(&mbar[s]))
                     : "memory");
    };

    // ---- main loop ------------------------------------------------------------------------------------------------------
    // The tensor core adds into its fp32 accumulator with truncation, an error that grows linearly with the length of
    // the chain (measured 1e-8 * K relative).  The reduction is therefore cut into chunks of two k tiles (64 products)
    // that alternate between two TMEM accumulators; a finished chunk is read back (tcgen05.ld) and added to fp32
    // registers with round-to-nearest while the next chunk's MMAs run.
    const int n_tiles = k_begin < k_end ? (k_end - k_begin + kBK - 1) / kBK : 0;
    const int q = warp & 3, half = warp >> 2;
    float racc[BN / 2];
#pragma unroll
    for (int c = 0; c < BN / 2; ++c)
        racc[c] = 0.f;
    auto drain_chunk = [&](int chunk)
    {
        const uint32_t base = tmem_acc + (chunk & 1) * 2 * BN + (static_cast<uint32_t>(32 * q) << 16) + half * (BN / 2);
#pragma unroll
        for (int c = 0; c < BN / 2; c += 8)
        {
            uint32_t r[8], r2[8];
            asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                         : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                         : "r"(base + c));
            asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                         : "=r"(r2[0]), "=r"(r2[1]), "=r"(r2[2]), "=r"(r2[3]), "=r"(r2[4]), "=r"(r2[5]), "=r"(r2[6]), "=r"(r2[7])
                         : "r"(base + BN + c));
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
            for (int i = 0; i < 8; ++i)
                racc[c + i] += __uint_as_float(r[i]) + __uint_as_float(r2[i]); // (hi*hi + lo*hi) + hi*lo
        }
    };
    uint32_t phase[2] = {0u, 0u};
    if (n_tiles > 0)
        fetch(k_begin, ra2[0], rb2[0]);
    if (n_tiles > 1)
        fetch(k_begin + kBK, ra2[1], rb2[1]);
    auto tile_step = [&](int t, const int s, float4 (&ra)[4], float4 (&rb)[4]) // s = t & 1, a literal at the call sites
    {
        if (t >= 2)
        {
            mbar_wait(&mbar[s], phase[s]); // the MMAs of tile t-2 (which read this stage) are done
            phase[s] ^= 1u;
            if (s == 1)
            {
                // tile t-2 closed chunk (t-3)/2
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                drain_chunk((t - 3) >> 1);
            }
        }
        stash(s, ra, rb);
        if (t + 2 < n_tiles)
            fetch(k_begin + (t + 2) * kBK, ra, rb); // two tiles of global loads stay in flight
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); // generic-proxy stores -> visible to the tensor core
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncthreads();
        if (tid == 0)
        {
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            issue(s, tmem_acc + ((t >> 1) & 1) * 2 * BN, s == 0);
        }
    };
    for (int t = 0; t < n_tiles; t += 2)
    {
        tile_step(t, 0, ra2[0], rb2[0]);
        if (t + 1 < n_tiles)
            tile_step(t + 1, 1, ra2[1], rb2[1]);
    }
    // drain: chunks whose closing tile has not been waited for yet
    if (n_tiles >= 1)
    {
        const int last = n_tiles - 1;
        // tile last-1 (if any) has not been waited for either
        if (n_tiles >= 2)
        {
            mbar_wait(&mbar[(last - 1) & 1], phase[(last - 1) & 1]);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            if (((last - 1) & 1) == 1)
                drain_chunk((last - 1) >> 1); // it closed its chunk
        }
        mbar_wait(&mbar[last & 1], phase[last & 1]);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        drain_chunk(last >> 1); // the last chunk (one or two tiles)
    }

    // ---- epilogue: registers -> shared memory (transpose) -> epi ------------------------------------------------------------
    // TMEM hands every thread ONE ROW of the tile; stored from there a warp request would touch 32 lines (16 bytes of
    // each).  The tile goes through the (now idle: every MMA has completed) stage memory instead and is read back with
    // the lanes along n, so that the epilogue's global loads / stores / atomics cover whole lines.
    if (n_tiles > 0)
    {
        float *tile = reinterpret_cast<float *>(tc_smem);
        constexpr int LD = BN + 4; // row stride 16 bytes off a multiple of 128: conflict-free 16-byte stores per row
        const int row = 32 * q + lane;
#pragma unroll
        for (int c = 0; c < BN / 2; c += 4)
            *reinterpret_cast<float4 *>(tile + row * LD + half * (BN / 2) + c) = make_float4(racc[c], racc[c + 1], racc[c + 2], racc[c + 3]);
        __syncthreads();
        float4 cs = make_float4(0.f, 0.f, 0.f, 0.f);
        (void)cs;
#pragma unroll
        for (int j = 0; j < BN / 8; ++j)
        {
            const int v = tid + j * kThreads, r = v / (BN / 4), c4 = v % (BN / 4);
            const int m = m0 + r, n = n0 + 4 * c4;
            if (m < M && n < N)
            {
                if constexpr (ColSumOf<Epi>::value)
                {
                    const float4 o = epi.store(m, n, *reinterpret_cast<const float4 *>(tile + r * LD + 4 * c4));
                    cs.x += o.x, cs.y += o.y, cs.z += o.z, cs.w += o.w;
                }
                else
                    epi(m, n, *reinterpret_cast<const float4 *>(tile + r * LD + 4 * c4));
            }
        }
        if constexpr (ColSumOf<Epi>::value)
        {
            if (epi.colsum) // (uniform over the grid)
            {
                // kThreads is a multiple of BN / 4, so a thread met the same four columns in every iteration above: the
                // threads of a column group combine through the (idle again) tile memory, one atomic per column
                constexpr int G = BN / 4;
                __syncthreads();
                float4 *red = reinterpret_cast<float4 *>(tc_smem);
                red[tid] = cs;
                __syncthreads();
                if (tid < G && n0 + 4 * tid < N)
                {
                    float4 sum = red[tid];
                    for (int t = tid + G; t < kThreads; t += G)
                    {
                        const float4 o = red[t];
                        sum.x += o.x, sum.y += o.y, sum.z += o.z, sum.w += o.w;
                    }
                    float *dst = epi.colsum + (n0 + 4 * tid) % epi.colmod;
                    atomicAdd(dst, sum.x), atomicAdd(dst + 1, sum.y), atomicAdd(dst + 2, sum.z), atomicAdd(dst + 3, sum.w);
                }
            }
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_acc), "r"(TMEM_COLS) : "memory");
}
} // namespace tc
} // namespace cpprl_b200
