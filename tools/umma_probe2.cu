// umma_probe2.cu — numerical probes for the tensor-core MLP update (tcgen05 kind::tf32, A operand from TMEM).
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o tools/umma_probe2 tools/umma_probe2.cu && tools/umma_probe2
//
// [1] D[128][64] = A[128][K] * B[64][K]^T with A in TMEM (lane = row, column = k), B in shared memory K-major
//     SWIZZLE_NONE with padded strides (LBO = 7 * 128 + 16: only 56 of the 64 rows exist, the last 8-row group reads
//     whatever follows), K = 64 as 8 chained MMAs with the A address advancing by 8 columns.  Checked against the host
//     for the 56 real columns, all 128 lanes.
// [2] the 3xTF32 split (hi * hi + hi * lo + lo * hi into one accumulator) against an fp64 host product: relative error.
// [3] N = 32 and a two-MMA "tower select": A rows 0..63 multiply B0, rows 64..127 multiply B1 (two MMAs into two column
//     ranges; the unused halves are ignored).
// Nothing here is on the product path.
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); return 1; } } while (0)

__device__ __forceinline__ uint32_t smem_addr(const void *p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ uint64_t make_desc(uint32_t addr, uint32_t lbo_bytes, uint32_t sbo_bytes)
{
    return static_cast<uint64_t>((addr >> 4) & 0x3fffu) | (static_cast<uint64_t>((lbo_bytes >> 4) & 0x3fffu) << 16) |
           (static_cast<uint64_t>((sbo_bytes >> 4) & 0x3fffu) << 32) | (1ull << 46);
}
__device__ __forceinline__ float tf32_hi(float x) { return __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xffffe000u); }

__device__ __forceinline__ void mma_ts(uint32_t d, uint32_t a_tmem, uint64_t db, uint32_t idesc, uint32_t acc)
{
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, {%5, %5, %5, %5}, p;\n\t"
        "}\n" ::"r"(d),
        "r"(a_tmem), "l"(db), "r"(idesc), "r"(acc), "r"(0u)
        : "memory");
}

constexpr int K = 64, NB = 64, RB = 56;           // B: 56 real rows, N = 64 instruction
constexpr uint32_t SBO = 128, LBO = 7 * 128 + 16; // K-major no-swizzle, 8-row groups 128 B apart, k chunks LBO apart
__host__ __device__ inline int b_off(int n, int k) { return (k / 4) * LBO + (n / 8) * SBO + (n % 8) * 16 + (k % 4) * 4; }

// mode 0: plain tf32 (A, B already on the tf32 grid); mode 1: 3xTF32 of arbitrary fp32 A, B
__global__ void __launch_bounds__(128) probe(const float *A, const float *B0, const float *B1, float *out, int mode)
{
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ uint64_t mbar;
    __shared__ uint32_t slot;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    unsigned char *b0h = smem, *b0l = smem + 16384, *b1h = smem + 32768, *b1l = smem + 49152;
    for (int i = tid; i < 65536 / 4; i += 128)
        reinterpret_cast<float *>(smem)[i] = 1e30f; // whatever follows the real rows must not matter
    __syncthreads();
    for (int i = tid; i < RB * K; i += 128)
    {
        const int n = i / K, k = i % K;
        const float x0 = B0[i], x1 = B1[i];
        const float h0 = mode ? tf32_hi(x0) : x0, h1 = mode ? tf32_hi(x1) : x1;
        *reinterpret_cast<float *>(b0h + b_off(n, k)) = h0;
        *reinterpret_cast<float *>(b0l + b_off(n, k)) = x0 - h0;
        *reinterpret_cast<float *>(b1h + b_off(n, k)) = h1;
        *reinterpret_cast<float *>(b1l + b_off(n, k)) = x1 - h1;
    }
    if (warp == 0)
    {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_addr(&slot)), "r"(512u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    if (tid == 0)
    {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(&mbar)), "r"(1u) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = slot;
    // A -> TMEM: lane = row (tid), columns [0, 64) hi, [64, 128) lo
    {
        const uint32_t base = tmem + (static_cast<uint32_t>(32 * warp) << 16);
        for (int c = 0; c < K; c += 8)
        {
            uint32_t h[8], l[8];
            for (int i = 0; i < 8; ++i)
            {
                const float x = A[tid * K + c + i];
                const float hi = mode ? tf32_hi(x) : x;
                h[i] = __float_as_uint(hi);
                l[i] = __float_as_uint(x - hi);
            }
            asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"r"(base + c), "r"(h[0]),
                         "r"(h[1]), "r"(h[2]), "r"(h[3]), "r"(h[4]), "r"(h[5]), "r"(h[6]), "r"(h[7])
                         : "memory");
            asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"r"(base + 64 + c), "r"(l[0]),
                         "r"(l[1]), "r"(l[2]), "r"(l[3]), "r"(l[4]), "r"(l[5]), "r"(l[6]), "r"(l[7])
                         : "memory");
        }
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (tid == 0)
    {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | (static_cast<uint32_t>(NB >> 3) << 17) | (static_cast<uint32_t>(128 >> 4) << 24);
        const uint32_t d0 = tmem + 128, d1 = tmem + 192; // accumulators: columns [128, 192) for B0, [192, 256) for B1
        for (int b = 0; b < 2; ++b)
        {
            const uint32_t d = b ? d1 : d0;
            unsigned char *bh = b ? b1h : b0h, *bl = b ? b1l : b0l;
            // mode 1: interleaved order (hi*hi, hi*lo, lo*hi per k-step); mode 2: cross terms first, then the hi*hi chain
            if (mode == 2)
                for (int ks = 0; ks < K / 8; ++ks)
                {
                    const uint64_t dbh = make_desc(smem_addr(bh) + ks * 2 * LBO, LBO, SBO);
                    const uint64_t dbl = make_desc(smem_addr(bl) + ks * 2 * LBO, LBO, SBO);
                    mma_ts(d, tmem + ks * 8, dbl, idesc, ks > 0 ? 1u : 0u); // hi * lo
                    mma_ts(d, tmem + 64 + ks * 8, dbh, idesc, 1u);          // lo * hi
                }
            for (int ks = 0; ks < K / 8; ++ks)
            {
                const uint64_t dbh = make_desc(smem_addr(bh) + ks * 2 * LBO, LBO, SBO);
                const uint64_t dbl = make_desc(smem_addr(bl) + ks * 2 * LBO, LBO, SBO);
                mma_ts(d, tmem + ks * 8, dbh, idesc, (ks > 0 || mode == 2) ? 1u : 0u); // hi * hi
                if (mode == 1)
                {
                    mma_ts(d, tmem + ks * 8, dbl, idesc, 1u);      // hi * lo
                    mma_ts(d, tmem + 64 + ks * 8, dbh, idesc, 1u); // lo * hi
                }
            }
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_addr(&mbar)) : "memory");
    }
    {
        uint32_t done = 0;
        while (!done)
            asm volatile(
                "{\n\t"
                ".reg .pred p;\n\t"
                "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                "selp.u32 %0, 1, 0, p;\n\t"
                "}\n"
                : "=r"(done)
                : "r"(smem_addr(&mbar)), "r"(0u)
                : "memory");
    }
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    {
        const uint32_t base = tmem + (static_cast<uint32_t>(32 * warp) << 16) + 128;
        for (int c = 0; c < 128; c += 8)
        {
            uint32_t r[8];
            asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                         : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                         : "r"(base + c));
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            for (int i = 0; i < 8; ++i)
                out[tid * 128 + c + i] = __uint_as_float(r[i]);
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u) : "memory");
}

int main()
{
    CK(cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
    std::vector<float> A(128 * K), B0(RB * K), B1(RB * K), out(128 * 128);
    float *dA, *dB0, *dB1, *dOut;
    CK(cudaMalloc(&dA, A.size() * 4));
    CK(cudaMalloc(&dB0, B0.size() * 4));
    CK(cudaMalloc(&dB1, B1.size() * 4));
    CK(cudaMalloc(&dOut, out.size() * 4));
    for (int mode = 0; mode < 3; ++mode)
    {
        srand(7 + mode);
        auto rnd = [&]() { return mode ? (rand() / (float)RAND_MAX * 2.f - 1.f) * 1.37f : static_cast<float>(rand() % 17 - 8); };
        for (auto &v : A) v = rnd();
        for (auto &v : B0) v = rnd();
        for (auto &v : B1) v = rnd();
        CK(cudaMemcpy(dA, A.data(), A.size() * 4, cudaMemcpyHostToDevice));
        CK(cudaMemcpy(dB0, B0.data(), B0.size() * 4, cudaMemcpyHostToDevice));
        CK(cudaMemcpy(dB1, B1.data(), B1.size() * 4, cudaMemcpyHostToDevice));
        CK(cudaMemset(dOut, 0, out.size() * 4));
        probe<<<1, 128, 65536>>>(dA, dB0, dB1, dOut, mode);
        CK(cudaDeviceSynchronize());
        CK(cudaMemcpy(out.data(), dOut, out.size() * 4, cudaMemcpyDeviceToHost));
        double max_rel = 0.0, max_abs = 0.0;
        int bad = 0;
        for (int b = 0; b < 2; ++b)
            for (int m = 0; m < 128; ++m)
                for (int n = 0; n < RB; ++n)
                {
                    double want = 0.0, mag = 0.0;
                    const std::vector<float> &B = b ? B1 : B0;
                    for (int k = 0; k < K; ++k)
                        want += (double)A[m * K + k] * B[n * K + k], mag += std::fabs((double)A[m * K + k] * B[n * K + k]);
                    const double got = out[m * 128 + b * 64 + n];
                    const double err = std::fabs(got - want);
                    max_abs = std::max(max_abs, err);
                    max_rel = std::max(max_rel, err / std::max(mag, 1e-30));
                    bad += mode == 0 && got != want;
                }
        if (mode == 0)
            printf("[1] TMEM-A x smem-B (56 real rows, N=64, K=64 chained), both B operands, all 128 lanes: %s (%d mismatches of %d)\n",
                   bad ? "MISMATCH" : "exact", bad, 2 * 128 * RB);
        else
            printf("[2] 3xTF32 (%s) vs fp64: max |err| / sum|a b| = %.3e, max abs %.3e  (fp32 rounding of a 64-term sum ~ 1e-7)\n",
                   mode == 1 ? "interleaved" : "cross terms first", max_rel, max_abs);
        printf("    sample D[5][3] = %g, D[100][64+40] = %g, garbage column D[5][60] = %g\n", out[5 * 128 + 3], out[100 * 128 + 64 + 40], out[5 * 128 + 60]);
    }
    printf("done\n");
    return 0;
}
