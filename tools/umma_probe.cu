// umma_probe.cu — layout discovery for tcgen05.mma kind::tf32 shared-memory descriptors on sm_100a.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o tools/umma_probe tools/umma_probe.cu && tools/umma_probe
//
// One MMA (K = 8) per experiment.  The PROBED operand's shared-memory region is filled with its own word index
// (as a float: integers < 2048 are exact in TF32), the other operand is an identity pattern in the known-good
// K-major SWIZZLE_NONE layout, so D shows which WORD the tensor core read for element (row, k) of the probed
// operand.  The host compares the observed map with the formula each descriptor is supposed to mean and prints
// a verdict per experiment; nothing here is on the product path.
#include <cstdio>
#include <cstdint>
#include <cstring>
#include <vector>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); return 1; } } while (0)

__device__ __forceinline__ uint32_t smem_addr(const void *p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

__device__ __forceinline__ uint64_t make_desc(uint32_t addr, uint32_t lbo_bytes, uint32_t sbo_bytes)
{
    return static_cast<uint64_t>((addr >> 4) & 0x3fffu) | (static_cast<uint64_t>((lbo_bytes >> 4) & 0x3fffu) << 16) |
           (static_cast<uint64_t>((sbo_bytes >> 4) & 0x3fffu) << 32) | (1ull << 46);
}

struct Exp
{
    int probe_a;       // 1: probe A, 0: probe B
    int mn_major;      // probed operand MN-major (instruction descriptor bit) ?
    uint32_t lbo, sbo; // descriptor fields of the probed operand (bytes)
    int M, N;          // instruction shape
    int d_lane;        // lane offset of the D address
    int region_words;  // words of the probed region filled with indices
};

// out: [128 lanes][256 cols] floats (whatever TMEM holds after the MMA; zeroed first)
__global__ void __launch_bounds__(128) probe_kernel(Exp e, float *out)
{
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ uint64_t mbar;
    __shared__ uint32_t tmem_slot;
    float *probed = reinterpret_cast<float *>(smem);               // 32 KB region
    float *other = reinterpret_cast<float *>(smem + 32768);        // 16 KB region
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    for (int i = tid; i < 8192; i += 128)
        probed[i] = i < e.region_words ? static_cast<float>(i) : 0.f;
    for (int i = tid; i < 4096; i += 128)
        other[i] = 0.f;
    __syncthreads();
    // identity in K-major no-swizzle: element (r, k) at (r/8)*128 + (r%8)*16 + (k/4)*LBO + (k%4)*4, LBO = rows/8*128
    const int other_rows = e.probe_a ? e.N : e.M;
    const uint32_t other_lbo = (other_rows / 8) * 128;
    if (tid < 8)
    {
        const int r = tid, k = tid;
        other[((r / 8) * 128 + (r % 8) * 16 + (k / 4) * other_lbo + (k % 4) * 4) / 4] = 1.f;
    }
    if (warp == 0)
    {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_addr(&tmem_slot)), "r"(256u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    if (tid == 0)
    {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(&mbar)), "r"(1u) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = tmem_slot;
    // zero the accumulator region (256 columns) with tcgen05.st
    {
        const uint32_t base = tmem + (static_cast<uint32_t>(32 * warp) << 16);
        for (int c = 0; c < 256; c += 8)
            asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %1, %1, %1, %1, %1, %1, %1};" ::"r"(base + c), "r"(0u) : "memory");
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (tid == 0)
    {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const uint64_t dp = make_desc(smem_addr(probed), e.lbo, e.sbo);
        const uint64_t dother = make_desc(smem_addr(other), other_lbo, 128);
        const uint64_t da = e.probe_a ? dp : dother, db = e.probe_a ? dother : dp;
        uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | (static_cast<uint32_t>(e.N >> 3) << 17) | (static_cast<uint32_t>(e.M >> 4) << 24);
        if (e.mn_major)
            idesc |= e.probe_a ? (1u << 15) : (1u << 16);
        const uint32_t d = tmem + (static_cast<uint32_t>(e.d_lane) << 16);
        asm volatile(
            "{\n\t"
            ".reg .pred p;\n\t"
            "setp.ne.b32 p, %4, 0;\n\t"
            "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n\t"
            "}\n" ::"r"(d), "l"(da), "l"(db), "r"(idesc), "r"(1u), "r"(0u)
            : "memory");
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
        const uint32_t base = tmem + (static_cast<uint32_t>(32 * warp) << 16);
        for (int c = 0; c < 256; c += 8)
        {
            uint32_t r[8];
            asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                         : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                         : "r"(base + c));
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            for (int i = 0; i < 8; ++i)
                out[(32 * warp + lane) * 256 + c + i] = __uint_as_float(r[i]);
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(256u) : "memory");
}

static std::vector<float> run(const Exp &e, float *d_out, bool &ok)
{
    std::vector<float> h(128 * 256, 0.f);
    cudaMemset(d_out, 0, h.size() * 4);
    probe_kernel<<<1, 128, 49152>>>(e, d_out);
    cudaError_t err = cudaDeviceSynchronize();
    ok = err == cudaSuccess;
    if (!ok)
        printf("   kernel failed: %s\n", cudaGetErrorString(err));
    cudaMemcpy(h.data(), d_out, h.size() * 4, cudaMemcpyDeviceToHost);
    return h;
}

int main()
{
    CK(cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 49152));
    float *d_out;
    CK(cudaMalloc(&d_out, 128 * 256 * 4));
    bool ok;
    // ---- 1. baseline: A K-major, M = 128, N = 16 ------------------------------------------------------------------------
    {
        Exp e{1, 0, 2048, 128, 128, 16, 0, 2048};
        auto h = run(e, d_out, ok);
        int bad = 0;
        for (int m = 0; m < 128; ++m)
            for (int k = 0; k < 8; ++k)
            {
                const int want = ((m / 8) * 128 + (m % 8) * 16 + (k / 4) * 2048 + (k % 4) * 4) / 4;
                bad += h[m * 256 + k] != static_cast<float>(want);
            }
        printf("[1] A K-major M128 (LBO=2048,SBO=128): %s (%d mismatches); D[5][0..7] =", bad ? "MISMATCH" : "as expected", bad);
        for (int k = 0; k < 8; ++k)
            printf(" %g", h[5 * 256 + k]);
        printf("\n");
    }
    // ---- 2. A MN-major: which word is read for (m, k)?  try both readings of LBO / SBO ------------------------------------
    for (int v = 0; v < 4; ++v)
    {
        // memory model we HOPE for: word(m, k) = (m%4) + (m/4)*X/4 + (k%8)*4 + ...   with X the stride between 4-row groups
        const uint32_t lbo = v == 0 ? 4096 : v == 1 ? 128 : v == 2 ? 128 : 256, sbo = v == 0 ? 128 : v == 1 ? 4096 : v == 2 ? 256 : 128;
        Exp e{1, 1, lbo, sbo, 128, 16, 0, 2048};
        auto h = run(e, d_out, ok);
        printf("[2.%d] A MN-major M128 LBO=%u SBO=%u: word(m,k) for m=0,1,4,5,8,32,64 x k=0..7:\n", v, lbo, sbo);
        const int ms[] = {0, 1, 4, 5, 8, 32, 64, 127};
        for (int m : ms)
        {
            printf("      m=%3d:", m);
            for (int k = 0; k < 8; ++k)
                printf(" %5g", h[m * 256 + k]);
            printf("\n");
        }
    }
    // ---- 3. B MN-major, N = 64 ----------------------------------------------------------------------------------------------
    for (int v = 0; v < 3; ++v)
    {
        const uint32_t lbo = v == 0 ? 2048 : v == 1 ? 128 : 256, sbo = v == 0 ? 128 : v == 1 ? 2048 : 128;
        Exp e{0, 1, lbo, sbo, 128, 64, 0, 2048};
        auto h = run(e, d_out, ok);
        printf("[3.%d] B MN-major N64 LBO=%u SBO=%u: word(n,k) = D[k][n] for n=0,1,4,5,8,32,63 x k=0..7:\n", v, lbo, sbo);
        const int ns[] = {0, 1, 4, 5, 8, 32, 63};
        for (int n : ns)
        {
            printf("      n=%3d:", n);
            for (int k = 0; k < 8; ++k)
                printf(" %5g", h[k * 256 + n]);
            printf("\n");
        }
    }
    // ---- 4. M = 64: where do the rows land?  D lane offset 0 and 16 ----------------------------------------------------------
    for (int v = 0; v < 2; ++v)
    {
        Exp e{1, 0, 1024, 128, 64, 16, v * 16, 1024};
        auto h = run(e, d_out, ok);
        printf("[4.%d] A K-major M64 N16, D lane offset %d: lane -> A word of column k=0 (row m expects word m/8*32 + m%%8*4):\n      ", v, v * 16);
        for (int l = 0; l < 128; ++l)
        {
            printf("%g ", h[l * 256 + 0]);
            if (l % 32 == 31)
                printf("\n      ");
        }
        printf("\n");
    }
    // ---- 5. N = 56 (B K-major), M = 128 ---------------------------------------------------------------------------------------
    {
        Exp e{0, 0, 7 * 128, 128, 128, 56, 0, 1024};
        auto h = run(e, d_out, ok);
        int bad = 0;
        for (int n = 0; n < 56; ++n)
            for (int k = 0; k < 8; ++k)
            {
                const int want = ((n / 8) * 128 + (n % 8) * 16 + (k / 4) * 7 * 128 + (k % 4) * 4) / 4;
                bad += h[k * 256 + n] != static_cast<float>(want);
            }
        printf("[5] B K-major N56 M128: %s (%d mismatches); column 56..63 of row 0 (must stay 0): %g %g\n", bad ? "MISMATCH" : "as expected", bad,
               h[56], h[63]);
    }
    // ---- 6. M = 128, N = 56 with B MN-major is covered by [3] + [5] if both behave ---------------------------------------------
    printf("done\n");
    return 0;
}
