// cnn_gemm.cuh — fp32 tiled GEMM with pluggable operand loaders and epilogue, the one contraction kernel behind every
// layer of CnnBase (src/model/cnn_base.cpp:13-22): convolutions as IMPLICIT GEMMs (the im2col matrix is never
// materialised: the A loader gathers it from the activation tensor), the 1568 -> 512 linear layer, and their
// data / weight gradients.
//
//   C[m][n] (+)= sum_k A(m, k) * B(k, n)          m < M, n < N, k in this CTA's K slice (grid.z when split_k)
//
// Loaders return 16 bytes:  XLoad(r, c) = X[r][c .. c+3] of a logical row-major matrix X (c % 4 == 0), zero outside it.
//   A_T = false:  A(m, k) = X[m][k]   (vectors along k, stored transposed into shared memory)
//   A_T = true :  A(m, k) = X[k][m]   (vectors along m; weight gradients: the reduction runs over X's rows)
//   B(k, n) = Y[k][n]                 (vectors along n)
// Epilogue: epi(m, n, float4 acc) for n % 4 == 0, called only for m < M (n may exceed N when N % 4 != 0: never here).
//
// 256 threads, BK = 16, double-buffered shared memory with register prefetch, FFMA2 (packed fp32x2) accumulators
// paired along m.  fp32 throughout: post-update parameters must match the reference to 1e-4 after hundreds of Adam
// steps, which TF32 / BF16 operands do not give.
#pragma once
#include <cuda_runtime.h>

namespace cpprl_b200
{
constexpr int kGemmThreads = 256;
constexpr int kGemmBK = 16;

template <int BM, int BN, int TM, int TN, bool A_T, class ALoad, class BLoad, class Epi>
__global__ void __launch_bounds__(kGemmThreads) cnn_gemm_kernel(const ALoad al, const BLoad bl, const Epi epi, int M, int N,
                                                                int K, int k_slice, int z_is_k_split, const int *stop_flag)
{
    constexpr int TX = BN / TN, TY = BM / TM;
    static_assert(TX * TY == kGemmThreads, "thread tile");
    static_assert(TM % 4 == 0 && TN % 4 == 0, "micro tile");
    constexpr int LDA = BM + 4, LDB = BN + 4;
    __shared__ __align__(16) float As[2][kGemmBK][LDA];
    __shared__ __align__(16) float Bs[2][kGemmBK][LDB];
    if (stop_flag && *stop_flag)
        return; // KL early stop already triggered in this update (ppo.cpp:100-104)
    const int tid = threadIdx.x, tx = tid % TX, ty = tid / TX;
    const int m0 = blockIdx.x * BM, n0 = blockIdx.y * BN;
    // grid.z either splits the reduction (weight gradients) or selects an operand set the loaders pick up themselves
    const int k_begin = z_is_k_split ? blockIdx.z * k_slice : 0, k_end = min(K, k_begin + k_slice);
    constexpr int A_VECS = BM * kGemmBK / 4, B_VECS = BN * kGemmBK / 4;
    constexpr int A_PER = (A_VECS + kGemmThreads - 1) / kGemmThreads, B_PER = (B_VECS + kGemmThreads - 1) / kGemmThreads;
    float4 ra[A_PER], rb[B_PER];

    auto fetch = [&](int k0)
    {
#pragma unroll
        for (int i = 0; i < A_PER; ++i)
        {
            const int idx = tid + i * kGemmThreads;
            ra[i] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (A_VECS % kGemmThreads == 0 || idx < A_VECS)
            {
                if (A_T)
                {
                    const int m4 = idx % (BM / 4), kk = idx / (BM / 4);
                    if (k0 + kk < k_end)
                        ra[i] = al(k0 + kk, m0 + 4 * m4);
                }
                else
                {
                    const int mm = idx % BM, k4 = idx / BM;
                    if (k0 + 4 * k4 < k_end)
                        ra[i] = al(m0 + mm, k0 + 4 * k4);
                }
            }
        }
#pragma unroll
        for (int i = 0; i < B_PER; ++i)
        {
            const int idx = tid + i * kGemmThreads;
            rb[i] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (B_VECS % kGemmThreads == 0 || idx < B_VECS)
            {
                const int n4 = idx % (BN / 4), kk = idx / (BN / 4);
                if (k0 + kk < k_end)
                    rb[i] = bl(k0 + kk, n0 + 4 * n4);
            }
        }
    };
    auto stash = [&](int buf)
    {
#pragma unroll
        for (int i = 0; i < A_PER; ++i)
        {
            const int idx = tid + i * kGemmThreads;
            if (A_VECS % kGemmThreads == 0 || idx < A_VECS)
            {
                if (A_T)
                {
                    const int m4 = idx % (BM / 4), kk = idx / (BM / 4);
                    *reinterpret_cast<float4 *>(&As[buf][kk][4 * m4]) = ra[i];
                }
                else
                {
                    const int mm = idx % BM, k4 = idx / BM;
                    As[buf][4 * k4 + 0][mm] = ra[i].x;
                    As[buf][4 * k4 + 1][mm] = ra[i].y;
                    As[buf][4 * k4 + 2][mm] = ra[i].z;
                    As[buf][4 * k4 + 3][mm] = ra[i].w;
                }
            }
        }
#pragma unroll
        for (int i = 0; i < B_PER; ++i)
        {
            const int idx = tid + i * kGemmThreads;
            if (B_VECS % kGemmThreads == 0 || idx < B_VECS)
            {
                const int n4 = idx % (BN / 4), kk = idx / (BN / 4);
                *reinterpret_cast<float4 *>(&Bs[buf][kk][4 * n4]) = rb[i];
            }
        }
    };

    // accumulators: pairs of consecutive m (one FFMA2 = two rows of one column)
    float2 acc[TM / 2][TN];
#pragma unroll
    for (int i = 0; i < TM / 2; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j)
            acc[i][j] = make_float2(0.f, 0.f);

    if (k_begin < k_end)
    {
        fetch(k_begin);
        stash(0);
    }
    __syncthreads();
    int buf = 0;
    for (int k0 = k_begin; k0 < k_end; k0 += kGemmBK)
    {
        const bool more = k0 + kGemmBK < k_end;
        if (more)
            fetch(k0 + kGemmBK); // global loads of the next tile fly during this tile's FMAs
#pragma unroll
        for (int kk = 0; kk < kGemmBK; ++kk)
        {
            float4 av[TM / 4], bv[TN / 4];
#pragma unroll
            for (int i = 0; i < TM / 4; ++i)
                av[i] = *reinterpret_cast<const float4 *>(&As[buf][kk][4 * ty + i * (4 * TY)]);
#pragma unroll
            for (int j = 0; j < TN / 4; ++j)
                bv[j] = *reinterpret_cast<const float4 *>(&Bs[buf][kk][4 * tx + j * (4 * TX)]);
#pragma unroll
            for (int i = 0; i < TM / 4; ++i)
            {
                const float2 alo = make_float2(av[i].x, av[i].y), ahi = make_float2(av[i].z, av[i].w);
#pragma unroll
                for (int j = 0; j < TN / 4; ++j)
                {
                    const float bs[4] = {bv[j].x, bv[j].y, bv[j].z, bv[j].w};
#pragma unroll
                    for (int c = 0; c < 4; ++c)
                    {
                        const float2 b2 = make_float2(bs[c], bs[c]);
                        acc[2 * i][4 * j + c] = __ffma2_rn(alo, b2, acc[2 * i][4 * j + c]);
                        acc[2 * i + 1][4 * j + c] = __ffma2_rn(ahi, b2, acc[2 * i + 1][4 * j + c]);
                    }
                }
            }
        }
        if (more)
            stash(buf ^ 1);
        __syncthreads();
        buf ^= 1;
    }

    // epilogue: rows m = m0 + 4 ty + i (4 TY) + r, columns n = n0 + 4 tx + j (4 TX) + c
#pragma unroll
    for (int i = 0; i < TM / 4; ++i)
#pragma unroll
        for (int r = 0; r < 4; ++r)
        {
            const int m = m0 + 4 * ty + i * (4 * TY) + r;
            if (m >= M)
                continue;
#pragma unroll
            for (int j = 0; j < TN / 4; ++j)
            {
                const int n = n0 + 4 * tx + j * (4 * TX);
                if (n >= N)
                    continue;
                const float2 *a2 = acc[2 * i + (r >> 1)];
                float4 v;
                v.x = (r & 1) ? a2[4 * j + 0].y : a2[4 * j + 0].x;
                v.y = (r & 1) ? a2[4 * j + 1].y : a2[4 * j + 1].x;
                v.z = (r & 1) ? a2[4 * j + 2].y : a2[4 * j + 2].x;
                v.w = (r & 1) ? a2[4 * j + 3].y : a2[4 * j + 3].x;
                epi(m, n, v);
            }
        }
}
} // namespace cpprl_b200
