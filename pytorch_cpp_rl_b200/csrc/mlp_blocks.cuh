// mlp_blocks.cuh — shared-memory GEMM building blocks of the fused update kernels (MLP towers of MlpBase, with or
// without the GRU front).  Activations are feature-major [feature][row] with row stride LD = R + 4; weights are
// pre-permuted "images" so that every operand fetch is a conflict-free 16-byte LDS; FMAs are packed FFMA2.
#pragma once
#include <cuda_runtime.h>

#include "mlp_update.cuh"

namespace cpprl_b200
{
namespace
{
__device__ __forceinline__ int perm16(int j) { return 4 * (j & 15) + (j >> 4); } // image column of feature j

template <int R>
struct Tile
{
    static constexpr int LD = R + 4;           // row stride of feature-major activations (LD % 32 == 4)
    static constexpr int RV = 4;               // rows per vector access
    static constexpr int NG = R / 32;          // float4 row groups per thread in the F/B GEMMs
    static_assert(R == 32 || R == 64, "tile rows");
};

// ---- GEMM building blocks (per tower: 128 threads, u = thread index inside the tower) ------------------
// out[j][r] = sum_k Wc[k][perm(j)] * in[k][r],  thread owns j in {jb, jb+16, jb+32, jb+48} and rows
// {32g + 4rb .. +3}.  Wc is contraction-major with the permuted output column, so the four weights are one
// 16-byte LDS; `in` rows are one 16-byte LDS per group (broadcast across the 16 jb lanes).
template <int R>
__device__ __forceinline__ void gemm_cf(const float *__restrict__ Wc, const float *__restrict__ in, int K, int jb,
                                        int rb, float (&acc)[4][Tile<R>::NG * 4])
{
    constexpr int LD = Tile<R>::LD, NG = Tile<R>::NG;
    // accumulators as row pairs: one FFMA2 (Blackwell packed fp32x2 FMA) does two rows of one output feature,
    // halving the issue slots the 32 FMAs of an iteration need
    float2 acc2[4][NG * 2];
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int c = 0; c < NG * 2; ++c)
            acc2[a][c] = make_float2(0.f, 0.f);
    const float *wp = Wc + 4 * jb;
    const float *ip = in + 4 * rb;
    // operands of iteration k+1 are fetched before the FMAs of iteration k (two warps per scheduler cannot
    // hide the 29-cycle LDS latency on their own)
    float4 w_next = *reinterpret_cast<const float4 *>(wp);
    float4 x_next[NG];
#pragma unroll
    for (int g = 0; g < NG; ++g)
        x_next[g] = *reinterpret_cast<const float4 *>(ip + 32 * g);
#pragma unroll 4
    for (int k = 0; k < K; ++k)
    {
        const float4 w = w_next;
        float4 x[NG];
#pragma unroll
        for (int g = 0; g < NG; ++g)
            x[g] = x_next[g];
        const int kn = k + 1 < K ? k + 1 : k;
        w_next = *reinterpret_cast<const float4 *>(wp + kn * kHid);
#pragma unroll
        for (int g = 0; g < NG; ++g)
            x_next[g] = *reinterpret_cast<const float4 *>(ip + kn * LD + 32 * g);
        const float2 wa[4] = {make_float2(w.x, w.x), make_float2(w.y, w.y), make_float2(w.z, w.z), make_float2(w.w, w.w)};
#pragma unroll
        for (int g = 0; g < NG; ++g)
        {
            const float2 xlo = make_float2(x[g].x, x[g].y), xhi = make_float2(x[g].z, x[g].w);
#pragma unroll
            for (int a = 0; a < 4; ++a)
            {
                acc2[a][g * 2] = __ffma2_rn(wa[a], xlo, acc2[a][g * 2]);
                acc2[a][g * 2 + 1] = __ffma2_rn(wa[a], xhi, acc2[a][g * 2 + 1]);
            }
        }
    }
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int c = 0; c < NG * 2; ++c)
        {
            acc[a][2 * c] = acc2[a][c].x;
            acc[a][2 * c + 1] = acc2[a][c].y;
        }
}

// h = tanh(acc + bias) stored feature-major.
template <int R>
__device__ __forceinline__ void store_tanh(float *__restrict__ out, const float *__restrict__ bias_perm, int jb, int rb,
                                           float (&acc)[4][Tile<R>::NG * 4])
{
    constexpr int LD = Tile<R>::LD, NG = Tile<R>::NG;
    const float4 b = *reinterpret_cast<const float4 *>(bias_perm + 4 * jb);
    const float bs[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int g = 0; g < NG; ++g)
        {
            float4 h;
            h.x = tanhf(acc[a][g * 4 + 0] + bs[a]);
            h.y = tanhf(acc[a][g * 4 + 1] + bs[a]);
            h.z = tanhf(acc[a][g * 4 + 2] + bs[a]);
            h.w = tanhf(acc[a][g * 4 + 3] + bs[a]);
            *reinterpret_cast<float4 *>(out + (jb + 16 * a) * LD + 32 * g + 4 * rb) = h;
        }
}

// dW[j][k] += sum_r dz[j][r] * act[k][r]  (64 x 64), thread owns j in {jq + 8a}, k in {kq + 16b}.
// Each accumulator is a pair (sum over even rows, sum over odd rows) fed by FFMA2 straight from the 16-byte
// operand loads; the halves are added once at the end.
template <int R>
__device__ __forceinline__ void grad_w2(const float *__restrict__ dz, const float *__restrict__ act, int u,
                                        float *__restrict__ gout)
{
    constexpr int LD = Tile<R>::LD;
    const int kq = u & 15, jq = u >> 4;
    float2 acc[8][4];
#pragma unroll
    for (int a = 0; a < 8; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b)
            acc[a][b] = make_float2(0.f, 0.f);
#pragma unroll 2
    for (int r = 0; r < R; r += 4)
    {
        float4 bv[4];
#pragma unroll
        for (int b = 0; b < 4; ++b)
            bv[b] = *reinterpret_cast<const float4 *>(act + (kq + 16 * b) * LD + r);
#pragma unroll
        for (int a = 0; a < 8; ++a)
        {
            const float4 av = *reinterpret_cast<const float4 *>(dz + (jq + 8 * a) * LD + r);
            const float2 alo = make_float2(av.x, av.y), ahi = make_float2(av.z, av.w);
#pragma unroll
            for (int b = 0; b < 4; ++b)
            {
                acc[a][b] = __ffma2_rn(alo, make_float2(bv[b].x, bv[b].y), acc[a][b]);
                acc[a][b] = __ffma2_rn(ahi, make_float2(bv[b].z, bv[b].w), acc[a][b]);
            }
        }
    }
#pragma unroll
    for (int a = 0; a < 8; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b)
            gout[(jq + 8 * a) * kHid + kq + 16 * b] += acc[a][b].x + acc[a][b].y;
}

// Same contraction as grad_w2, accumulated into registers the caller keeps across tiles (no shared-memory
// gradient image): acc[a][b] pairs (even rows, odd rows) of element (jq + 8a, kq + 16b).
template <int R>
__device__ __forceinline__ void grad_w2_acc(const float *__restrict__ dz, const float *__restrict__ act, int u,
                                            float2 (&acc)[8][4])
{
    constexpr int LD = Tile<R>::LD;
    const int kq = u & 15, jq = u >> 4;
#pragma unroll 2
    for (int r = 0; r < R; r += 4)
    {
        float4 bv[4];
#pragma unroll
        for (int b = 0; b < 4; ++b)
            bv[b] = *reinterpret_cast<const float4 *>(act + (kq + 16 * b) * LD + r);
#pragma unroll
        for (int a = 0; a < 8; ++a)
        {
            const float4 av = *reinterpret_cast<const float4 *>(dz + (jq + 8 * a) * LD + r);
            const float2 alo = make_float2(av.x, av.y), ahi = make_float2(av.z, av.w);
#pragma unroll
            for (int b = 0; b < 4; ++b)
            {
                acc[a][b] = __ffma2_rn(alo, make_float2(bv[b].x, bv[b].y), acc[a][b]);
                acc[a][b] = __ffma2_rn(ahi, make_float2(bv[b].z, bv[b].w), acc[a][b]);
            }
        }
    }
}

// dW1[j][k] += sum_r dz[j][r] * x[k][r]  (64 x D).  Thread owns j in {jq + 8a} and k in {kb + kq, kb + 16 + kq}
// for successive 32-wide blocks kb of the input features: the dz rows are broadcast loads shared by the 16 kq
// lanes and by both k columns, so 12 LDS.128 feed 32 FFMA2 per 4 rows.
template <int R>
__device__ __forceinline__ void grad_w1(const float *__restrict__ dz, const float *__restrict__ x, int u, int D, int Dp,
                                        float *__restrict__ gout)
{
    constexpr int LD = Tile<R>::LD;
    const int kq = u & 15, jq = u >> 4;
    for (int kb = 0; kb < Dp; kb += 32)
    {
        const int k0 = kb + kq, k1 = kb + 16 + kq;
        const bool two = kb + 16 < Dp; // warp-uniform
        const float *x0 = x + (k0 < Dp ? k0 : 0) * LD;
        const float *x1 = x + (k1 < Dp ? k1 : 0) * LD;
        float2 acc0[8], acc1[8];
#pragma unroll
        for (int a = 0; a < 8; ++a)
            acc0[a] = acc1[a] = make_float2(0.f, 0.f);
#pragma unroll 2
        for (int r = 0; r < R; r += 4)
        {
            const float4 xa = *reinterpret_cast<const float4 *>(x0 + r);
            const float4 xb = two ? *reinterpret_cast<const float4 *>(x1 + r) : make_float4(0.f, 0.f, 0.f, 0.f);
            const float2 xalo = make_float2(xa.x, xa.y), xahi = make_float2(xa.z, xa.w);
            const float2 xblo = make_float2(xb.x, xb.y), xbhi = make_float2(xb.z, xb.w);
#pragma unroll
            for (int a = 0; a < 8; ++a)
            {
                const float4 av = *reinterpret_cast<const float4 *>(dz + (jq + 8 * a) * LD + r);
                const float2 alo = make_float2(av.x, av.y), ahi = make_float2(av.z, av.w);
                acc0[a] = __ffma2_rn(alo, xalo, acc0[a]);
                acc0[a] = __ffma2_rn(ahi, xahi, acc0[a]);
                if (two)
                {
                    acc1[a] = __ffma2_rn(alo, xblo, acc1[a]);
                    acc1[a] = __ffma2_rn(ahi, xbhi, acc1[a]);
                }
            }
        }
#pragma unroll
        for (int a = 0; a < 8; ++a)
        {
            if (k0 < D)
                gout[(jq + 8 * a) * D + k0] += acc0[a].x + acc0[a].y;
            if (two && k1 < D)
                gout[(jq + 8 * a) * D + k1] += acc1[a].x + acc1[a].y;
        }
    }
}

} // namespace
} // namespace cpprl_b200
