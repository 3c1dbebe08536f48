// cnn_update.cu — CnnBase forward / backward / optimiser step on the GPU.
//
// Replaces, for the Nature-DQN convolutional base:
//   CnnBase::forward               src/model/cnn_base.cpp:33-47   (x / 255, three conv + ReLU, flatten, linear + ReLU,
//                                                                  critic_linear)
//   heads, log-prob / entropy, PPO / A2C loss, autograd backward, Adam / RMSprop step, KL stop, statistics
//                                  src/model/output_layers.cpp:39-59, src/algorithms/ppo.cpp:86-146, a2c.cpp:58-85
//   FeedForwardGenerator gathers   src/generators/feed_forward_generator.cpp:56-76 (rows are read through the
//                                  permutation inside the first convolution's loader; no minibatch copy exists)
//
// Every dense contraction is one instantiation of cnn_gemm_kernel (cnn_gemm.cuh): the three convolutions as implicit
// GEMMs over NHWC activations, their data gradients (conv2's stride-2 gradient split into the four input-pixel
// parity classes so that no tap is wasted), their weight gradients (reduction over rows x pixels split across the
// grid, fp32 atomics into the gradient image), and the 1568 -> 512 linear layer.  Weights are kept as "images"
// (B operands contiguous along n) rebuilt by the optimiser kernel each step.  A minibatch is processed in passes of
// ws.chunk rows so that the activation scratch stays bounded; gradients accumulate across the passes and the KL
// decision (ppo.cpp:100-104) is taken on the device before the optimiser kernel runs.
#include <math.h>

#include <cstdlib>

#include "cnn_gemm.cuh"
#include "cnn_gemm_tc.cuh"
#include "cnn_update.cuh"
#include "common.cuh"

namespace cpprl_b200
{
CnnParamOffsets cnn_param_offsets(int C, int A, int head)
{
    CnnParamOffsets po;
    int o = 0;
    po.w1 = o, o += kCnnC1 * C * 64;
    po.b1 = o, o += kCnnC1;
    po.w2 = o, o += kCnnC2 * kCnnK2;
    po.b2 = o, o += kCnnC2;
    po.w3 = o, o += kCnnC3 * kCnnK3;
    po.b3 = o, o += kCnnC3;
    po.wfc = o, o += kCnnHid * kCnnFlat;
    po.bfc = o, o += kCnnHid;
    po.wv = o, o += kCnnHid;
    po.bv = o, o += 1;
    if (head == 1)
        po.slog = o, o += A;
    else
        po.slog = -1;
    po.wo = o, o += A * kCnnHid;
    po.bo = o, o += A;
    po.P = o;
    return po;
}

CnnImageLayout cnn_image_layout(int C)
{
    CnnImageLayout il;
    int o = 0;
    il.w1 = o, o += C * 64 * kCnnC1;
    il.w2 = o, o += kCnnK2 * kCnnC2;
    il.w3 = o, o += kCnnK3 * kCnnC3;
    il.wfc = o, o += kCnnFlat * kCnnHid;
    il.grad_total = o;
    il.wfd = o, o += kCnnHid * kCnnFlat;
    il.w3d = o, o += 9 * kCnnC3 * kCnnC2;
    il.w2d = o, o += 4 * 4 * kCnnC2 * kCnnC1;
    il.total = o;
    return il;
}

size_t cnn_workspace_floats(int C, int A, int chunk, bool training)
{
    const CnnImageLayout il = cnn_image_layout(C);
    const CnnParamOffsets po = cnn_param_offsets(C, A, 1);
    size_t n = il.total + 64;
    const size_t acts = static_cast<size_t>(chunk) * (kCnnP1 * kCnnC1 + kCnnP2 * kCnnC2 + kCnnP3 * kCnnC3 + kCnnHid);
    n += acts;
    if (training)
        n += acts + il.grad_total + po.P + 64;
    return n + 64;
}

void cnn_workspace_carve(CnnWorkspace &ws, float *base, int C, int A, int chunk, bool training)
{
    const CnnImageLayout il = cnn_image_layout(C);
    const CnnParamOffsets po = cnn_param_offsets(C, A, 1);
    auto take = [&](size_t n) { float *p = base; base += (n + 15) & ~static_cast<size_t>(15); return p; };
    const size_t c = static_cast<size_t>(chunk);
    ws.chunk = chunk;
    ws.sums = take(8);
    ws.consts = take(4);
    ws.img = take(il.total);
    ws.act1 = take(c * kCnnP1 * kCnnC1);
    ws.act2 = take(c * kCnnP2 * kCnnC2);
    ws.act3 = take(c * kCnnP3 * kCnnC3);
    ws.h = take(c * kCnnHid);
    ws.gimg = ws.gflat = ws.dz1 = ws.dz2 = ws.dz3 = ws.dh = nullptr;
    if (training)
    {
        ws.gimg = take(il.grad_total);
        ws.gflat = take(po.P);
        ws.dz1 = take(c * kCnnP1 * kCnnC1);
        ws.dz2 = take(c * kCnnP2 * kCnnC2);
        ws.dz3 = take(c * kCnnP3 * kCnnC3);
        ws.dh = take(c * kCnnHid);
    }
}

namespace
{
// ---- operand loaders (see cnn_gemm.cuh: X(r, c) = X[r][c .. c+3]) -----------------------------------------------------
// Every loader splits into row(r) -> RowCtx (the divisions, the permutation look-up, the validity of the row) and
// load(ctx, c): the tensor-core kernel's K-major A operand keeps ONE row per thread for the whole kernel, so the
// context is formed once and the k loop only adds offsets.
struct RowCtx
{
    const float *p; // base pointer of the row
    int a, b;       // loader specific (pixel coordinates for the data-gradient gathers)
    int lim;        // columns of this row that exist (0: the row lies outside the matrix); a tap bit mask for the
                    // data-gradient gathers.  ONE compare per load decides between the element and the zero page.
};

// Out-of-range operand elements are read from this zero page: every load is UNCONDITIONAL with a selected address
// (no branch, no zero-initialised destination registers in the GEMM main loops).
__device__ __align__(16) float g_zero_page[4];
// base + unsigned 32-bit element offset (one IMAD.WIDE.U32), one predicate, one pointer select
__device__ __forceinline__ float4 ld_or_zero(const float *base, int off, bool ok)
{
    const float *q = base + static_cast<unsigned>(off);
    return __ldg(reinterpret_cast<const float4 *>(ok ? q : g_zero_page));
}

// im2col of the raw observations for conv1: X[(i, oy, ox)][(c, ky, kx)].  Rows come straight out of the rollout through
// the minibatch permutation; pixels stay raw 0..255 here, the division by 255 (cnn_base.cpp:37) is applied to the
// accumulated sums in the epilogues (conv is linear in its input; raw integer pixels are exact in TF32).
struct ObsIm2col
{
    static constexpr bool kLanesAlongK = false; // see cnn_gemm_tc.cuh

    const float *obs;
    const long long *perm;
    long long row_base;
    int rows, C;
    __device__ __forceinline__ RowCtx row(int r) const
    {
        const bool ok = r < rows * kCnnP1;
        r = ok ? r : 0;
        const int i = r / kCnnP1, p = r - i * kCnnP1, oy = p / kCnnO1, ox = p - oy * kCnnO1;
        const long long row = perm ? __ldg(perm + row_base + i) : row_base + i;
        return RowCtx{obs + row * C * (kCnnIn * kCnnIn) + (4 * oy) * kCnnIn + 4 * ox, 0, 0, ok ? C * 64 : 0};
    }
    __device__ __forceinline__ void clip(RowCtx &ctx, int k_end) const { ctx.lim = min(ctx.lim, k_end); }
    __device__ __forceinline__ float4 load(const RowCtx &ctx, int k) const
    {
        const int c = k >> 6, ky = (k >> 3) & 7, kx = k & 7;
        return ld_or_zero(ctx.p, c * (kCnnIn * kCnnIn) + ky * kCnnIn + kx, k < ctx.lim);
    }
    // the same offset split into a per-thread constant (column kc of a 32-column k tile) and a per-tile part (tile
    // base k0, a multiple of 32: uniform over the CTA): 64 columns per channel, so no carry between the two
    __device__ __forceinline__ int col_part(int kc) const { return (kc >> 3) * kCnnIn + (kc & 7); }
    __device__ __forceinline__ int tile_part(int k0) const { return (k0 >> 6) * (kCnnIn * kCnnIn) + ((k0 & 32) >> 3) * kCnnIn; }
    __device__ __forceinline__ float4 load_parts(const RowCtx &ctx, int part, int k) const { return ld_or_zero(ctx.p, part, k < ctx.lim); }
    __device__ __forceinline__ float4 operator()(int r, int k) const { return load(row(r), k); }
};

// the same loader for frames known to lie on the TF32 grid (see ExactOf in cnn_gemm_tc.cuh)
struct ObsIm2colExact : ObsIm2col
{
    static constexpr bool kTf32Exact = true;
};

// im2col of an NHWC activation: X[(i, oy, ox)][(ky, kx, c)]
template <int IH, int OH, int KW, int S, int CIN>
struct ActIm2col
{
    static constexpr bool kLanesAlongK = true; // see cnn_gemm_tc.cuh

    const float *act;
    int rows;
    __device__ __forceinline__ RowCtx row(int r) const
    {
        const bool ok = r < rows * OH * OH;
        r = ok ? r : 0;
        const int i = r / (OH * OH), p = r - i * (OH * OH), oy = p / OH, ox = p - oy * OH;
        return RowCtx{act + (static_cast<size_t>(i) * IH * IH + (S * oy) * IH + S * ox) * CIN, 0, 0, ok ? KW * KW * CIN : 0};
    }
    __device__ __forceinline__ void clip(RowCtx &ctx, int k_end) const { ctx.lim = min(ctx.lim, k_end); }
    __device__ __forceinline__ float4 load(const RowCtx &ctx, int k) const
    {
        const int tap = k / CIN, c = k - tap * CIN, ky = tap / KW, kx = tap - ky * KW;
        return ld_or_zero(ctx.p, (ky * IH + kx) * CIN + c, k < ctx.lim);
    }
    static_assert(CIN == 32 || CIN == 64, "a 32-column k tile lies inside one tap");
    __device__ __forceinline__ int col_part(int kc) const { return kc; }
    __device__ __forceinline__ int tile_part(int k0) const
    {
        const int tap = k0 / CIN, ky = tap / KW, kx = tap - ky * KW;
        return (ky * IH + kx) * CIN + (k0 - tap * CIN);
    }
    __device__ __forceinline__ float4 load_parts(const RowCtx &ctx, int part, int k) const { return ld_or_zero(ctx.p, part, k < ctx.lim); }
    __device__ __forceinline__ float4 operator()(int r, int k) const { return load(row(r), k); }
};

// plain row-major matrix [rows][cols]
struct MatLoad
{
    static constexpr bool kLanesAlongK = true; // see cnn_gemm_tc.cuh

    const float *p;
    int rows, cols;
    __device__ __forceinline__ RowCtx row(int r) const
    {
        const bool ok = r < rows;
        return RowCtx{p + static_cast<size_t>(ok ? r : 0) * cols, 0, 0, ok ? cols : 0};
    }
    __device__ __forceinline__ void clip(RowCtx &ctx, int k_end) const { ctx.lim = min(ctx.lim, k_end); }
    __device__ __forceinline__ float4 load(const RowCtx &ctx, int c) const { return ld_or_zero(ctx.p, c, c < ctx.lim); }
    __device__ __forceinline__ int col_part(int kc) const { return kc; }
    __device__ __forceinline__ int tile_part(int k0) const { return k0; }
    __device__ __forceinline__ float4 load_parts(const RowCtx &ctx, int part, int k) const { return ld_or_zero(ctx.p, part, k < ctx.lim); }
    __device__ __forceinline__ float4 operator()(int r, int c) const { return load(row(r), c); }
    // B operand: element (r, c) for r < row_lim (the caller folds the column check and the k range into row_lim)
    __device__ __forceinline__ int row_limit(int c, int k_end) const { return c < cols ? min(rows, k_end) : 0; }
    __device__ __forceinline__ float4 at(int r, int c, int row_lim) const { return ld_or_zero(p, r * cols + c, r < row_lim); }
};

// one matrix per blockIdx.z (conv2 data gradient: one weight slice per input-pixel parity class)
struct ClassMatLoad
{
    const float *p;
    int rows, cols;
    __device__ __forceinline__ int row_limit(int c, int k_end) const { return c < cols ? min(rows, k_end) : 0; }
    __device__ __forceinline__ float4 at(int r, int c, int row_lim) const
    {
        return ld_or_zero(p, (static_cast<int>(blockIdx.z) * rows + r) * cols + c, r < row_lim);
    }
    __device__ __forceinline__ float4 operator()(int r, int c) const { return at(r, c, row_limit(c, rows)); }
};

// conv3 data gradient: X[(i, y, x) on the 9x9 input][(tap, n)] = dz3[i][y - ky][x - kx][n]; ctx.lim = mask of the taps
// (ky, kx) whose output pixel (y - ky, x - kx) exists.  (No k split with the mask loaders: K = taps * channels.)
struct Dgrad3Load
{
    static constexpr bool kLanesAlongK = true; // see cnn_gemm_tc.cuh

    const float *dz;
    int rows;
    __device__ __forceinline__ RowCtx row(int r) const
    {
        const bool ok = r < rows * kCnnP2;
        r = ok ? r : 0;
        const int i = r / kCnnP2, p = r - i * kCnnP2, y = p / kCnnO2, x = p - y * kCnnO2;
        int mask = 0;
#pragma unroll
        for (int tap = 0; tap < 9; ++tap)
        {
            const int oy = y - tap / 3, ox = x - tap % 3;
            mask |= (oy >= 0 && oy < kCnnO3 && ox >= 0 && ox < kCnnO3) ? (1 << tap) : 0;
        }
        return RowCtx{dz + static_cast<size_t>(i) * kCnnP3 * kCnnC3, y, x, ok ? mask : 0};
    }
    __device__ __forceinline__ void clip(RowCtx &, int) const {}
    __device__ __forceinline__ float4 load(const RowCtx &ctx, int k) const
    {
        const int tap = k >> 5, n = k & 31, ky = tap / 3, kx = tap - 3 * ky;
        return ld_or_zero(ctx.p, ((ctx.a - ky) * kCnnO3 + (ctx.b - kx)) * kCnnC3 + n, (ctx.lim >> tap) & 1);
    }
    // parts: the row's own pixel offset (y * 7 + x) * 32 is added on the fly (two's complement: the sum is exact)
    __device__ __forceinline__ int col_part(int kc) const { return kc; }
    __device__ __forceinline__ int tile_part(int k0) const
    {
        const int tap = k0 >> 5, ky = tap / 3, kx = tap - 3 * ky;
        return -(ky * kCnnO3 + kx) * kCnnC3;
    }
    __device__ __forceinline__ float4 load_parts(const RowCtx &ctx, int part, int k) const
    {
        return ld_or_zero(ctx.p, (ctx.a * kCnnO3 + ctx.b) * kCnnC3 + part, (ctx.lim >> (k >> 5)) & 1);
    }
    __device__ __forceinline__ float4 operator()(int r, int k) const { return load(row(r), k); }
};

// conv2 data gradient for the input pixels (y, x) = (2 py + cy, 2 px + cx) of parity class blockIdx.z = 2 cy + cx:
// only the taps ky = cy + 2a, kx = cx + 2b reach them: X[(i, py, px)][(a, b, n)] = dz2[i][py - a][px - b][n]
struct Dgrad2Load
{
    static constexpr bool kLanesAlongK = true; // see cnn_gemm_tc.cuh

    const float *dz;
    int rows;
    __device__ __forceinline__ RowCtx row(int r) const
    {
        const bool ok = r < rows * 100;
        r = ok ? r : 0;
        const int i = r / 100, p = r - i * 100, py = p / 10, px = p - py * 10;
        int mask = 0;
#pragma unroll
        for (int tap = 0; tap < 4; ++tap)
        {
            const int oy = py - (tap >> 1), ox = px - (tap & 1);
            mask |= (oy >= 0 && oy < kCnnO2 && ox >= 0 && ox < kCnnO2) ? (1 << tap) : 0;
        }
        return RowCtx{dz + static_cast<size_t>(i) * kCnnP2 * kCnnC2, py, px, ok ? mask : 0};
    }
    __device__ __forceinline__ void clip(RowCtx &, int) const {}
    __device__ __forceinline__ float4 load(const RowCtx &ctx, int k) const
    {
        const int tap = k >> 6, n = k & 63;
        return ld_or_zero(ctx.p, ((ctx.a - (tap >> 1)) * kCnnO2 + (ctx.b - (tap & 1))) * kCnnC2 + n, (ctx.lim >> tap) & 1);
    }
    __device__ __forceinline__ int col_part(int kc) const { return kc; }
    __device__ __forceinline__ int tile_part(int k0) const
    {
        const int tap = k0 >> 6;
        return -((tap >> 1) * kCnnO2 + (tap & 1)) * kCnnC2 + (k0 & 32);
    }
    __device__ __forceinline__ float4 load_parts(const RowCtx &ctx, int part, int k) const
    {
        return ld_or_zero(ctx.p, (ctx.a * kCnnO2 + ctx.b) * kCnnC2 + part, (ctx.lim >> (k >> 6)) & 1);
    }
    __device__ __forceinline__ float4 operator()(int r, int k) const { return load(row(r), k); }
};

// ---- epilogues ------------------------------------------------------------------------------------------------------------
struct BiasReluStore // out[m][n] = relu(acc / div + bias[n])   (div = 255 for conv1: raw pixels in, see ObsIm2col)
{
    float *out;
    const float *bias;
    int ld;
    float div;
    __device__ __forceinline__ void operator()(int m, int n, float4 v) const
    {
        const float4 b = __ldg(reinterpret_cast<const float4 *>(bias + n));
        if (div != 1.f)
            v.x = v.x / div, v.y = v.y / div, v.z = v.z / div, v.w = v.w / div;
        v.x = fmaxf(v.x + b.x, 0.f), v.y = fmaxf(v.y + b.y, 0.f), v.z = fmaxf(v.z + b.z, 0.f), v.w = fmaxf(v.w + b.w, 0.f);
        *reinterpret_cast<float4 *>(out + static_cast<size_t>(m) * ld + n) = v;
    }
};

struct AtomicAddStore // gradient image, reduction split across the grid (div as above)
{
    float *out;
    int ld;
    float div;
    __device__ __forceinline__ void operator()(int m, int n, float4 v) const
    {
        float *o = out + static_cast<size_t>(m) * ld + n;
        if (div != 1.f)
            v.x = v.x / div, v.y = v.y / div, v.z = v.z / div, v.w = v.w / div;
        atomicAdd(o, v.x), atomicAdd(o + 1, v.y), atomicAdd(o + 2, v.z), atomicAdd(o + 3, v.w);
    }
};

// The two epilogues that write an activation gradient dz can also form its column sums -- the bias gradient of the layer
// that produced the activation -- on the way out (tensor-core kernel only: tc::ColSumOf): `colsum` (may be null) receives
// sum_m dz[m][n] at index n % colmod with one atomic per column and CTA, instead of a separate pass that re-reads dz from
// HBM (dz1 alone is 210 MB per 4096-row pass).
struct ReluMaskStore // dz[m][n] = acc where the forward activation was positive
{
    static constexpr bool kColSum = true;
    float *out;
    const float *act;
    int ld;
    float *colsum;
    int colmod;
    __device__ __forceinline__ float4 store(int m, int n, float4 v) const
    {
        const size_t o = static_cast<size_t>(m) * ld + n;
        const float4 a = __ldg(reinterpret_cast<const float4 *>(act + o));
        v.x = a.x > 0.f ? v.x : 0.f, v.y = a.y > 0.f ? v.y : 0.f, v.z = a.z > 0.f ? v.z : 0.f, v.w = a.w > 0.f ? v.w : 0.f;
        *reinterpret_cast<float4 *>(out + o) = v;
        return v;
    }
    __device__ __forceinline__ void operator()(int m, int n, float4 v) const { store(m, n, v); }
};

struct Dgrad2Store // rows are (i, py, px) of parity class blockIdx.z: scatter to the NHWC pixel (2 py + cy, 2 px + cx)
{
    static constexpr bool kColSum = true;
    float *out;
    const float *act;
    float *colsum;
    int colmod;
    __device__ __forceinline__ float4 store(int m, int n, float4 v) const
    {
        const int cy = blockIdx.z >> 1, cx = blockIdx.z & 1;
        const int i = m / 100, p = m - i * 100, py = p / 10, px = p - py * 10;
        const size_t o = (static_cast<size_t>(i) * kCnnP1 + (2 * py + cy) * kCnnO1 + 2 * px + cx) * kCnnC1 + n;
        const float4 a = __ldg(reinterpret_cast<const float4 *>(act + o));
        v.x = a.x > 0.f ? v.x : 0.f, v.y = a.y > 0.f ? v.y : 0.f, v.z = a.z > 0.f ? v.z : 0.f, v.w = a.w > 0.f ? v.w : 0.f;
        *reinterpret_cast<float4 *>(out + o) = v;
        return v;
    }
    __device__ __forceinline__ void operator()(int m, int n, float4 v) const { store(m, n, v); }
};

// ---- parameter index -> image index ------------------------------------------------------------------------------------
// tensor of flat parameter index i: 0 w1, 1 w2, 2 w3, 3 wfc, -1 everything else (biases, critic_linear, head)
__device__ __forceinline__ int cnn_tensor_of(const CnnParamOffsets &po, int i)
{
    if (i < po.b1)
        return 0;
    if (i >= po.w2 && i < po.b2)
        return 1;
    if (i >= po.w3 && i < po.b3)
        return 2;
    if (i >= po.wfc && i < po.bfc)
        return 3;
    return -1;
}

// position of weight element e of tensor t inside its forward image (also the gradient image)
__device__ __forceinline__ int cnn_fwd_image_index(const CnnImageLayout &il, int C, int t, int e)
{
    if (t == 0)
    {
        const int K1 = C * 64, n = e / K1, k = e - n * K1;
        return il.w1 + k * kCnnC1 + n;
    }
    if (t == 1)
    {
        const int n = e / kCnnK2, r = e - n * kCnnK2, c = r >> 4, tap = r & 15; // [n][c][ky][kx]
        return il.w2 + (tap * kCnnC1 + c) * kCnnC2 + n;
    }
    if (t == 2)
    {
        const int n = e / kCnnK3, r = e - n * kCnnK3, c = r / 9, tap = r - 9 * c;
        return il.w3 + (tap * kCnnC2 + c) * kCnnC3 + n;
    }
    const int j = e / kCnnFlat, r = e - j * kCnnFlat, c = r / kCnnP3, px = r - c * kCnnP3; // flatten order (c, y, x)
    return il.wfc + (px * kCnnC3 + c) * kCnnHid + j;
}

// scatter one weight value into every image that holds it
__device__ __forceinline__ void cnn_image_scatter(float *img, const CnnImageLayout &il, int C, int t, int e, float v)
{
    img[cnn_fwd_image_index(il, C, t, e)] = v;
    if (t == 1)
    {
        const int n = e / kCnnK2, r = e - n * kCnnK2, c = r >> 4, ky = (r >> 2) & 3, kx = r & 3;
        const int cls = (ky & 1) * 2 + (kx & 1), tap = (ky >> 1) * 2 + (kx >> 1);
        img[il.w2d + ((cls * 4 + tap) * kCnnC2 + n) * kCnnC1 + c] = v;
    }
    else if (t == 2)
    {
        const int n = e / kCnnK3, r = e - n * kCnnK3, c = r / 9, tap = r - 9 * c;
        img[il.w3d + (tap * kCnnC3 + n) * kCnnC2 + c] = v;
    }
    else if (t == 3)
    {
        const int j = e / kCnnFlat, r = e - j * kCnnFlat, c = r / kCnnP3, px = r - c * kCnnP3;
        img[il.wfd + j * kCnnFlat + px * kCnnC3 + c] = v;
    }
}

__device__ __forceinline__ int cnn_tensor_base(const CnnParamOffsets &po, int t)
{
    return t == 0 ? po.w1 : t == 1 ? po.w2 : t == 2 ? po.w3 : po.wfc;
}

__global__ void cnn_build_images_kernel(const float *__restrict__ params, CnnParamOffsets po, CnnImageLayout il, int C,
                                        float *__restrict__ img)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < po.P; i += gridDim.x * blockDim.x)
    {
        const int t = cnn_tensor_of(po, i);
        if (t >= 0)
            cnn_image_scatter(img, il, C, t, i - cnn_tensor_base(po, t), params[i]);
    }
}

// ---- bias gradients: column sums of a [rows][N] matrix (N in {32, 64, 512}) --------------------------------------------------
__global__ void __launch_bounds__(256) cnn_colsum_kernel(const float *__restrict__ x, long long rows, int N,
                                                         float *__restrict__ out, const int *stop_flag)
{
    if (stop_flag && *stop_flag)
        return;
    __shared__ float part[256];
    const int tid = threadIdx.x;
    const int cols4 = N / 4;                 // float4 columns
    const int lanes = cols4 < 256 ? cols4 : 256; // threads along n
    const int groups = 256 / lanes;          // row groups per CTA
    const int c4 = tid % lanes, g = tid / lanes;
    for (int cbase = 0; cbase < cols4; cbase += lanes)
    {
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        if (g < groups && cbase + c4 < cols4)
            for (long long r = static_cast<long long>(blockIdx.x) * groups + g; r < rows; r += static_cast<long long>(gridDim.x) * groups)
            {
                const float4 v = __ldg(reinterpret_cast<const float4 *>(x + r * N) + cbase + c4);
                acc.x += v.x, acc.y += v.y, acc.z += v.z, acc.w += v.w;
            }
        // combine the row groups of this CTA, then one atomic per column
        for (int comp = 0; comp < 4; ++comp)
        {
            part[tid] = comp == 0 ? acc.x : comp == 1 ? acc.y : comp == 2 ? acc.z : acc.w;
            __syncthreads();
            if (g == 0 && cbase + c4 < cols4)
            {
                float s = 0.f;
                for (int q = 0; q < groups; ++q)
                    s += part[q * lanes + c4];
                atomicAdd(out + 4 * (cbase + c4) + comp, s);
            }
            __syncthreads();
        }
    }
}

// ---- heads, loss and their gradients: one warp per row of the 512-wide feature matrix ----------------------------------
// critic_linear and the head linear (cnn_base.cpp:46, output_layers.cpp:39-59), Categorical / Normal log-prob and
// entropy, the PPO / A2C loss terms and d loss / d h (masked by the ReLU of the linear layer).  Head weight gradients
// accumulate in shared memory per CTA, everything per-row-scalar in lane 0's registers; one atomic per value per CTA.
constexpr int kHeadsMaxOut = kMaxActions + 1;

__global__ void __launch_bounds__(256, 1) cnn_heads_kernel(const CnnUpdateArgs a, const float *__restrict__ h,
                                                        float *__restrict__ dh, float *__restrict__ gflat,
                                                        float *__restrict__ sums, int chunk_base, int rows, float inv_b)
{
    extern __shared__ __align__(16) float hsm[];
    const int A = a.A, nout = A + 1;
    const bool training = a.algo != kAlgoForward, ppo = a.algo == kAlgoPPO;
    if (training && a.state->stopped)
        return;
    float *W = hsm, *G = hsm + nout * kCnnHid;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < nout * kCnnHid; i += 256)
    {
        const int o = i / kCnnHid, j = i - o * kCnnHid;
        W[i] = a.params[(o < A ? a.po.wo + o * kCnnHid : a.po.wv) + j];
        if (training)
            G[i] = 0.f;
    }
    float adv_mean = 0.f, adv_denom = 1.f;
    if (ppo && a.adv_sums)
    {
        const double cnt = static_cast<double>(a.rows);
        const double m = a.adv_sums[0] / cnt;
        double var = (a.adv_sums[1] - a.adv_sums[0] * m) / (cnt - 1.0); // unbiased (ppo.cpp:57)
        if (var < 0.0)
            var = 0.0;
        adv_mean = static_cast<float>(m);
        adv_denom = static_cast<float>(sqrt(var)) + 1e-5f;
    }
    const float ce_b = a.c_entropy * inv_b;
    const float bv = a.params[a.po.bv];
    __syncthreads();

    float gb[kHeadsMaxOut], gs[kMaxActions], st[5] = {0.f, 0.f, 0.f, 0.f, 0.f}, ent_sum = 0.f;
#pragma unroll
    for (int o = 0; o < kHeadsMaxOut; ++o)
        gb[o] = 0.f;
#pragma unroll
    for (int o = 0; o < kMaxActions; ++o)
        gs[o] = 0.f;

    for (int r = blockIdx.x * 8 + warp; r < rows; r += gridDim.x * 8)
    {
        float4 hv[4];
#pragma unroll
        for (int q = 0; q < 4; ++q)
            hv[q] = __ldg(reinterpret_cast<const float4 *>(h + static_cast<size_t>(r) * kCnnHid + 128 * q) + lane);
        float out[kHeadsMaxOut];
#pragma unroll
        for (int o = 0; o < kHeadsMaxOut; ++o)
        {
            out[o] = 0.f;
            if (o < nout)
            {
                float s = 0.f;
#pragma unroll
                for (int q = 0; q < 4; ++q)
                {
                    const float4 w = *reinterpret_cast<const float4 *>(W + o * kCnnHid + 128 * q + 4 * lane);
                    s = fmaf(hv[q].x, w.x, s), s = fmaf(hv[q].y, w.y, s), s = fmaf(hv[q].z, w.z, s), s = fmaf(hv[q].w, w.w, s);
                }
                out[o] = warp_sum(s);
            }
        }
        const long long row = a.perm ? __ldg(a.perm + a.row_base + chunk_base + r) : a.row_base + chunk_base + r;
        float value = bv, logit[kMaxActions];
#pragma unroll
        for (int o = 0; o < kHeadsMaxOut; ++o)
            if (o == A)
                value += out[o];
#pragma unroll
        for (int o = 0; o < kMaxActions; ++o)
            logit[o] = o < A ? out[o] + __ldg(a.params + a.po.bo + o) : 0.f;
        // log-prob, entropy and d logp / d head outputs (every lane computes the same values)
        float newlp = 0.f, ent = 0.f, dl[kMaxActions], dsl[kMaxActions], pr[kMaxActions];
        int act = 0;
        if (a.head == 0)
        {
            act = static_cast<int>(__ldg(static_cast<const long long *>(a.actions) + row));
            float m = -INFINITY;
#pragma unroll
            for (int o = 0; o < kMaxActions; ++o)
                if (o < A)
                    m = fmaxf(m, logit[o]);
            float se = 0.f;
#pragma unroll
            for (int o = 0; o < kMaxActions; ++o)
                if (o < A)
                    se += expf(logit[o] - m);
            const float lse = m + logf(se);
#pragma unroll
            for (int o = 0; o < kMaxActions; ++o)
            {
                dl[o] = dsl[o] = pr[o] = 0.f;
                if (o < A)
                {
                    const float l = logit[o] - lse; // categorical.cpp:35-36
                    pr[o] = expf(l);
                    ent -= l * pr[o];               // categorical.cpp:48-52
                    newlp += o == act ? l : 0.f;    // categorical.cpp:54-61
                    logit[o] = l;
                }
            }
        }
        else
        {
#pragma unroll
            for (int o = 0; o < kMaxActions; ++o)
            {
                dl[o] = dsl[o] = pr[o] = 0.f;
                if (o < A)
                {
                    const float sl = __ldg(a.params + a.po.slog + o), sigma = expf(sl), var = sigma * sigma;
                    const float diff = __ldg(static_cast<const float *>(a.actions) + row * A + o) - logit[o];
                    newlp += -(diff * diff) / (2.f * var) - logf(sigma) - 0.91893853320467274178f; // normal.cpp:24-37
                    ent += 0.5f + 0.5f * 1.8378770664093454836f + logf(sigma);
                    dl[o] = diff / var;                 // d logp / d loc
                    dsl[o] = diff * diff / var - 1.f;   // d logp / d scale_log
                }
            }
        }
        float g_lp = 0.f, dval = 0.f;
        if (training)
        {
            const float ret = __ldg(a.returns + row);
            if (ppo)
            {
                const float olp = __ldg(a.old_logp + row);
                const float adv = (ret - __ldg(a.value_preds + row) - adv_mean) / adv_denom;
                const float ratio = expf(newlp - olp);                       // ppo.cpp:107
                const float s1 = ratio * adv, s2 = fminf(fmaxf(ratio, a.clip_lo), a.clip_hi) * adv;
                const bool in_range = ratio >= a.clip_lo && ratio <= a.clip_hi;
                g_lp = (in_range || s1 < s2) ? -a.c_actor * inv_b * adv * ratio : 0.f;
                const float dv = ret - value;
                dval = -a.c_value * inv_b * dv;
                st[0] += dv * dv, st[1] += fminf(s1, s2), st[2] += ent, st[3] += olp - newlp;
                st[4] += fabsf(ratio - 1.f) > a.clip_param ? 1.f : 0.f;     // ppo.cpp:116-122
            }
            else
            {
                const float adv = ret - value;                               // a2c.cpp:69-75
                g_lp = -inv_b * adv;
                dval = -2.f * a.c_value * inv_b * adv;
                st[0] += adv * adv, st[1] += adv * newlp, st[2] += ent;
            }
        }
        else
        {
            if (lane == 0)
            {
                a.out_value[chunk_base + r] = value;
                if (a.out_logp)
                    a.out_logp[chunk_base + r] = newlp;
                if (a.out_dist)
                {
                    // Categorical: logit[] already holds the normalised log-probabilities; Normal: the means
#pragma unroll
                    for (int o = 0; o < kMaxActions; ++o)
                        if (o < A)
                            a.out_dist[static_cast<size_t>(chunk_base + r) * A + o] = logit[o];
                }
            }
            ent_sum += ent;
            continue;
        }
        float dout[kHeadsMaxOut];
#pragma unroll
        for (int o = 0; o < kHeadsMaxOut; ++o)
        {
            dout[o] = 0.f;
            if (o < A)
            {
                if (a.head == 0)
                    dout[o] = g_lp * ((o == act ? 1.f : 0.f) - pr[o]) + ce_b * pr[o] * (logit[o] + ent);
                else
                {
                    dout[o] = g_lp * dl[o];
                    gs[o] += g_lp * dsl[o] - ce_b;
                }
            }
            else if (o == A)
                dout[o] = dval;
            gb[o] += dout[o];
        }
        // d loss / d h (ReLU mask of the linear layer) and the head weight gradients
        float4 dv4[4];
#pragma unroll
        for (int q = 0; q < 4; ++q)
            dv4[q] = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int o = 0; o < kHeadsMaxOut; ++o)
            if (o < nout)
            {
#pragma unroll
                for (int q = 0; q < 4; ++q)
                {
                    const int j = o * kCnnHid + 128 * q + 4 * lane;
                    const float4 w = *reinterpret_cast<const float4 *>(W + j);
                    dv4[q].x = fmaf(dout[o], w.x, dv4[q].x), dv4[q].y = fmaf(dout[o], w.y, dv4[q].y);
                    dv4[q].z = fmaf(dout[o], w.z, dv4[q].z), dv4[q].w = fmaf(dout[o], w.w, dv4[q].w);
                    atomicAdd(G + j, dout[o] * hv[q].x), atomicAdd(G + j + 1, dout[o] * hv[q].y);
                    atomicAdd(G + j + 2, dout[o] * hv[q].z), atomicAdd(G + j + 3, dout[o] * hv[q].w);
                }
            }
#pragma unroll
        for (int q = 0; q < 4; ++q)
        {
            float4 d = dv4[q];
            d.x = hv[q].x > 0.f ? d.x : 0.f, d.y = hv[q].y > 0.f ? d.y : 0.f;
            d.z = hv[q].z > 0.f ? d.z : 0.f, d.w = hv[q].w > 0.f ? d.w : 0.f;
            *(reinterpret_cast<float4 *>(dh + static_cast<size_t>(r) * kCnnHid + 128 * q) + lane) = d;
        }
    }
    if (!training)
    {
        if (lane == 0 && a.out_entropy_sum)
            atomicAdd(a.out_entropy_sum, ent_sum);
        return;
    }
    __syncthreads();
    for (int i = tid; i < nout * kCnnHid; i += 256)
    {
        const int o = i / kCnnHid, j = i - o * kCnnHid;
        atomicAdd(gflat + (o < A ? a.po.wo + o * kCnnHid : a.po.wv) + j, G[i]);
    }
    if (lane == 0)
    {
#pragma unroll
        for (int o = 0; o < kHeadsMaxOut; ++o)
            if (o < A)
            {
                atomicAdd(gflat + a.po.bo + o, gb[o]);
                if (a.head == 1)
                    atomicAdd(gflat + a.po.slog + o, gs[o]);
            }
            else if (o == A)
                atomicAdd(gflat + a.po.bv, gb[o]);
#pragma unroll
        for (int i = 0; i < 5; ++i)
            atomicAdd(sums + i, st[i]);
    }
}

// ---- KL stop / statistics / optimiser constants (one thread), then the optimiser step over the flat parameters ------------
__global__ void cnn_decide_kernel(UpdateState *st, const float *sums, float *consts, int algo, float inv_b, double kl_limit,
                                  double lr, float beta1, float beta2)
{
    if (st->stopped)
        return;
    const float sum_v = sums[0], sum_a = sums[1], sum_e = sums[2], sum_kl = sums[3], sum_clip = sums[4];
    if (algo == kAlgoPPO)
    {
        const float kl = sum_kl * inv_b;
        st->last_kl = kl;
        if (static_cast<double>(kl) > kl_limit) // ppo.cpp:100
        {
            st->stopped = 1;
            st->kl_stop_update = static_cast<float>(st->num_updates);
            return;
        }
        st->total_value_loss += 0.5f * (sum_v * inv_b);
        st->total_action_loss += -(sum_a * inv_b);
        st->total_entropy += sum_e * inv_b;
        st->total_clip_fraction += sum_clip * inv_b;
        st->num_updates += 1;
        const double t = static_cast<double>(st->opt_step + 1);
        st->opt_step += 1;
        consts[0] = static_cast<float>(lr / (1.0 - pow(static_cast<double>(beta1), t)));
        consts[1] = static_cast<float>(sqrt(1.0 - pow(static_cast<double>(beta2), t)));
    }
    else
    {
        st->a2c_value_loss = sum_v * inv_b;
        st->a2c_action_loss = -(sum_a * inv_b);
        st->a2c_entropy = sum_e * inv_b;
        st->num_updates += 1;
    }
}

__global__ void __launch_bounds__(256) cnn_optimizer_kernel(const CnnUpdateArgs a, const float *__restrict__ gimg,
                                                            const float *__restrict__ gflat, const float *__restrict__ consts,
                                                            float *__restrict__ img)
{
    if (a.state->stopped)
        return;
    const float step_size = consts[0], bc2_sqrt = consts[1], lr = static_cast<float>(a.lr);
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < a.po.P; i += gridDim.x * blockDim.x)
    {
        const int t = cnn_tensor_of(a.po, i);
        const int e = t >= 0 ? i - cnn_tensor_base(a.po, t) : 0;
        const float g = t >= 0 ? gimg[cnn_fwd_image_index(a.il, a.C, t, e)] : gflat[i];
        float p = a.params[i];
        if (a.opt == kOptAdam)
        {
            const float m = a.opt_m[i] * a.beta1 + g * (1.f - a.beta1);
            const float v = a.opt_v[i] * a.beta2 + (g * g) * (1.f - a.beta2);
            a.opt_m[i] = m;
            a.opt_v[i] = v;
            p -= step_size * (m / (sqrtf(v) / bc2_sqrt + a.eps));
        }
        else
        {
            const float s = a.opt_v[i] * a.alpha + (g * g) * (1.f - a.alpha);
            a.opt_v[i] = s;
            p -= lr * (g / (sqrtf(s) + a.eps));
        }
        a.params[i] = p;
        if (t >= 0)
            cnn_image_scatter(img, a.il, a.C, t, e, p);
    }
}
} // namespace

// ---- TF32 exactness of the observations ------------------------------------------------------------------------------------
namespace
{
__global__ void __launch_bounds__(256) cnn_obs_scan_kernel(const float *__restrict__ obs, long long n4, long long n, int *__restrict__ flag)
{
    unsigned int low = 0u;
    const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n4; i += stride)
    {
        const float4 v = ld_stream4(obs + 4 * i);
        low |= __float_as_uint(v.x) | __float_as_uint(v.y) | __float_as_uint(v.z) | __float_as_uint(v.w);
    }
    if (blockIdx.x == 0 && threadIdx.x < n - 4 * n4)
        low |= __float_as_uint(obs[4 * n4 + threadIdx.x]);
    if (__any_sync(0xffffffffu, (low & 0x1fffu) != 0u) && (threadIdx.x & 31) == 0)
        *flag = 1;
}
} // namespace

int cnn_scan_observations(const float *obs, long long n, int *flag, int *exact, cudaStream_t s)
{
    *exact = 0;
    const char *off = std::getenv("CPPRL_B200_CNN_NO_EXACT"); // A/B switch for tests and measurements
    if (n <= 0 || (reinterpret_cast<uintptr_t>(obs) & 15u) != 0 || (off && off[0] == '1'))
        return 0; // nothing to scan, an unaligned view, or switched off: keep the full path
    CPPRL_CUDA_TRY(cudaMemsetAsync(flag, 0, sizeof(int), s));
    const long long n4 = n / 4;
    long long blocks = (n4 + 256 * 8 - 1) / (256 * 8);
    const long long cap = static_cast<long long>(sm_count()) * 8;
    blocks = blocks > cap ? cap : (blocks < 1 ? 1 : blocks);
    cnn_obs_scan_kernel<<<static_cast<unsigned>(blocks), 256, 0, s>>>(obs, n4, n, flag);
    CPPRL_LAUNCH_CHECK("cnn_obs_scan_kernel");
    int inexact = 1;
    CPPRL_CUDA_TRY(cudaMemcpyAsync(&inexact, flag, sizeof(int), cudaMemcpyDeviceToHost, s));
    CPPRL_CUDA_TRY(cudaStreamSynchronize(s));
    *exact = inexact == 0;
    return 0;
}

// ---- host orchestration -------------------------------------------------------------------------------------------------
namespace
{
struct PlainStore // C[m][n] = acc
{
    float *out;
    int ld;
    __device__ __forceinline__ void operator()(int m, int n, float4 v) const
    {
        *reinterpret_cast<float4 *>(out + static_cast<size_t>(m) * ld + n) = v;
    }
};

template <int BN, bool A_T, class AL, class BL, class EP>
int launch_gemm_tc(const AL &al, const BL &bl, const EP &ep, int M, int N, int K, int grid_z, bool z_is_k_split, const int *stop,
                   cudaStream_t s)
{
    auto kernel = tc::gemm_tc_kernel<BN, A_T, AL, BL, EP>;
    static std::atomic<uint64_t> configured{0};
    if (!configured_on_device(configured))
    {
        CPPRL_CUDA_TRY(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(tc::smem_bytes(BN))));
        // two CTAs per SM (BN <= 64) need the whole 228 KB carve-out; the driver's default sizes it for one
        CPPRL_CUDA_TRY(cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
        mark_configured_on_device(configured);
    }
    int k_slice = K;
    if (z_is_k_split)
        k_slice = ((K + grid_z - 1) / grid_z + tc::kBK - 1) / tc::kBK * tc::kBK;
    const dim3 grid((M + tc::kBM - 1) / tc::kBM, (N + BN - 1) / BN, grid_z);
    kernel<<<grid, tc::kThreads, tc::smem_bytes(BN), s>>>(al, bl, ep, M, N, K, k_slice, z_is_k_split ? 1 : 0, stop);
    CPPRL_CUDA_TRY(cudaGetLastError());
    count_launch();
    return 0;
}

// Tensor cores (tcgen05, 3xTF32) by default; CPPRL_B200_CNN_FP32=1 selects the fp32 FFMA kernels (same loaders and
// epilogues), kept for A/B measurements.
inline bool use_tensor_cores()
{
    static const bool on = [] {
        const char *e = std::getenv("CPPRL_B200_CNN_FP32");
        return !(e && e[0] == '1');
    }();
    return on;
}

// CPPRL_B200_CNN_COLSUM_KERNEL=1 keeps the separate column-sum launches on the tensor-core path (A/B runs)
inline bool cnn_fold_colsum()
{
    static const bool on = [] {
        const char *e = std::getenv("CPPRL_B200_CNN_COLSUM_KERNEL");
        return !(e && e[0] == '1');
    }();
    return on;
}

template <int BM, int BN, int TM, int TN, bool A_T, class AL, class BL, class EP>
int launch_gemm(const AL &al, const BL &bl, const EP &ep, int M, int N, int K, int grid_z, bool z_is_k_split, const int *stop,
                cudaStream_t s)
{
    if (use_tensor_cores())
        return launch_gemm_tc<BN, A_T>(al, bl, ep, M, N, K, grid_z, z_is_k_split, stop, s);
    int k_slice = K;
    if (z_is_k_split)
        k_slice = ((K + grid_z - 1) / grid_z + kGemmBK - 1) / kGemmBK * kGemmBK;
    const dim3 grid((M + BM - 1) / BM, (N + BN - 1) / BN, grid_z);
    cnn_gemm_kernel<BM, BN, TM, TN, A_T, AL, BL, EP><<<grid, kGemmThreads, 0, s>>>(al, bl, ep, M, N, K, k_slice,
                                                                                    z_is_k_split ? 1 : 0, stop);
    CPPRL_CUDA_TRY(cudaGetLastError());
    count_launch();
    return 0;
}

inline int splits_for(int tiles, long long k_total)
{
    // enough CTAs to fill the machine twice, each reducing at least 2048 rows
    const int sms = sm_count();
    long long z = (2ll * sms + tiles - 1) / tiles;
    const long long max_z = (k_total + 2047) / 2048;
    if (z > max_z)
        z = max_z;
    return z < 1 ? 1 : static_cast<int>(z);
}

// forward of one pass of `rows` rows starting at minibatch offset `base`
int cnn_forward_pass(const CnnUpdateArgs &a, const CnnWorkspace &ws, int base, int rows, const int *stop, cudaStream_t s)
{
    const int K1 = a.C * 64;
    const float *img = ws.img;
    const ObsIm2col obs_rows{a.obs, a.perm, a.row_base + base, rows, a.C};
    const MatLoad w1{img + a.il.w1, K1, kCnnC1};
    const BiasReluStore act1_out{ws.act1, a.params + a.po.b1, kCnnC1, 255.f};
    if (int rc = a.obs_exact ? launch_gemm<256, 32, 8, 4, false>(ObsIm2colExact{obs_rows}, w1, act1_out, rows * kCnnP1, kCnnC1, K1, 1,
                                                                 false, stop, s)
                             : launch_gemm<256, 32, 8, 4, false>(obs_rows, w1, act1_out, rows * kCnnP1, kCnnC1, K1, 1, false, stop, s))
        return rc;
    if (int rc = launch_gemm<128, 64, 8, 4, false>(ActIm2col<kCnnO1, kCnnO2, 4, 2, kCnnC1>{ws.act1, rows},
                                                   MatLoad{img + a.il.w2, kCnnK2, kCnnC2},
                                                   BiasReluStore{ws.act2, a.params + a.po.b2, kCnnC2, 1.f}, rows * kCnnP2, kCnnC2,
                                                   kCnnK2, 1, false, stop, s))
        return rc;
    if (int rc = launch_gemm<256, 32, 8, 4, false>(ActIm2col<kCnnO2, kCnnO3, 3, 1, kCnnC2>{ws.act2, rows},
                                                   MatLoad{img + a.il.w3, kCnnK3, kCnnC3},
                                                   BiasReluStore{ws.act3, a.params + a.po.b3, kCnnC3, 1.f}, rows * kCnnP3, kCnnC3,
                                                   kCnnK3, 1, false, stop, s))
        return rc;
    return launch_gemm<128, 128, 8, 8, false>(MatLoad{ws.act3, rows, kCnnFlat}, MatLoad{img + a.il.wfc, kCnnFlat, kCnnHid},
                                              BiasReluStore{ws.h, a.params + a.po.bfc, kCnnHid, 1.f}, rows, kCnnHid, kCnnFlat, 1,
                                              false, stop, s);
}

int launch_heads(const CnnUpdateArgs &a, const CnnWorkspace &ws, int base, int rows, float inv_b, cudaStream_t s)
{
    static std::atomic<uint64_t> configured{0};
    const size_t smem = static_cast<size_t>(2) * (a.A + 1) * kCnnHid * sizeof(float);
    if (!configured_on_device(configured))
    {
        CPPRL_CUDA_TRY(cudaFuncSetAttribute(cnn_heads_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            2 * kHeadsMaxOut * kCnnHid * static_cast<int>(sizeof(float))));
        mark_configured_on_device(configured);
    }
    const int grid = std::max(1, std::min(2 * sm_count(), (rows + 7) / 8));
    cnn_heads_kernel<<<grid, 256, smem, s>>>(a, ws.h, ws.dh, ws.gflat, ws.sums, base, rows, inv_b);
    CPPRL_CUDA_TRY(cudaGetLastError());
    count_launch();
    return 0;
}

int launch_colsum(const float *x, long long rows, int N, float *out, const int *stop, cudaStream_t s)
{
    const int lanes = std::min(N / 4, 256), groups = 256 / lanes;
    const int grid = static_cast<int>(std::max<long long>(1, std::min<long long>(4 * sm_count(), (rows + groups * 64 - 1) / (groups * 64))));
    cnn_colsum_kernel<<<grid, 256, 0, s>>>(x, rows, N, out, stop);
    CPPRL_CUDA_TRY(cudaGetLastError());
    count_launch();
    return 0;
}

int cnn_backward_pass(const CnnUpdateArgs &a, const CnnWorkspace &ws, int base, int rows, const int *stop, cudaStream_t s)
{
    const int K1 = a.C * 64;
    const float *img = ws.img;
    // the tensor-core kernels form the conv layers' bias gradients in the epilogue that writes dz (ReluMaskStore above)
    const bool fold = use_tensor_cores() && cnn_fold_colsum();
    // linear 1568 -> 512: weight gradient (reduction over rows), bias, data gradient masked by conv3's ReLU
    if (int rc = launch_gemm<128, 128, 8, 8, true>(MatLoad{ws.act3, rows, kCnnFlat}, MatLoad{ws.dh, rows, kCnnHid},
                                                   AtomicAddStore{ws.gimg + a.il.wfc, kCnnHid, 1.f}, kCnnFlat, kCnnHid, rows,
                                                   splits_for(13 * 4, rows), true, stop, s))
        return rc;
    if (int rc = launch_colsum(ws.dh, rows, kCnnHid, ws.gflat + a.po.bfc, stop, s))
        return rc;
    if (int rc = launch_gemm<128, 128, 8, 8, false>(MatLoad{ws.dh, rows, kCnnHid}, MatLoad{img + a.il.wfd, kCnnHid, kCnnFlat},
                                                    ReluMaskStore{ws.dz3, ws.act3, kCnnFlat, fold ? ws.gflat + a.po.b3 : nullptr, kCnnC3},
                                                    rows, kCnnFlat, kCnnHid, 1, false, stop, s))
        return rc;
    // conv3
    if (int rc = launch_gemm<256, 32, 8, 4, true>(ActIm2col<kCnnO2, kCnnO3, 3, 1, kCnnC2>{ws.act2, rows},
                                                  MatLoad{ws.dz3, rows * kCnnP3, kCnnC3},
                                                  AtomicAddStore{ws.gimg + a.il.w3, kCnnC3, 1.f}, kCnnK3, kCnnC3, rows * kCnnP3,
                                                  splits_for(use_tensor_cores() ? 5 : 3, static_cast<long long>(rows) * kCnnP3), true, stop, s))
        return rc;
    if (!fold)
        if (int rc = launch_colsum(ws.dz3, static_cast<long long>(rows) * kCnnP3, kCnnC3, ws.gflat + a.po.b3, stop, s))
            return rc;
    if (int rc = launch_gemm<128, 64, 8, 4, false>(Dgrad3Load{ws.dz3, rows}, MatLoad{img + a.il.w3d, 9 * kCnnC3, kCnnC2},
                                                   ReluMaskStore{ws.dz2, ws.act2, kCnnC2, fold ? ws.gflat + a.po.b2 : nullptr, kCnnC2},
                                                   rows * kCnnP2, kCnnC2, 9 * kCnnC3, 1, false, stop, s))
        return rc;
    // conv2
    if (int rc = launch_gemm<128, 64, 8, 4, true>(ActIm2col<kCnnO1, kCnnO2, 4, 2, kCnnC1>{ws.act1, rows},
                                                  MatLoad{ws.dz2, rows * kCnnP2, kCnnC2},
                                                  AtomicAddStore{ws.gimg + a.il.w2, kCnnC2, 1.f}, kCnnK2, kCnnC2, rows * kCnnP2,
                                                  splits_for(4, static_cast<long long>(rows) * kCnnP2), true, stop, s))
        return rc;
    if (!fold)
        if (int rc = launch_colsum(ws.dz2, static_cast<long long>(rows) * kCnnP2, kCnnC2, ws.gflat + a.po.b2, stop, s))
            return rc;
    if (int rc = launch_gemm<256, 32, 8, 4, false>(Dgrad2Load{ws.dz2, rows}, ClassMatLoad{img + a.il.w2d, 4 * kCnnC2, kCnnC1},
                                                   Dgrad2Store{ws.dz1, ws.act1, fold ? ws.gflat + a.po.b1 : nullptr, kCnnC1}, rows * 100,
                                                   kCnnC1, 4 * kCnnC2, 4, false, stop, s))
        return rc;
    // conv1 (its input needs no gradient)
    const ObsIm2col obs_rows{a.obs, a.perm, a.row_base + base, rows, a.C};
    const MatLoad dz1{ws.dz1, rows * kCnnP1, kCnnC1};
    const AtomicAddStore gw1{ws.gimg + a.il.w1, kCnnC1, 255.f};
    const int z1 = splits_for((K1 + 255) / 256 * (use_tensor_cores() ? 2 : 1), static_cast<long long>(rows) * kCnnP1);
    if (int rc = a.obs_exact ? launch_gemm<256, 32, 8, 4, true>(ObsIm2colExact{obs_rows}, dz1, gw1, K1, kCnnC1, rows * kCnnP1, z1, true, stop, s)
                             : launch_gemm<256, 32, 8, 4, true>(obs_rows, dz1, gw1, K1, kCnnC1, rows * kCnnP1, z1, true, stop, s))
        return rc;
    return fold ? 0 : launch_colsum(ws.dz1, static_cast<long long>(rows) * kCnnP1, kCnnC1, ws.gflat + a.po.b1, stop, s);
}
} // namespace

// Test hook: C[M][N] = A * B on the tensor-core path with plain row-major operands: B is [K][N];
//   variant bit 0: A is given transposed ([K][M]); bits 4..: tile columns 32 / 64 / 128
int cnn_debug_tc_gemm(int variant, const float *A, const float *B, float *C, int M, int N, int K, cudaStream_t s)
{
    const bool a_t = variant & 1;
    const int bn = (variant >> 4) & 0xff;
    const MatLoad al{A, a_t ? K : M, a_t ? M : K}, bl{B, K, N};
    const PlainStore ep{C, N};
#define CPPRL_TC_CASE(BN_, A_) \
    if (bn == BN_ && a_t == A_) \
        return launch_gemm_tc<BN_, A_>(al, bl, ep, M, N, K, 1, false, nullptr, s);
    CPPRL_TC_CASE(32, false)
    CPPRL_TC_CASE(32, true)
    CPPRL_TC_CASE(64, false)
    CPPRL_TC_CASE(64, true)
    CPPRL_TC_CASE(128, false)
    CPPRL_TC_CASE(128, true)
#undef CPPRL_TC_CASE
    return fail(CPPRL_B200_EINVAL, "debug_tc_gemm: unknown variant");
}

namespace
{
// ---- env-sharded update: sum of the gradient vector over the GPUs of the box (NVLink peer memory, comm.cuh) ---------------
// Pull protocol with one flag per CTA: CTA b copies slice b of {grad[0, n), sums[0, 8)} into this rank's exchange buffer
// (parity = sequence & 1), publishes step_flag[b] = seq (system fence + release store), then for every rank IN RANK
// ORDER waits for that rank's flag b and adds its slice (peer reads over NVLink; the own slice from local memory) —
// every rank forms the same sum in the same order, so the optimiser steps stay bit-identical box-wide.  A buffer is
// rewritten two exchanges later, which no peer can still be reading: finishing exchange s+1 needs every peer's
// record s+1, published only after that peer finished reading s.
__global__ void __launch_bounds__(256) cnn_allreduce_kernel(const CommView comm, unsigned int seq, float *__restrict__ grad,
                                                            long long n, float *__restrict__ sums, UpdateState *st)
{
    if (st->stopped)
        return;
    const int world = comm.world, me = comm.rank, b = blockIdx.x, G = gridDim.x;
    const long long total = n + 8; // n is a multiple of 4 (workspace carve), the sums ride at the end
    const long long per = ((total / 4 + G - 1) / G) * 4;
    const long long lo = static_cast<long long>(b) * per, hi = lo + per < total ? lo + per : total;
    float *mine = exchange_fbuf(comm.area[me], seq & 1u, world, comm.x_stride);
    auto src = [&](long long i) -> const float * { return i < n ? grad + i : sums + (i - n); };
    for (long long i = lo + 4ll * threadIdx.x; i < hi; i += 4ll * blockDim.x)
        *reinterpret_cast<float4 *>(mine + i) = *reinterpret_cast<const float4 *>(src(i));
    __syncthreads();
    __shared__ int ok;
    if (threadIdx.x == 0)
    {
        flag_publish(&comm.area[me]->step_flag[b], seq);
        int good = 1;
        for (int r = 0; r < world; ++r)
            if (r != me && !flag_wait(&comm.area[r]->step_flag[b], seq))
                good = 0;
        ok = good;
        if (!good)
            st->comm_error = 1;
    }
    __syncthreads();
    if (!ok)
        return;
    for (long long i = lo + 4ll * threadIdx.x; i < hi; i += 4ll * blockDim.x)
    {
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int r = 0; r < world; ++r)
        {
            const float4 v = r == me ? *reinterpret_cast<const float4 *>(mine + i)
                                     : peer_load4(exchange_fbuf(comm.area[r], seq & 1u, world, comm.x_stride) + i);
            acc.x += v.x, acc.y += v.y, acc.z += v.z, acc.w += v.w;
        }
        *reinterpret_cast<float4 *>(const_cast<float *>(src(i))) = acc;
    }
}

// box-wide advantage moments (ppo.cpp:57 over the GLOBAL rollout): the local {sum, sum of squares} sit in this rank's
// exchange area; read every rank's in rank order
__global__ void cnn_adv_exchange_kernel(const CommView comm, unsigned int seq, double *out, UpdateState *st)
{
    flag_publish(&comm.area[comm.rank]->adv_flag, seq);
    double s1 = 0.0, s2 = 0.0;
    for (int r = 0; r < comm.world; ++r)
    {
        if (r != comm.rank && !flag_wait(&comm.area[r]->adv_flag, seq))
            st->comm_error = 1;
        s1 += peer_load(comm.area[r]->adv);
        s2 += peer_load(comm.area[r]->adv + 1);
    }
    out[0] = s1;
    out[1] = s2;
}
} // namespace

int cnn_adv_exchange(const CommView &comm, unsigned int seq, double *out, UpdateState *st, cudaStream_t s)
{
    cnn_adv_exchange_kernel<<<1, 1, 0, s>>>(comm, seq, out, st);
    CPPRL_CUDA_TRY(cudaGetLastError());
    count_launch();
    return 0;
}

// floats one exchange record of the CnnBase update takes (checked against the communicator's capacity)
long long cnn_exchange_floats(int C, int A)
{
    const CnnImageLayout il = cnn_image_layout(C);
    const CnnParamOffsets po = cnn_param_offsets(C, A, 1);
    return static_cast<long long>((il.grad_total + 15) & ~15ll) + ((po.P + 15) & ~15ll) + 8;
}

int cnn_build_images(const CnnUpdateArgs &a, const CnnWorkspace &ws, cudaStream_t s)
{
    cnn_build_images_kernel<<<std::min(4 * sm_count(), (a.po.P + 255) / 256), 256, 0, s>>>(a.params, a.po, a.il, a.C, ws.img);
    CPPRL_CUDA_TRY(cudaGetLastError());
    count_launch();
    return 0;
}

int cnn_minibatch(const CnnUpdateArgs &a, const CnnWorkspace &ws, cudaStream_t s)
{
    const int *stop = &a.state->stopped;
    const bool sharded = a.comm.world > 1;
    const float inv_b = 1.f / static_cast<float>(sharded ? a.B_norm : a.B); // every mean divides by the box-wide row count
    // gradient image and flat gradients are adjacent in the workspace: one clear; then the five partial sums
    const long long grad_floats = (ws.gflat - ws.gimg) + ((a.po.P + 15) & ~15ll);
    CPPRL_CUDA_TRY(cudaMemsetAsync(ws.gimg, 0, grad_floats * sizeof(float), s));
    CPPRL_CUDA_TRY(cudaMemsetAsync(ws.sums, 0, 8 * sizeof(float), s));
    for (int base = 0; base < a.B; base += ws.chunk)
    {
        const int rows = std::min(ws.chunk, a.B - base);
        if (int rc = cnn_forward_pass(a, ws, base, rows, stop, s))
            return rc;
        if (int rc = launch_heads(a, ws, base, rows, inv_b, s))
            return rc;
        if (int rc = cnn_backward_pass(a, ws, base, rows, stop, s))
            return rc;
    }
    if (sharded)
    {
        // (a rank without rows in this minibatch contributes zeros)
        cnn_allreduce_kernel<<<std::min(sm_count(), kMaxCtas), 256, 0, s>>>(a.comm, a.xseq, ws.gimg, grad_floats, ws.sums, a.state);
        CPPRL_CUDA_TRY(cudaGetLastError());
        count_launch();
    }
    cnn_decide_kernel<<<1, 1, 0, s>>>(a.state, ws.sums, ws.consts, a.algo, inv_b, a.kl_limit, a.lr, a.beta1, a.beta2);
    CPPRL_CUDA_TRY(cudaGetLastError());
    cnn_optimizer_kernel<<<std::min(8 * sm_count(), (a.po.P + 255) / 256), 256, 0, s>>>(a, ws.gimg, ws.gflat, ws.consts, ws.img);
    CPPRL_CUDA_TRY(cudaGetLastError());
    count_launch(2);
    return 0;
}

int cnn_forward(const CnnUpdateArgs &a, const CnnWorkspace &ws, cudaStream_t s)
{
    for (int base = 0; base < a.B; base += ws.chunk)
    {
        const int rows = std::min(ws.chunk, a.B - base);
        if (int rc = cnn_forward_pass(a, ws, base, rows, nullptr, s))
            return rc;
        if (int rc = launch_heads(a, ws, base, rows, 0.f, s))
            return rc;
    }
    return 0;
}
} // namespace cpprl_b200
