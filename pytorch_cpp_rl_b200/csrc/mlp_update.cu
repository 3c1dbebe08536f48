// mlp_update.cu — fused on-policy update for the non-recurrent MlpBase policy (PPO and A2C).
//
// Replaces, in ONE persistent cooperative kernel per epoch:
//   minibatch gather            src/generators/feed_forward_generator.cpp:42-80
//   observation normalise/clamp src/observation_normalizer.cpp:69-74
//   actor/critic towers + heads src/model/mlp_base.cpp:47-64, src/model/output_layers.cpp:39-59
//   log-prob / entropy          src/distributions/categorical.cpp:29-61, src/distributions/normal.cpp:24-37
//   PPO / A2C loss              src/algorithms/ppo.cpp:95-134, src/algorithms/a2c.cpp:63-80
//   autograd backward           src/algorithms/ppo.cpp:138, src/algorithms/a2c.cpp:84   (hand-derived here)
//   Adam / RMSprop step         src/algorithms/ppo.cpp:140, src/algorithms/a2c.cpp:85   (libtorch optimisers)
//   KL early stop, statistics   src/algorithms/ppo.cpp:96-104,143-145,156-177
//
// Structure: the grid is G co-resident CTAs.  Each minibatch ("step") is
//   phase 1  every CTA gathers tiles of R rows through the permutation, runs forward, loss and backward
//            entirely out of shared memory (weights resident as a pre-permuted "image", activations
//            feature-major [feature][row] so every GEMM operand is a conflict-free 16-byte LDS) and
//            accumulates its gradient partial in shared memory, then stores it to gpart[cta][*];
//   barrier
//   phase 2  the partials (with the loss/KL partial sums appended) are reduced in a fixed order, the KL
//            stop is decided identically by every CTA, and each CTA applies the optimiser to its slice of
//            the flat parameter vector, writing both the parameter and its image location(s);
//   barrier, image reload.
// fp32 CUDA-core FMAs throughout: the 64-wide contractions stay inside the 1e-4 parity budget only in
// full fp32; tensor cores are reserved for the CNN base.
#include <math.h>

#include "common.cuh"
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

// Shared-memory carve-up (offsets in floats), identical on host and device.
struct SmemPlan
{
    int img, gsm, x, h1[2], h2[2], logit, dlogit, val, dval, ret, oldlp, adv, act, rowid, misc, total;
};

__host__ __device__ inline SmemPlan make_plan(int R, int img_floats, int Pstride, int Dp, int Ap)
{
    const int LD = R + 4;
    SmemPlan p;
    int o = 0;
    p.img = o, o += img_floats;
    p.gsm = o, o += (Pstride + 3) & ~3;
    p.x = o, o += Dp * LD;
    for (int t = 0; t < 2; ++t)
        p.h1[t] = o, o += kHid * LD;
    for (int t = 0; t < 2; ++t)
        p.h2[t] = o, o += kHid * LD;
    p.logit = o, o += Ap * LD;
    p.dlogit = o, o += Ap * LD;
    p.val = o, o += LD;
    p.dval = o, o += LD;
    p.ret = o, o += LD;
    p.oldlp = o, o += LD;
    p.adv = o, o += LD;
    p.act = o, o += Ap * LD;
    p.rowid = o, o += 2 * LD; // long long per row
    p.misc = o, o += 32;
    p.total = o;
    return p;
}

// ---- weight image ---------------------------------------------------------------------------------
// Writes parameter i (value v) to its location(s) in the image.
__device__ __forceinline__ void image_scatter(const MlpUpdateArgs &a, float *__restrict__ image, int i, float v)
{
    const ParamOffsets &po = a.po;
    const ImageLayout &il = a.il;
    const int D = a.D;
#pragma unroll
    for (int t = 0; t < 2; ++t)
    {
        if (i >= po.w1[t] && i < po.w1[t] + kHid * D)
        {
            const int e = i - po.w1[t], j = e / D, k = e - j * D;
            image[il.w1t[t] + k * kHid + perm16(j)] = v;
            return;
        }
        if (i >= po.b1[t] && i < po.b1[t] + kHid)
        {
            image[il.b1[t] + perm16(i - po.b1[t])] = v;
            return;
        }
        if (i >= po.w2[t] && i < po.w2[t] + kHid * kHid)
        {
            const int e = i - po.w2[t], j = e >> 6, k = e & 63;
            image[il.w2t[t] + k * kHid + perm16(j)] = v;
            image[il.w2n[t] + j * kHid + perm16(k)] = v;
            return;
        }
        if (i >= po.b2[t] && i < po.b2[t] + kHid)
        {
            image[il.b2[t] + perm16(i - po.b2[t])] = v;
            return;
        }
    }
    if (i >= po.wv && i < po.wv + kHid)
        image[il.wv + (i - po.wv)] = v;
    else if (i == po.bv)
        image[il.bv] = v;
    else if (po.slog >= 0 && i >= po.slog && i < po.slog + a.A)
        image[il.slog + (i - po.slog)] = v;
    else if (i >= po.wo && i < po.wo + a.A * kHid)
        image[il.wo + (i - po.wo)] = v;
    else if (i >= po.bo && i < po.bo + a.A)
        image[il.bo + (i - po.bo)] = v;
}

// ---- GEMM building blocks (per tower: 128 threads, u = thread index inside the tower) ------------------
// out[j][r] = sum_k Wc[k][perm(j)] * in[k][r],  thread owns j in {jb, jb+16, jb+32, jb+48} and rows
// {32g + 4rb .. +3}.  Wc is contraction-major with the permuted output column, so the four weights are one
// 16-byte LDS; `in` rows are one 16-byte LDS per group (broadcast across the 16 jb lanes).
template <int R>
__device__ __forceinline__ void gemm_cf(const float *__restrict__ Wc, const float *__restrict__ in, int K, int jb,
                                        int rb, float (&acc)[4][Tile<R>::NG * 4])
{
    constexpr int LD = Tile<R>::LD, NG = Tile<R>::NG;
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int c = 0; c < NG * 4; ++c)
            acc[a][c] = 0.f;
    const float *wp = Wc + 4 * jb;
    const float *ip = in + 4 * rb;
#pragma unroll 4
    for (int k = 0; k < K; ++k)
    {
        const float4 w = *reinterpret_cast<const float4 *>(wp + k * kHid);
#pragma unroll
        for (int g = 0; g < NG; ++g)
        {
            const float4 x = *reinterpret_cast<const float4 *>(ip + k * LD + 32 * g);
            const float xs[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
            for (int i = 0; i < 4; ++i)
            {
                acc[0][g * 4 + i] = fmaf(w.x, xs[i], acc[0][g * 4 + i]);
                acc[1][g * 4 + i] = fmaf(w.y, xs[i], acc[1][g * 4 + i]);
                acc[2][g * 4 + i] = fmaf(w.z, xs[i], acc[2][g * 4 + i]);
                acc[3][g * 4 + i] = fmaf(w.w, xs[i], acc[3][g * 4 + i]);
            }
        }
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
template <int R>
__device__ __forceinline__ void grad_w2(const float *__restrict__ dz, const float *__restrict__ act, int u,
                                        float *__restrict__ gout)
{
    constexpr int LD = Tile<R>::LD;
    const int kq = u & 15, jq = u >> 4;
    float acc[8][4];
#pragma unroll
    for (int a = 0; a < 8; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b)
            acc[a][b] = 0.f;
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
#pragma unroll
            for (int b = 0; b < 4; ++b)
            {
                acc[a][b] = fmaf(av.x, bv[b].x, acc[a][b]);
                acc[a][b] = fmaf(av.y, bv[b].y, acc[a][b]);
                acc[a][b] = fmaf(av.z, bv[b].z, acc[a][b]);
                acc[a][b] = fmaf(av.w, bv[b].w, acc[a][b]);
            }
        }
    }
#pragma unroll
    for (int a = 0; a < 8; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b)
            gout[(jq + 8 * a) * kHid + kq + 16 * b] += acc[a][b];
}

// dW1[j][k] += sum_r dz[j][r] * x[k][r]  (64 x D), thread owns j = u % 64 and half of the k range.
template <int R>
__device__ __forceinline__ void grad_w1(const float *__restrict__ dz, const float *__restrict__ x, int u, int D, int Dp,
                                        float *__restrict__ gout)
{
    constexpr int LD = Tile<R>::LD;
    const int j = u & 63, half = u >> 6;
    const int k_begin = half * (Dp >> 1), k_end = k_begin + (Dp >> 1);
    for (int kc = k_begin; kc < k_end; kc += 4)
    {
        float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll 4
        for (int r = 0; r < R; r += 4)
        {
            const float4 av = *reinterpret_cast<const float4 *>(dz + j * LD + r);
#pragma unroll
            for (int kk = 0; kk < 4; ++kk)
            {
                const float4 xv = *reinterpret_cast<const float4 *>(x + (kc + kk) * LD + r);
                acc[kk] = fmaf(av.x, xv.x, acc[kk]);
                acc[kk] = fmaf(av.y, xv.y, acc[kk]);
                acc[kk] = fmaf(av.z, xv.z, acc[kk]);
                acc[kk] = fmaf(av.w, xv.w, acc[kk]);
            }
        }
#pragma unroll
        for (int kk = 0; kk < 4; ++kk)
            if (kc + kk < D)
                gout[j * D + kc + kk] += acc[kk];
    }
}

__device__ __forceinline__ float block_bcast_sum_stats(const float *gpart, int G, int Pstride, int slot, int lane)
{
    float s = 0.f;
    for (int g = lane; g < G; g += 32)
        s += __ldcg(gpart + static_cast<size_t>(g) * Pstride + slot);
    return warp_sum(s);
}

// ---- the kernel -------------------------------------------------------------------------------------
template <int R>
__global__ void __launch_bounds__(kUpdThreads, 1) mlp_update_kernel(const MlpUpdateArgs a)
{
    constexpr int LD = Tile<R>::LD;
    extern __shared__ __align__(16) float smem[];
    const SmemPlan sp = make_plan(R, a.il.floats, a.po.Pstride, a.Dp, a.Ap);
    float *img = smem + sp.img;
    float *gsm = smem + sp.gsm;
    float *X = smem + sp.x;
    float *H1[2] = {smem + sp.h1[0], smem + sp.h1[1]};
    float *H2[2] = {smem + sp.h2[0], smem + sp.h2[1]};
    float *LOGIT = smem + sp.logit, *DLOGIT = smem + sp.dlogit;
    float *VAL = smem + sp.val, *DVAL = smem + sp.dval, *RET = smem + sp.ret, *OLDLP = smem + sp.oldlp,
          *ADV = smem + sp.adv, *ACT = smem + sp.act;
    long long *ROWID = reinterpret_cast<long long *>(smem + sp.rowid);
    float *MISC = smem + sp.misc;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tower = tid >> 7, u = tid & 127; // warps 0-3 actor, 4-7 critic
    const int jb = u & 15, rb = u >> 4;
    const int G = gridDim.x;
    const int P = a.po.P, Pstride = a.po.Pstride, Pstat = Pstride - kNumStats;
    const int D = a.D, Dp = a.Dp, A = a.A;
    const bool training = a.algo != kAlgoForward;

    // ---- prologue: read persistent state, build the weight image from the parameters -----------------
    int stopped = 0, applied = 0;
    long long opt_step0 = 0;
    if (training)
    {
        stopped = a.state->stopped;
        opt_step0 = a.state->opt_step;
    }
    for (int i = blockIdx.x * kUpdThreads + tid; i < P; i += G * kUpdThreads)
        image_scatter(a, a.image, i, a.params[i]);
    if (tid == 0)
    {
        float mean = 0.f, denom = 1.f;
        if (a.adv_sums)
        {
            const double cnt = static_cast<double>(a.rows);
            const double m = a.adv_sums[0] / cnt;
            double var = (a.adv_sums[1] - a.adv_sums[0] * m) / (cnt - 1.0); // unbiased (ppo.cpp:57)
            if (var < 0.0)
                var = 0.0;
            mean = static_cast<float>(m);
            denom = static_cast<float>(sqrt(var)) + 1e-5f;
        }
        MISC[0] = mean;
        MISC[1] = denom;
    }
    grid_barrier(a.barrier, G);
    if (stopped)
        return;

    const float inv_b = 1.f / static_cast<float>(a.B);
    const int n_tiles = (a.B + R - 1) / R;

    for (int step = 0; step < a.n_steps; ++step)
    {
        // (re)load the weight image, clear the gradient accumulator
        {
            const float4 *src = reinterpret_cast<const float4 *>(a.image);
            float4 *dst = reinterpret_cast<float4 *>(img);
            for (int i = tid; i < a.il.floats / 4; i += kUpdThreads)
                dst[i] = __ldcg(src + i);
            for (int i = tid; i < Pstride; i += kUpdThreads)
                gsm[i] = 0.f;
        }
        __syncthreads();
        const float adv_mean = MISC[0], adv_denom = MISC[1];

        for (int tile = blockIdx.x; tile < n_tiles; tile += G)
        {
            const int row0 = tile * R;
            const int nrows = min(R, a.B - row0);
            // ---- P0: gather the tile through the permutation ---------------------------------------
            if (tid < R)
            {
                long long row = -1;
                if (tid < nrows)
                {
                    const long long pos = static_cast<long long>(step) * a.B + row0 + tid;
                    row = a.perm ? a.perm[pos] : pos;
                }
                ROWID[tid] = row;
                float ret = 0.f, vold = 0.f, olp = 0.f;
                if (row >= 0 && training)
                {
                    ret = a.returns[row];
                    olp = a.old_logp[row];
                    if (a.adv_sums)
                        vold = a.value_preds[row];
                }
                RET[tid] = ret;
                OLDLP[tid] = olp;
                ADV[tid] = (ret - vold - adv_mean) / adv_denom; // (adv - mean) / (std + 1e-5)
                if (a.head == 0)
                {
                    long long act = row >= 0 ? static_cast<const long long *>(a.actions)[row] : 0;
                    reinterpret_cast<int *>(ACT)[tid] = static_cast<int>(act);
                }
            }
            __syncthreads();
            for (int e = tid; e < Dp * R; e += kUpdThreads)
            {
                const int k = e / R, r = e - k * R; // R is a power of two
                const long long row = ROWID[r];
                float v = 0.f;
                if (row >= 0 && k < D)
                {
                    v = a.obs[row * D + k];
                    if (a.norm_mean)
                    {
                        // clamp((x - mean) / sqrt(var + 1e-8), -clip, clip)   (observation_normalizer.cpp:71-73)
                        const float c = a.norm_clip[0];
                        v = (v - a.norm_mean[k]) / sqrtf(a.norm_var[k] + 1e-8f);
                        v = fminf(fmaxf(v, -c), c);
                    }
                }
                X[k * LD + r] = v;
            }
            if (a.head == 1)
                for (int e = tid; e < A * R; e += kUpdThreads)
                {
                    const int k = e / R, r = e - k * R;
                    const long long row = ROWID[r];
                    ACT[k * LD + r] = row >= 0 ? static_cast<const float *>(a.actions)[row * A + k] : 0.f;
                }
            __syncthreads();

            // ---- P1: layer 1, both towers ------------------------------------------------------------
            {
                float acc[4][Tile<R>::NG * 4];
                gemm_cf<R>(img + a.il.w1t[tower], X, Dp, jb, rb, acc);
                store_tanh<R>(H1[tower], img + a.il.b1[tower], jb, rb, acc);
            }
            __syncthreads();
            // ---- P2: layer 2 ---------------------------------------------------------------------------
            {
                float acc[4][Tile<R>::NG * 4];
                gemm_cf<R>(img + a.il.w2t[tower], H1[tower], kHid, jb, rb, acc);
                store_tanh<R>(H2[tower], img + a.il.b2[tower], jb, rb, acc);
            }
            __syncthreads();
            // ---- P3: heads: logits / loc [A][R] from the actor tower, value [R] from the critic tower ----
            for (int e = tid; e < (A + 1) * R; e += kUpdThreads)
            {
                const int o = e / R, r = e - o * R;
                const bool is_value = o == A;
                const float *w = is_value ? img + a.il.wv : img + a.il.wo + o * kHid;
                const float *h = is_value ? H2[1] : H2[0];
                float s0 = 0.f, s1 = 0.f;
#pragma unroll 8
                for (int k = 0; k < kHid; k += 2)
                {
                    s0 = fmaf(w[k], h[k * LD + r], s0);
                    s1 = fmaf(w[k + 1], h[(k + 1) * LD + r], s1);
                }
                const float s = s0 + s1 + (is_value ? img[a.il.bv] : img[a.il.bo + o]);
                if (is_value)
                    VAL[r] = s;
                else
                    LOGIT[o * LD + r] = s;
            }
            __syncthreads();
            // ---- P4: log-prob, entropy, loss and its gradient w.r.t. head outputs (one thread per row) ----
            if (warp < R / 32)
            {
                const int r = tid;
                const bool valid = r < nrows;
                float st_v = 0.f, st_a = 0.f, st_e = 0.f, st_kl = 0.f, st_clip = 0.f;
                float z[kMaxActions], dz[kMaxActions];
                float newlp = 0.f, ent = 0.f;
                const float value = VAL[r];
                if (a.head == 0)
                {
                    // Categorical(logits): l = z - logsumexp(z); p = softmax(l)      (categorical.cpp:35-36)
                    float m = -INFINITY;
                    for (int i = 0; i < A; ++i)
                    {
                        z[i] = LOGIT[i * LD + r];
                        m = fmaxf(m, z[i]);
                    }
                    float se = 0.f;
                    for (int i = 0; i < A; ++i)
                        se += expf(z[i] - m);
                    const float lse = m + logf(se);
                    const int act = reinterpret_cast<const int *>(ACT)[r];
                    for (int i = 0; i < A; ++i)
                    {
                        z[i] -= lse;          // normalised log-probabilities
                        dz[i] = expf(z[i]);   // probabilities
                        ent -= z[i] * dz[i];  // entropy = -sum p log p                  (categorical.cpp:48-52)
                    }
                    newlp = z[min(max(act, 0), A - 1)]; // log_prob = gather(l, action)     (categorical.cpp:54-61)
                }
                else
                {
                    // Normal(loc, exp(scale_log))                                        (normal.cpp:24-37)
                    for (int i = 0; i < A; ++i)
                    {
                        const float sl = img[a.il.slog + i];
                        const float sigma = expf(sl);
                        const float var = sigma * sigma;
                        const float log_sigma = logf(sigma);
                        const float diff = ACT[i * LD + r] - LOGIT[i * LD + r];
                        newlp += -(diff * diff) / (2.f * var) - log_sigma - 0.91893853320467274178f;
                        ent += 0.5f + 0.5f * 1.8378770664093454836f + log_sigma;
                        z[i] = diff / var;                  // d logp / d loc
                        dz[i] = diff * diff / var - 1.f;    // d logp / d scale_log
                    }
                }
                float g_lp = 0.f, dval = 0.f;
                if (training && valid)
                {
                    const float ret = RET[r];
                    if (a.algo == kAlgoPPO)
                    {
                        const float olp = OLDLP[r], adv = ADV[r];
                        st_kl = olp - newlp;                                   // ppo.cpp:96
                        const float ratio = expf(newlp - olp);                 // ppo.cpp:107
                        const float s1 = ratio * adv;
                        const float rc = fminf(fmaxf(ratio, a.clip_lo), a.clip_hi);
                        const float s2 = rc * adv;
                        st_clip = fabsf(ratio - 1.f) > a.clip_param ? 1.f : 0.f; // ppo.cpp:116-122
                        st_a = fminf(s1, s2);                                  // action_loss = -mean(min)
                        const bool in_range = ratio >= a.clip_lo && ratio <= a.clip_hi;
                        const bool active = in_range || s1 < s2;
                        g_lp = active ? -a.c_actor * inv_b * adv * ratio : 0.f;
                        const float dv = ret - value;
                        st_v = dv * dv;                                        // value_loss = 0.5*mean
                        dval = -a.c_value * inv_b * dv;
                    }
                    else
                    {
                        const float adv = ret - value;                         // a2c.cpp:69
                        st_v = adv * adv;                                      // a2c.cpp:72
                        st_a = adv * newlp;                                    // a2c.cpp:75
                        g_lp = -inv_b * adv;
                        dval = -2.f * a.c_value * inv_b * adv;
                    }
                    st_e = ent;
                }
                const float ce_b = a.c_entropy * inv_b;
                if (a.head == 0)
                {
                    const int act = reinterpret_cast<const int *>(ACT)[r];
                    for (int i = 0; i < A; ++i)
                    {
                        const float p = dz[i];
                        float d = g_lp * ((i == act ? 1.f : 0.f) - p) + ce_b * p * (z[i] + ent);
                        DLOGIT[i * LD + r] = (training && valid) ? d : 0.f;
                    }
                }
                else
                {
                    for (int i = 0; i < A; ++i)
                    {
                        DLOGIT[i * LD + r] = (training && valid) ? g_lp * z[i] : 0.f;
                        float ds = (training && valid) ? g_lp * dz[i] - ce_b : 0.f;
                        ds = warp_sum(ds);
                        if (lane == 0 && training)
                            atomicAdd(&gsm[a.po.slog + i], ds);
                    }
                }
                DVAL[r] = dval;
                if (!training)
                {
                    if (valid)
                    {
                        a.out_value[row0 + r] = value;
                        a.out_logp[row0 + r] = newlp;
                    }
                    float e = valid ? ent : 0.f;
                    e = warp_sum(e);
                    if (lane == 0)
                        atomicAdd(a.out_entropy_sum, e);
                }
                else
                {
                    st_v = warp_sum(st_v), st_a = warp_sum(st_a), st_e = warp_sum(st_e);
                    st_kl = warp_sum(st_kl), st_clip = warp_sum(st_clip);
                    float db = warp_sum(dval);
                    if (lane == 0)
                    {
                        atomicAdd(&gsm[Pstat + 0], st_v);
                        atomicAdd(&gsm[Pstat + 1], st_a);
                        atomicAdd(&gsm[Pstat + 2], st_e);
                        atomicAdd(&gsm[Pstat + 3], st_kl);
                        atomicAdd(&gsm[Pstat + 4], st_clip);
                        atomicAdd(&gsm[a.po.bv], db);
                    }
                    for (int i = 0; i < A; ++i)
                    {
                        float d = warp_sum(DLOGIT[i * LD + r]);
                        if (lane == 0)
                            atomicAdd(&gsm[a.po.bo + i], d);
                    }
                }
            }
            if (!training)
            {
                __syncthreads();
                continue;
            }
            __syncthreads();
            // ---- P5: head weight gradients and dz2 = (head^T dout) * (1 - h2^2), in place over H2 ----------
            {
                const int k = u & 63, half = u >> 6;
                const int nout = tower == 0 ? A : 1;
                const float *dout = tower == 0 ? DLOGIT : DVAL;
                const float *wh = tower == 0 ? img + a.il.wo : img + a.il.wv;
                float *h2 = H2[tower] + k * LD;
                float wk[kMaxActions], gacc[kMaxActions];
                for (int o = 0; o < nout; ++o)
                {
                    wk[o] = wh[o * kHid + k];
                    gacc[o] = 0.f;
                }
                float bsum = 0.f;
                const int r_begin = half * (R / 2);
                for (int r = r_begin; r < r_begin + R / 2; r += 4)
                {
                    const float4 h = *reinterpret_cast<const float4 *>(h2 + r);
                    float4 dh = make_float4(0.f, 0.f, 0.f, 0.f);
                    for (int o = 0; o < nout; ++o)
                    {
                        const float4 d = *reinterpret_cast<const float4 *>(dout + o * LD + r);
                        gacc[o] += d.x * h.x + d.y * h.y + d.z * h.z + d.w * h.w;
                        dh.x = fmaf(d.x, wk[o], dh.x);
                        dh.y = fmaf(d.y, wk[o], dh.y);
                        dh.z = fmaf(d.z, wk[o], dh.z);
                        dh.w = fmaf(d.w, wk[o], dh.w);
                    }
                    float4 dz;
                    dz.x = dh.x * (1.f - h.x * h.x);
                    dz.y = dh.y * (1.f - h.y * h.y);
                    dz.z = dh.z * (1.f - h.z * h.z);
                    dz.w = dh.w * (1.f - h.w * h.w);
                    bsum += dz.x + dz.y + dz.z + dz.w;
                    *reinterpret_cast<float4 *>(h2 + r) = dz;
                }
                const int wbase = tower == 0 ? a.po.wo : a.po.wv;
                for (int o = 0; o < nout; ++o)
                    atomicAdd(&gsm[wbase + o * kHid + k], gacc[o]);
                atomicAdd(&gsm[a.po.b2[tower] + k], bsum);
            }
            __syncthreads();
            // ---- P6: dW2 += dz2 h1^T ------------------------------------------------------------------------
            grad_w2<R>(H2[tower], H1[tower], u, gsm + a.po.w2[tower]);
            __syncthreads();
            // ---- P7: dz1 = (W2^T dz2) * (1 - h1^2), in place over H1; db1 -----------------------------------
            {
                float acc[4][Tile<R>::NG * 4];
                gemm_cf<R>(img + a.il.w2n[tower], H2[tower], kHid, jb, rb, acc);
#pragma unroll
                for (int q = 0; q < 4; ++q)
                {
                    float bsum = 0.f;
#pragma unroll
                    for (int g = 0; g < Tile<R>::NG; ++g)
                    {
                        float *hp = H1[tower] + (jb + 16 * q) * LD + 32 * g + 4 * rb;
                        const float4 h = *reinterpret_cast<const float4 *>(hp);
                        float4 dz;
                        dz.x = acc[q][g * 4 + 0] * (1.f - h.x * h.x);
                        dz.y = acc[q][g * 4 + 1] * (1.f - h.y * h.y);
                        dz.z = acc[q][g * 4 + 2] * (1.f - h.z * h.z);
                        dz.w = acc[q][g * 4 + 3] * (1.f - h.w * h.w);
                        bsum += dz.x + dz.y + dz.z + dz.w;
                        *reinterpret_cast<float4 *>(hp) = dz;
                    }
                    atomicAdd(&gsm[a.po.b1[tower] + jb + 16 * q], bsum);
                }
            }
            __syncthreads();
            // ---- P8: dW1 += dz1 x^T ---------------------------------------------------------------------------
            grad_w1<R>(H1[tower], X, u, D, Dp, gsm + a.po.w1[tower]);
            __syncthreads();
        } // tiles

        if (!training)
            break;

        // ---- publish this CTA's partial ----------------------------------------------------------------------
        {
            float *dst = a.gpart + static_cast<size_t>(blockIdx.x) * Pstride;
            for (int i = tid; i < Pstride; i += kUpdThreads)
                __stcg(dst + i, gsm[i]);
        }
        grid_barrier(a.barrier, G);

        // ---- phase 2: statistics, KL stop, optimiser on this CTA's slice --------------------------------------
        if (warp < 5)
        {
            const float s = block_bcast_sum_stats(a.gpart, G, Pstride, Pstat + warp, lane);
            if (lane == 0)
                MISC[8 + warp] = s;
        }
        if (tid == 32 * 5)
        {
            // bias corrections as host-style doubles (libtorch Adam), t = steps applied so far + 1
            const double t = static_cast<double>(opt_step0 + applied + 1);
            const double bc1 = 1.0 - pow(static_cast<double>(a.beta1), t);
            const double bc2 = 1.0 - pow(static_cast<double>(a.beta2), t);
            MISC[16] = static_cast<float>(a.lr / bc1);
            MISC[17] = static_cast<float>(sqrt(bc2));
        }
        __syncthreads();
        const float sum_v = MISC[8], sum_a = MISC[9], sum_e = MISC[10], sum_kl = MISC[11], sum_clip = MISC[12];
        if (a.algo == kAlgoPPO)
        {
            const float kl = sum_kl * inv_b;
            const bool stop = static_cast<double>(kl) > a.kl_limit; // ppo.cpp:100
            if (blockIdx.x == 0 && tid == 0)
            {
                UpdateState *st = a.state;
                st->last_kl = kl;
                if (stop)
                {
                    st->stopped = 1;
                    st->kl_stop_update = static_cast<float>(st->num_updates);
                }
                else
                {
                    st->total_value_loss += 0.5f * (sum_v * inv_b);
                    st->total_action_loss += -(sum_a * inv_b);
                    st->total_entropy += sum_e * inv_b;
                    st->total_clip_fraction += sum_clip * inv_b;
                    st->num_updates += 1;
                    st->opt_step = opt_step0 + applied + 1;
                }
            }
            if (stop)
                return; // identical decision in every CTA (same data, same summation order)
        }
        else if (blockIdx.x == 0 && tid == 0)
        {
            UpdateState *st = a.state;
            st->a2c_value_loss = sum_v * inv_b;
            st->a2c_action_loss = -(sum_a * inv_b);
            st->a2c_entropy = sum_e * inv_b;
            st->num_updates += 1;
        }
        {
            const float step_size = MISC[16], bc2_sqrt = MISC[17];
            const float lr = static_cast<float>(a.lr);
            for (int i = blockIdx.x * kUpdThreads + tid; i < P; i += G * kUpdThreads)
            {
                float g = 0.f;
                int c = 0;
                for (; c + 4 <= G; c += 4)
                {
                    const float g0 = __ldcg(a.gpart + static_cast<size_t>(c) * Pstride + i);
                    const float g1 = __ldcg(a.gpart + static_cast<size_t>(c + 1) * Pstride + i);
                    const float g2 = __ldcg(a.gpart + static_cast<size_t>(c + 2) * Pstride + i);
                    const float g3 = __ldcg(a.gpart + static_cast<size_t>(c + 3) * Pstride + i);
                    g += (g0 + g1) + (g2 + g3);
                }
                for (; c < G; ++c)
                    g += __ldcg(a.gpart + static_cast<size_t>(c) * Pstride + i);
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
                image_scatter(a, a.image, i, p);
            }
        }
        ++applied;
        if (step + 1 < a.n_steps)
            grid_barrier(a.barrier, G);
    } // steps
}

template <int R>
int launch_update(const MlpUpdateArgs &args, int grid, size_t smem_bytes, cudaStream_t s)
{
    static bool configured = false;
    if (!configured)
    {
        CPPRL_CUDA_TRY(cudaFuncSetAttribute(mlp_update_kernel<R>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        configured = true;
    }
    void *kargs[] = {const_cast<MlpUpdateArgs *>(&args)};
    CPPRL_CUDA_TRY(cudaLaunchCooperativeKernel(reinterpret_cast<const void *>(mlp_update_kernel<R>), dim3(grid),
                                               dim3(kUpdThreads), kargs, smem_bytes, s));
    count_launch();
    return 0;
}
} // namespace

// ---- host-side helpers used by cabi_update.cu ------------------------------------------------------------
size_t mlp_update_smem_bytes(int R, int img_floats, int Pstride, int Dp, int Ap)
{
    return sizeof(float) * static_cast<size_t>(make_plan(R, img_floats, Pstride, Dp, Ap).total);
}

int mlp_update_launch(int R, const MlpUpdateArgs &args, int grid, cudaStream_t s)
{
    const size_t smem = mlp_update_smem_bytes(R, args.il.floats, args.po.Pstride, args.Dp, args.Ap);
    if (smem > 227 * 1024)
        return fail(CPPRL_B200_EUNSUPPORTED, "mlp_update: shared-memory plan exceeds 227 KB for this obs/action size");
    if (R == 64)
        return launch_update<64>(args, grid, smem, s);
    if (R == 32)
        return launch_update<32>(args, grid, smem, s);
    return fail(CPPRL_B200_EINVAL, "mlp_update: unsupported tile size");
}
} // namespace cpprl_b200
