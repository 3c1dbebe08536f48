// gru_update.cu — fused PPO update (and forward evaluation) for the GRU-fronted MlpBase.
//
// Replaces, for recurrent policies:
//   RecurrentGenerator::next       src/generators/recurrent_generator.cpp:47-98   (one environment per minibatch)
//   NNBase::forward_gru            src/model/nn_base.cpp:47-114   (run splitting at mask zeros, state never carried
//                                                                  across runs: every run restarts from hxs * mask)
//   MlpBase::forward, heads, log-prob / entropy, PPO loss, autograd backward (BPTT), Adam step, KL stop, statistics
//                                  src/model/mlp_base.cpp:47-64, src/algorithms/ppo.cpp:86-146
//
// The reference's recurrent PPO is a strictly serial chain: N minibatches per epoch, each ONE environment's T steps,
// each followed by an optimiser step.  One CTA therefore walks the whole epoch (no grid barrier, no partial-gradient
// exchange): per minibatch it
//   1. runs the GRU forward over the T steps (W_hh rows in registers, one matrix-vector product per step), keeping
//      h, r, z, n and W_hn h + b_hn on a tape in L2;
//   2. pushes the T outputs through both towers, heads, loss and the hand-derived backward in tiles of 32 rows out of
//      shared memory (same building blocks as mlp_update.cu), accumulating the 64x64 weight gradients in registers
//      and writing dL/dh_t to the tape;
//   3. decides the KL stop (ppo.cpp:100-104), applies Adam to the tower / head parameters from registers;
//   4. back-propagates through time (W_hh columns in registers; dW_hh accumulated in the same registers' twins) and
//      applies Adam to the GRU parameters.
// Forward evaluation (Policy::evaluate_actions / get_values) launches one CTA per agent.
#include <math.h>

#include "common.cuh"
#include "gru_update.cuh"
#include "mlp_blocks.cuh"

namespace cpprl_b200
{
namespace
{
constexpr int GR = 32; // tile rows of the tower phases
constexpr int GLD = Tile<GR>::LD;

struct GruSmem
{
    int w1t[2], w1n[2], w2t[2], w2n[2], b1[2], b2[2], wo, bo, wv, bv, slog, sig; // weight image
    int bih, bhh, gwih;                                                          // GRU biases, gradient of W_ih
    int x, h1[2], h2[2], logit, dlogit, val, dval, ret, oldlp, adv, act, stat;    // tower tile
    int gs;                                                                       // small gradients + statistics
    int hcur, gate, dgi, dgh, dhz, dhn, hprev, xseq, mseq, cut, misc, total;      // GRU step state
};

// offsets (in floats) inside `gs`
struct GsLayout
{
    int b1[2], b2[2], wv, bv, wo, bo, slog, stat, bih, bhh, total;
};

__host__ __device__ inline GsLayout make_gs(int A)
{
    GsLayout g;
    int o = 0;
    for (int t = 0; t < 2; ++t)
        g.b1[t] = o, o += kHid, g.b2[t] = o, o += kHid;
    g.wv = o, o += kHid;
    g.bv = o, o += 4;
    g.wo = o, o += A * kHid;
    g.bo = o, o += (A + 3) & ~3;
    g.slog = o, o += (A + 3) & ~3;
    g.stat = o, o += 8;
    g.bih = o, o += kGruGates;
    g.bhh = o, o += kGruGates;
    g.total = o;
    return g;
}

__host__ __device__ inline GruSmem make_gru_smem(int Dp, int A, int Ap, int T)
{
    GruSmem s;
    int o = 0;
    for (int t = 0; t < 2; ++t)
    {
        s.w1t[t] = o, o += kHid * kHid;
        s.w1n[t] = o, o += kHid * kHid;
        s.w2t[t] = o, o += kHid * kHid;
        s.w2n[t] = o, o += kHid * kHid;
        s.b1[t] = o, o += kHid;
        s.b2[t] = o, o += kHid;
    }
    s.wo = o, o += Ap * kHid;
    s.bo = o, o += Ap;
    s.wv = o, o += kHid;
    s.bv = o, o += 4;
    s.slog = o, o += Ap;
    s.sig = o, o += 2 * Ap;
    s.bih = o, o += kGruGates;
    s.bhh = o, o += kGruGates;
    s.x = o, o += kHid * GLD;
    s.gwih = s.x; // dW_ih accumulates during BPTT, when the tower tiles (5 * 64 * GLD floats >= 192 * Dp) are dead
    for (int t = 0; t < 2; ++t)
        s.h1[t] = o, o += kHid * GLD;
    for (int t = 0; t < 2; ++t)
        s.h2[t] = o, o += kHid * GLD;
    s.logit = o, o += Ap * GLD;
    s.dlogit = o, o += Ap * GLD;
    s.val = o, o += GLD;
    s.dval = o, o += GLD;
    s.ret = o, o += GLD;
    s.oldlp = o, o += GLD;
    s.adv = o, o += GLD;
    s.act = o, o += Ap * GLD;
    s.stat = o, o += (5 + Ap) * GLD;
    s.gs = o, o += (make_gs(A).total + 3) & ~3;
    s.hcur = o, o += kHid;
    s.gate = o, o += 4 * kHid;
    s.dgi = o, o += kGruGates;
    s.dgh = o, o += kGruGates;
    s.dhz = o, o += kHid;
    s.dhn = o, o += kHid;
    s.hprev = o, o += kHid;
    s.xseq = o, o += ((T * Dp) + 3) & ~3;
    s.mseq = o, o += (T + 3) & ~3;
    s.cut = o, o += (T + 4) & ~3;
    s.misc = o, o += 32;
    s.total = o;
    return s;
}

__device__ __forceinline__ float sigmoidf_(float x) { return 1.f / (1.f + expf(-x)); }

__global__ void __launch_bounds__(kGruThreads, 1) gru_update_kernel(const GruUpdateArgs a)
{
    extern __shared__ __align__(16) float smem[];
    const GruSmem sp = make_gru_smem(a.Dp, a.A, a.Ap, a.T);
    const GsLayout gl = make_gs(a.A);
    const GruParamOffsets &po = a.po;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tower = tid >> 7, u = tid & 127, jb = u & 15, rb = u >> 4;
    const int D = a.D, Dp = a.Dp, A = a.A, T = a.T, N = a.N;
    const bool training = a.algo != kAlgoForward;
    float *GS = smem + sp.gs;
    float *HCUR = smem + sp.hcur, *GATE = smem + sp.gate, *DGI = smem + sp.dgi, *DGH = smem + sp.dgh;
    float *DHZ = smem + sp.dhz, *DHN = smem + sp.dhn, *HPREV = smem + sp.hprev;
    float *XSEQ = smem + sp.xseq, *MSEQ = smem + sp.mseq, *MISC = smem + sp.misc;
    int *CUT = reinterpret_cast<int *>(smem + sp.cut);
    float *BIH = smem + sp.bih, *BHH = smem + sp.bhh, *GWIH = smem + sp.gwih;
    float *X = smem + sp.x;
    float *h1t = smem + sp.h1[0] + tower * (kHid * GLD), *h2t = smem + sp.h2[0] + tower * (kHid * GLD);
    float *LOGIT = smem + sp.logit, *DLOGIT = smem + sp.dlogit, *VAL = smem + sp.val, *DVAL = smem + sp.dval;
    float *RET = smem + sp.ret, *OLDLP = smem + sp.oldlp, *ADV = smem + sp.adv, *ACT = smem + sp.act, *STAT = smem + sp.stat;
    float *SIG = smem + sp.sig;
    const float *w1t_img = smem + sp.w1t[tower], *w1n_img = smem + sp.w1n[tower];
    const float *w2t_img = smem + sp.w2t[tower], *w2n_img = smem + sp.w2n[tower];
    const float *b1_img = smem + sp.b1[tower], *b2_img = smem + sp.b2[tower];
    const float *wo_img = smem + sp.wo, *bo_img = smem + sp.bo, *wv_img = smem + sp.wv, *bv_img = smem + sp.bv;
    // forward evaluation keeps only the hidden states (one tape per CTA), training all six
    float *tape = a.tape + static_cast<size_t>(blockIdx.x) * (training ? kGruTapes : 1) * T * kHid;
    float *TH = tape, *TR = tape + T * kHid, *TZ = tape + 2 * T * kHid, *TN = tape + 3 * T * kHid,
          *TG = tape + 4 * T * kHid, *TDX = tape + 5 * T * kHid;

    float adv_mean = 0.f, adv_denom = 1.f;
    long long opt_step0 = 0;
    int applied = 0;
    if (training)
    {
        if (a.state->stopped)
            return;
        opt_step0 = a.state->opt_step;
        if (a.adv_sums)
        {
            const double cnt = static_cast<double>(a.rows);
            const double m = a.adv_sums[0] / cnt;
            double var = (a.adv_sums[1] - a.adv_sums[0] * m) / (cnt - 1.0); // unbiased (ppo.cpp:57)
            if (var < 0.0)
                var = 0.0;
            adv_mean = static_cast<float>(m);
            adv_denom = static_cast<float>(sqrt(var)) + 1e-5f;
        }
    }
    // PPO: one environment per minibatch, T rows.  A2C: one batch of all T * N rows, one CTA per environment, the
    // gradient summed over the CTAs in global memory (grad_out) and applied by the RMSprop kernel that follows.
    const bool ppo = a.algo == kAlgoPPO;
    const float inv_b = 1.f / static_cast<float>(ppo ? T : T * N);
    const float ce_b = a.c_entropy * inv_b;
    const int n_seq = ppo ? a.n_mb : 1;

    for (int mb = 0; mb < n_seq; ++mb)
    {
        const int agent = ppo ? static_cast<int>(a.perm[mb]) : static_cast<int>(blockIdx.x);
        // ---- (0) weight image from the current parameters (L2 loads: the previous minibatch's Adam wrote them) ----
        for (int i = tid; i < 2 * 2 * kHid * kHid; i += kGruThreads)
        {
            const int t = i >> 13, layer = (i >> 12) & 1, e = i & 4095, j = e >> 6, k = e & 63;
            const float v = __ldcg(a.params + (layer ? po.w2[t] : po.w1[t]) + e);
            smem[(layer ? sp.w2t[t] : sp.w1t[t]) + k * kHid + perm16(j)] = v; // forward:  out[j] += W[j][k] in[k]
            smem[(layer ? sp.w2n[t] : sp.w1n[t]) + j * kHid + perm16(k)] = v; // backward: out[k] += W[j][k] dz[j]
        }
        for (int i = tid; i < 2 * kHid; i += kGruThreads)
        {
            const int t = i >> 6, j = i & 63;
            smem[sp.b1[t] + perm16(j)] = __ldcg(a.params + po.b1[t] + j);
            smem[sp.b2[t] + perm16(j)] = __ldcg(a.params + po.b2[t] + j);
        }
        for (int i = tid; i < A * kHid; i += kGruThreads)
            smem[sp.wo + i] = __ldcg(a.params + po.wo + i);
        for (int i = tid; i < kHid; i += kGruThreads)
            smem[sp.wv + i] = __ldcg(a.params + po.wv + i);
        if (tid < A)
        {
            smem[sp.bo + tid] = __ldcg(a.params + po.bo + tid);
            if (a.head == 1)
            {
                const float sl = __ldcg(a.params + po.slog + tid);
                const float sigma = expf(sl);
                smem[sp.slog + tid] = sl;
                SIG[2 * tid] = sigma * sigma;
                SIG[2 * tid + 1] = logf(sigma);
            }
        }
        if (tid == 0)
            smem[sp.bv] = __ldcg(a.params + po.bv);
        for (int i = tid; i < kGruGates; i += kGruThreads)
        {
            BIH[i] = __ldcg(a.params + po.bih + i);
            BHH[i] = __ldcg(a.params + po.bhh + i);
        }
        for (int i = tid; i < gl.total; i += kGruThreads)
            GS[i] = 0.f;
        // the sequence's observations, masks and run boundaries
        for (int i = tid; i < T * Dp; i += kGruThreads)
        {
            const int t = i / Dp, d = i - t * Dp;
            float x = 0.f;
            if (d < D)
            {
                x = __ldg(a.obs + (static_cast<size_t>(t) * N + agent) * D + d);
                if (a.norm_mean)
                {
                    // ObservationNormalizer::process_observation (observation_normalizer.cpp:69-74)
                    const float clip = __ldg(a.norm_clip);
                    x = (x - __ldg(a.norm_mean + d)) / sqrtf(__ldg(a.norm_var + d) + 1e-8f);
                    x = fminf(fmaxf(x, -clip), clip);
                }
            }
            XSEQ[i] = x;
        }
        for (int t = tid; t < T; t += kGruThreads)
        {
            MSEQ[t] = __ldg(a.masks + static_cast<size_t>(t) * N + agent);
            int cut = 0;
            if (t > 0)
            {
                if (a.cut_over_all_agents)
                    for (int n = 0; n < N && !cut; ++n)
                        cut = __ldg(a.masks + static_cast<size_t>(t) * N + n) == 0.f;
                else
                    cut = __ldg(a.masks + static_cast<size_t>(t) * N + agent) == 0.f;
            }
            CUT[t] = cut;
        }
        if (tid == 0)
            CUT[T] = 1;
        const float hx0 = tid < kHid ? __ldg(a.hxs + static_cast<size_t>(agent) * kHid + tid) : 0.f;
        __syncthreads();

        // ---- (1) GRU forward: thread g < 192 owns row g of W_hh (registers) ---------------------------------------
        {
            float whh[kHid], wih[kGruMaxObs];
#pragma unroll
            for (int d = 0; d < kGruMaxObs; ++d)
                wih[d] = (tid < kGruGates && d < D) ? __ldcg(a.params + po.wih + tid * D + d) : 0.f;
            if (tid < kGruGates)
            {
#pragma unroll
                for (int k = 0; k < kHid; k += 4)
                {
                    const float4 w = __ldcg(reinterpret_cast<const float4 *>(a.params + po.whh + tid * kHid + k));
                    whh[k] = w.x, whh[k + 1] = w.y, whh[k + 2] = w.z, whh[k + 3] = w.w;
                }
            }
            for (int t = 0; t < T; ++t)
            {
                if (t == 0 || CUT[t])
                {
                    // every run restarts from the ORIGINAL hxs times the run's first mask (nn_base.cpp:98-102)
                    if (tid < kHid)
                        HCUR[tid] = hx0 * MSEQ[t];
                    __syncthreads();
                }
                if (tid < kGruGates)
                {
                    float gi = BIH[tid];
#pragma unroll
                    for (int d = 0; d < kGruMaxObs; d += 4)
                        if (d < Dp)
                        {
                            const float4 x = *reinterpret_cast<const float4 *>(XSEQ + t * Dp + d);
                            gi = fmaf(wih[d], x.x, gi);
                            gi = fmaf(wih[d + 1], x.y, gi);
                            gi = fmaf(wih[d + 2], x.z, gi);
                            gi = fmaf(wih[d + 3], x.w, gi);
                        }
                    float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll
                    for (int k = 0; k < kHid; k += 4)
                    {
                        const float4 h = *reinterpret_cast<const float4 *>(HCUR + k);
                        a0 = fmaf(whh[k], h.x, a0);
                        a1 = fmaf(whh[k + 1], h.y, a1);
                        a2 = fmaf(whh[k + 2], h.z, a2);
                        a3 = fmaf(whh[k + 3], h.w, a3);
                    }
                    const float gh = BHH[tid] + ((a0 + a1) + (a2 + a3));
                    if (tid < 2 * kHid)
                        GATE[tid] = gi + gh; // r, z pre-activations
                    else
                    {
                        GATE[tid] = gi;       // W_in x + b_in
                        GATE[tid + kHid] = gh; // W_hn h + b_hn
                    }
                }
                __syncthreads();
                if (tid < kHid)
                {
                    const float r = sigmoidf_(GATE[tid]), z = sigmoidf_(GATE[kHid + tid]);
                    const float ghn = GATE[3 * kHid + tid];
                    const float n = tanhf(GATE[2 * kHid + tid] + r * ghn);
                    const float hp = HCUR[tid];
                    const float hn = (1.f - z) * n + z * hp;
                    const int o = t * kHid + tid;
                    TH[o] = hn;
                    if (training)
                        TR[o] = r, TZ[o] = z, TN[o] = n, TG[o] = ghn;
                    HCUR[tid] = hn;
                }
                __syncthreads();
            }
            if (!training && a.step_mode && a.out_hxs && tid < kHid)
                a.out_hxs[static_cast<size_t>(agent) * kHid + tid] = HCUR[tid];
        }
        __syncthreads();

        // ---- (2) towers over tiles of 32 steps: forward, loss, backward; dW1 / dW2 accumulate in registers -----------
        float2 accW1[8][4], accW2[8][4];
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j)
                accW1[i][j] = accW2[i][j] = make_float2(0.f, 0.f);
        for (int t0 = 0; t0 < T; t0 += GR)
        {
            const int nrows = min(GR, T - t0);
            for (int i = tid; i < GR * kHid; i += kGruThreads)
            {
                const int r = i >> 6, k = i & 63;
                X[k * GLD + r] = r < nrows ? __ldcg(TH + (t0 + r) * kHid + k) : 0.f;
            }
            if (tid < GR)
            {
                const int r = tid;
                const bool valid = r < nrows;
                const size_t row = static_cast<size_t>(t0 + (valid ? r : 0)) * N + agent;
                if (training)
                {
                    const float ret = __ldg(a.returns + row);
                    RET[r] = ret;
                    if (ppo)
                    {
                        OLDLP[r] = __ldg(a.old_logp + row);
                        ADV[r] = (ret - __ldg(a.value_preds + row) - adv_mean) / adv_denom;
                    }
                }
                if (a.head == 0)
                    reinterpret_cast<int *>(ACT)[r] = static_cast<int>(__ldg(static_cast<const long long *>(a.actions) + row));
                else
                    for (int i = 0; i < A; ++i)
                        ACT[i * GLD + r] = __ldg(static_cast<const float *>(a.actions) + row * A + i);
            }
            __syncthreads();
            // P1 / P2: both layers of both towers
            {
                float acc[4][4];
                gemm_cf<GR>(w1t_img, X, kHid, jb, rb, acc);
                store_tanh<GR>(h1t, b1_img, jb, rb, acc);
            }
            __syncthreads();
            {
                float acc[4][4];
                gemm_cf<GR>(w2t_img, h1t, kHid, jb, rb, acc);
                store_tanh<GR>(h2t, b2_img, jb, rb, acc);
            }
            __syncthreads();
            // P3: heads
            for (int e = tid; e < (A + 1) * GR; e += kGruThreads)
            {
                const int o = e / GR, r = e - o * GR;
                const bool is_value = o == A;
                const float *w = is_value ? wv_img : wo_img + o * kHid;
                const float *h = smem + (is_value ? sp.h2[1] : sp.h2[0]) + r;
                float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
#pragma unroll 4
                for (int k = 0; k < kHid; k += 4)
                {
                    const float4 wq = *reinterpret_cast<const float4 *>(w + k);
                    s0 = fmaf(wq.x, h[k * GLD], s0);
                    s1 = fmaf(wq.y, h[(k + 1) * GLD], s1);
                    s2 = fmaf(wq.z, h[(k + 2) * GLD], s2);
                    s3 = fmaf(wq.w, h[(k + 3) * GLD], s3);
                }
                const float s = (s0 + s1) + (s2 + s3) + (is_value ? bv_img[0] : bo_img[o]);
                if (is_value)
                    VAL[r] = s;
                else
                    LOGIT[o * GLD + r] = s;
            }
            __syncthreads();
            // P4: log-prob, entropy, loss, gradient w.r.t. the head outputs (four lanes per row)
            if (tid < 4 * GR)
            {
                const int r = tid >> 2, q = tid & 3;
                const bool valid = r < nrows;
                const bool live = training && valid;
                float newlp = 0.f, ent = 0.f;
                const float value = VAL[r];
                int act = 0;
                if (a.head == 0)
                {
                    act = reinterpret_cast<const int *>(ACT)[r];
                    float m = -INFINITY;
                    for (int i = q; i < A; i += 4)
                        m = fmaxf(m, LOGIT[i * GLD + r]);
                    m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 1));
                    m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 2));
                    float se = 0.f;
                    for (int i = q; i < A; i += 4)
                        se += expf(LOGIT[i * GLD + r] - m);
                    se += __shfl_xor_sync(0xffffffffu, se, 1);
                    se += __shfl_xor_sync(0xffffffffu, se, 2);
                    const float lse = m + logf(se);
                    for (int i = q; i < A; i += 4)
                    {
                        const float l = LOGIT[i * GLD + r] - lse;
                        const float pr = expf(l);
                        ent -= l * pr;
                        newlp += i == act ? l : 0.f;
                        LOGIT[i * GLD + r] = l;
                        DLOGIT[i * GLD + r] = pr;
                    }
                }
                else
                {
                    for (int i = q; i < A; i += 4)
                    {
                        const float var = SIG[2 * i], log_sigma = SIG[2 * i + 1];
                        const float diff = ACT[i * GLD + r] - LOGIT[i * GLD + r];
                        newlp += -(diff * diff) / (2.f * var) - log_sigma - 0.91893853320467274178f;
                        ent += 0.5f + 0.5f * 1.8378770664093454836f + log_sigma;
                        LOGIT[i * GLD + r] = diff / var;
                        STAT[(5 + i) * GLD + r] = diff * diff / var - 1.f;
                    }
                }
                newlp += __shfl_xor_sync(0xffffffffu, newlp, 1);
                newlp += __shfl_xor_sync(0xffffffffu, newlp, 2);
                ent += __shfl_xor_sync(0xffffffffu, ent, 1);
                ent += __shfl_xor_sync(0xffffffffu, ent, 2);
                float g_lp = 0.f, dval = 0.f, st_v = 0.f, st_a = 0.f, st_kl = 0.f, st_clip = 0.f;
                if (live && !ppo)
                {
                    const float adv = RET[r] - value;                        // a2c.cpp:69
                    st_v = adv * adv;                                        // a2c.cpp:72
                    st_a = adv * newlp;                                      // a2c.cpp:75
                    g_lp = -inv_b * adv;
                    dval = -2.f * a.c_value * inv_b * adv;
                }
                else if (live)
                {
                    const float olp = OLDLP[r], adv = ADV[r];
                    st_kl = olp - newlp;                                     // ppo.cpp:96
                    const float ratio = expf(newlp - olp);                   // ppo.cpp:107
                    const float s1 = ratio * adv;
                    const float s2 = fminf(fmaxf(ratio, a.clip_lo), a.clip_hi) * adv;
                    st_clip = fabsf(ratio - 1.f) > a.clip_param ? 1.f : 0.f; // ppo.cpp:116-122
                    st_a = fminf(s1, s2);
                    const bool in_range = ratio >= a.clip_lo && ratio <= a.clip_hi;
                    g_lp = (in_range || s1 < s2) ? -a.c_actor * inv_b * adv * ratio : 0.f;
                    const float dv = RET[r] - value;
                    st_v = dv * dv;
                    dval = -a.c_value * inv_b * dv;
                }
                if (a.head == 0)
                {
                    for (int i = q; i < A; i += 4)
                    {
                        const float l = LOGIT[i * GLD + r], pr = DLOGIT[i * GLD + r];
                        const float d = g_lp * ((i == act ? 1.f : 0.f) - pr) + ce_b * pr * (l + ent);
                        DLOGIT[i * GLD + r] = live ? d : 0.f;
                    }
                }
                else
                {
                    for (int i = q; i < A; i += 4)
                    {
                        DLOGIT[i * GLD + r] = live ? g_lp * LOGIT[i * GLD + r] : 0.f;
                        STAT[(5 + i) * GLD + r] = live ? g_lp * STAT[(5 + i) * GLD + r] - ce_b : 0.f;
                    }
                }
                if (q == 0)
                {
                    DVAL[r] = dval;
                    STAT[0 * GLD + r] = st_v;
                    STAT[1 * GLD + r] = st_a;
                    STAT[2 * GLD + r] = live ? ent : 0.f;
                    STAT[3 * GLD + r] = st_kl;
                    STAT[4 * GLD + r] = st_clip;
                }
                if (!training)
                {
                    if (valid && q == 0)
                    {
                        const size_t row = static_cast<size_t>(t0 + r) * N + agent;
                        a.out_value[row] = value;
                        if (a.out_logp)
                            a.out_logp[row] = newlp;
                    }
                    const float e = warp_sum(valid && q == 0 ? ent : 0.f);
                    if (lane == 0 && a.out_entropy_sum)
                        atomicAdd(a.out_entropy_sum, e);
                }
            }
            __syncthreads();
            if (!training)
                continue;
            // P5a: per-row terms -> loss sums, head biases, scale_log (one item per warp, one owner per destination)
            {
                const int n_items = 5 + A + 1 + (a.head == 1 ? A : 0);
                for (int item = warp; item < n_items; item += kGruThreads / 32)
                {
                    const float *src;
                    int dst;
                    if (item < 5)
                        src = STAT + item * GLD, dst = gl.stat + item;
                    else if (item < 5 + A)
                        src = DLOGIT + (item - 5) * GLD, dst = gl.bo + (item - 5);
                    else if (item == 5 + A)
                        src = DVAL, dst = gl.bv;
                    else
                        src = STAT + (5 + item - (6 + A)) * GLD, dst = gl.slog + (item - (6 + A));
                    const float sum = warp_sum(src[lane]);
                    if (lane == 0)
                        GS[dst] += sum;
                }
            }
            // P5b: head weight gradients and dz2 = (head^T dout) * (1 - h2^2), in place over H2
            {
                constexpr int NQ = GR / 8;
                const int k = u & 63, half = u >> 6;
                const int nout = tower == 0 ? A : 1;
                const float *dout = (tower == 0 ? DLOGIT : DVAL) + half * (GR / 2);
                const float *wh = (tower == 0 ? wo_img : wv_img) + k;
                float *h2 = h2t + k * GLD + half * (GR / 2);
                const int gbase = (tower == 0 ? gl.wo : gl.wv) + k;
                float4 h[NQ], dh[NQ];
#pragma unroll
                for (int qd = 0; qd < NQ; ++qd)
                {
                    h[qd] = *reinterpret_cast<const float4 *>(h2 + 4 * qd);
                    dh[qd] = make_float4(0.f, 0.f, 0.f, 0.f);
                }
                for (int o = 0; o < nout; ++o)
                {
                    const float wk = wh[o * kHid];
                    float g0 = 0.f, g1 = 0.f;
#pragma unroll
                    for (int qd = 0; qd < NQ; ++qd)
                    {
                        const float4 d = *reinterpret_cast<const float4 *>(dout + o * GLD + 4 * qd);
                        g0 = fmaf(d.x, h[qd].x, g0);
                        g1 = fmaf(d.y, h[qd].y, g1);
                        g0 = fmaf(d.z, h[qd].z, g0);
                        g1 = fmaf(d.w, h[qd].w, g1);
                        dh[qd].x = fmaf(d.x, wk, dh[qd].x);
                        dh[qd].y = fmaf(d.y, wk, dh[qd].y);
                        dh[qd].z = fmaf(d.z, wk, dh[qd].z);
                        dh[qd].w = fmaf(d.w, wk, dh[qd].w);
                    }
                    atomicAdd(&GS[gbase + o * kHid], g0 + g1);
                }
                float bsum = 0.f;
#pragma unroll
                for (int qd = 0; qd < NQ; ++qd)
                {
                    float4 dz;
                    dz.x = dh[qd].x * (1.f - h[qd].x * h[qd].x);
                    dz.y = dh[qd].y * (1.f - h[qd].y * h[qd].y);
                    dz.z = dh[qd].z * (1.f - h[qd].z * h[qd].z);
                    dz.w = dh[qd].w * (1.f - h[qd].w * h[qd].w);
                    bsum += (dz.x + dz.y) + (dz.z + dz.w);
                    *reinterpret_cast<float4 *>(h2 + 4 * qd) = dz;
                }
                atomicAdd(&GS[gl.b2[tower] + k], bsum);
            }
            __syncthreads();
            // P6: dW2 += dz2 h1^T
            grad_w2_acc<GR>(h2t, h1t, u, accW2);
            __syncthreads();
            // P7: dz1 = (W2^T dz2) * (1 - h1^2), in place over H1; db1
            {
                float acc[4][4];
                gemm_cf<GR>(w2n_img, h2t, kHid, jb, rb, acc);
#pragma unroll
                for (int q = 0; q < 4; ++q)
                {
                    float *hp = h1t + (jb + 16 * q) * GLD + 4 * rb;
                    const float4 h = *reinterpret_cast<const float4 *>(hp);
                    float4 dz;
                    dz.x = acc[q][0] * (1.f - h.x * h.x);
                    dz.y = acc[q][1] * (1.f - h.y * h.y);
                    dz.z = acc[q][2] * (1.f - h.z * h.z);
                    dz.w = acc[q][3] * (1.f - h.w * h.w);
                    *reinterpret_cast<float4 *>(hp) = dz;
                    atomicAdd(&GS[gl.b1[tower] + jb + 16 * q], (dz.x + dz.y) + (dz.z + dz.w));
                }
            }
            __syncthreads();
            // P8: dW1 += dz1 h_gru^T;  P9: dL/dh_gru = W1^T dz1, summed over the two towers, to the tape
            grad_w2_acc<GR>(h1t, X, u, accW1);
            {
                float acc[4][4];
                gemm_cf<GR>(w1n_img, h1t, kHid, jb, rb, acc);
#pragma unroll
                for (int q = 0; q < 4; ++q)
                    *reinterpret_cast<float4 *>(h2t + (jb + 16 * q) * GLD + 4 * rb) =
                        make_float4(acc[q][0], acc[q][1], acc[q][2], acc[q][3]);
            }
            __syncthreads();
            for (int i = tid; i < GR * kHid; i += kGruThreads)
            {
                const int r = i >> 6, k = i & 63;
                if (r < nrows)
                    TDX[(t0 + r) * kHid + k] = smem[sp.h2[0] + k * GLD + r] + smem[sp.h2[1] + k * GLD + r];
            }
            __syncthreads();
        } // tiles
        if (!training)
            break;

        // ---- (3) KL early stop (ppo.cpp:100-104), statistics, Adam on the tower / head parameters -----------------------
        const float sum_v = GS[gl.stat + 0], sum_a = GS[gl.stat + 1], sum_e = GS[gl.stat + 2], sum_kl = GS[gl.stat + 3],
                    sum_clip = GS[gl.stat + 4];
        const float kl = sum_kl * inv_b;
        const bool stop = ppo && static_cast<double>(kl) > a.kl_limit;
        if (tid == 0 && !ppo)
        {
            UpdateState *st = a.state;
            atomicAdd(&st->a2c_value_loss, sum_v * inv_b);
            atomicAdd(&st->a2c_action_loss, -(sum_a * inv_b));
            atomicAdd(&st->a2c_entropy, sum_e * inv_b);
            if (blockIdx.x == 0)
                st->num_updates += 1;
        }
        if (tid == 0 && ppo)
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
            return;
        // bias corrections as host-style doubles (libtorch Adam), t = steps applied so far + 1
        const double tstep = static_cast<double>(opt_step0 + applied + 1);
        const float step_size = static_cast<float>(a.lr / (1.0 - pow(static_cast<double>(a.beta1), tstep)));
        const float bc2_sqrt = static_cast<float>(sqrt(1.0 - pow(static_cast<double>(a.beta2), tstep)));
        auto adam = [&](int gi, float g)
        {
            if (!ppo)
            {
                atomicAdd(a.grad_out + gi, g);
                return;
            }
            const float m = __ldcg(a.opt_m + gi) * a.beta1 + g * (1.f - a.beta1);
            const float v = __ldcg(a.opt_v + gi) * a.beta2 + (g * g) * (1.f - a.beta2);
            a.opt_m[gi] = m;
            a.opt_v[gi] = v;
            a.params[gi] = __ldcg(a.params + gi) - step_size * (m / (sqrtf(v) / bc2_sqrt + a.eps));
        };
        auto adam_range = [&](int gs_off, int p_off, int n)
        {
            for (int i = tid; i < n; i += kGruThreads)
                adam(p_off + i, GS[gs_off + i]);
        };
        {
            const int kq = u & 15, jq = u >> 4;
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j)
                {
                    const int e = (jq + 8 * i) * kHid + kq + 16 * j;
                    adam(po.w1[tower] + e, accW1[i][j].x + accW1[i][j].y);
                    adam(po.w2[tower] + e, accW2[i][j].x + accW2[i][j].y);
                }
        }
        for (int t = 0; t < 2; ++t)
        {
            adam_range(gl.b1[t], po.b1[t], kHid);
            adam_range(gl.b2[t], po.b2[t], kHid);
        }
        adam_range(gl.wv, po.wv, kHid);
        adam_range(gl.bv, po.bv, 1);
        adam_range(gl.wo, po.wo, A * kHid);
        adam_range(gl.bo, po.bo, A);
        if (a.head == 1)
            adam_range(gl.slog, po.slog, A);

        // ---- (4) back-propagation through time ------------------------------------------------------------------------
        // phase A (threads k < 64): gate derivatives of step t from the tape; phase B (all threads, k = tid / 4, p = tid % 4):
        // dh_{t-1}[k] = dh_t[k] z[k] + sum_g W_hh[g][k] dgh[g] with gates [48p, 48p + 48) per thread (two shuffles),
        // dW_hh[g][k] += dgh[g] h_{t-1}[k] in registers.
        for (int i = tid; i < kGruGates * Dp; i += kGruThreads)
            GWIH[i] = 0.f; // aliases the tower tiles, dead from here on
        {
            const int kk = tid >> 2, pp = tid & 3;
            float wcol[48], gw[48];
#pragma unroll
            for (int i = 0; i < 48; ++i)
            {
                wcol[i] = __ldcg(a.params + po.whh + (pp * 48 + i) * kHid + kk);
                gw[i] = 0.f;
            }
            float gbi[3] = {0.f, 0.f, 0.f}, gbh[3] = {0.f, 0.f, 0.f};
            // software pipeline: the tape values of step t - 1 are in flight while step t computes
            float c_r = 0.f, c_z = 0.f, c_n = 0.f, c_g = 0.f, c_dx = 0.f, c_hp = 0.f;
            auto load_step = [&](int t, float &r, float &z, float &n, float &g, float &dx, float &hp)
            {
                const int o = t * kHid + tid;
                r = __ldcg(TR + o), z = __ldcg(TZ + o), n = __ldcg(TN + o), g = __ldcg(TG + o), dx = __ldcg(TDX + o);
                hp = (t == 0 || CUT[t]) ? hx0 * MSEQ[t] : __ldcg(TH + o - kHid);
            };
            if (tid < kHid)
            {
                load_step(T - 1, c_r, c_z, c_n, c_g, c_dx, c_hp);
                DHN[tid] = 0.f; // carry into step T-1
            }
            __syncthreads();
            for (int t = T - 1; t >= 0; --t)
            {
                float n_r = 0.f, n_z = 0.f, n_n = 0.f, n_g = 0.f, n_dx = 0.f, n_hp = 0.f;
                if (tid < kHid)
                {
                    if (t > 0)
                        load_step(t - 1, n_r, n_z, n_n, n_g, n_dx, n_hp);
                    const float dh = c_dx + DHN[tid];
                    const float dn = dh * (1.f - c_z);
                    const float dz = dh * (c_hp - c_n);
                    const float dpn = dn * (1.f - c_n * c_n);
                    const float dpr = dpn * c_g * c_r * (1.f - c_r);
                    const float dpz = dz * c_z * (1.f - c_z);
                    DGI[tid] = dpr, DGI[kHid + tid] = dpz, DGI[2 * kHid + tid] = dpn;
                    DGH[tid] = dpr, DGH[kHid + tid] = dpz, DGH[2 * kHid + tid] = dpn * c_r;
                    DHZ[tid] = dh * c_z;
                    HPREV[tid] = c_hp;
                    gbi[0] += dpr, gbi[1] += dpz, gbi[2] += dpn;
                    gbh[0] += dpr, gbh[1] += dpz, gbh[2] += dpn * c_r;
                }
                __syncthreads();
                {
                    const float hp = HPREV[kk];
                    float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
#pragma unroll
                    for (int i = 0; i < 48; i += 4)
                    {
                        const float4 d = *reinterpret_cast<const float4 *>(DGH + pp * 48 + i);
                        s0 = fmaf(wcol[i], d.x, s0);
                        s1 = fmaf(wcol[i + 1], d.y, s1);
                        s2 = fmaf(wcol[i + 2], d.z, s2);
                        s3 = fmaf(wcol[i + 3], d.w, s3);
                        gw[i] = fmaf(d.x, hp, gw[i]);
                        gw[i + 1] = fmaf(d.y, hp, gw[i + 1]);
                        gw[i + 2] = fmaf(d.z, hp, gw[i + 2]);
                        gw[i + 3] = fmaf(d.w, hp, gw[i + 3]);
                    }
                    float s = (s0 + s1) + (s2 + s3);
                    s += __shfl_xor_sync(0xffffffffu, s, 1);
                    s += __shfl_xor_sync(0xffffffffu, s, 2);
                    // a run that starts at step t takes its state from hxs, not from step t-1: no gradient crosses
                    const float carry = (t == 0 || CUT[t]) ? 0.f : DHZ[kk] + s;
                    for (int i = tid; i < kGruGates * Dp; i += kGruThreads)
                    {
                        const int g = i / Dp, d = i - g * Dp;
                        GWIH[i] = fmaf(DGI[g], XSEQ[t * Dp + d], GWIH[i]);
                    }
                    if (pp == 0)
                        DHN[kk] = carry; // read by phase A of step t-1, after the barrier below
                }
                c_r = n_r, c_z = n_z, c_n = n_n, c_g = n_g, c_dx = n_dx, c_hp = n_hp;
                __syncthreads();
            }
            // Adam on the GRU parameters
#pragma unroll
            for (int i = 0; i < 48; ++i)
                adam(po.whh + (pp * 48 + i) * kHid + kk, gw[i]);
            if (tid < kHid)
            {
#pragma unroll
                for (int g = 0; g < 3; ++g)
                {
                    adam(po.bih + g * kHid + tid, gbi[g]);
                    adam(po.bhh + g * kHid + tid, gbh[g]);
                }
            }
            for (int i = tid; i < kGruGates * D; i += kGruThreads)
            {
                const int g = i / D, d = i - g * D;
                adam(po.wih + i, GWIH[g * Dp + d]);
            }
        }
        ++applied;
        __syncthreads(); // parameter stores of this minibatch are visible to the whole CTA before the next image build
    } // minibatches
}
} // namespace

size_t gru_update_smem_bytes(int Dp, int A, int Ap, int T) { return static_cast<size_t>(make_gru_smem(Dp, A, Ap, T).total) * sizeof(float); }

int launch_gru_update(const GruUpdateArgs &args, int grid, cudaStream_t s)
{
    static bool configured = false;
    if (!configured)
    {
        CPPRL_CUDA_TRY(cudaFuncSetAttribute(gru_update_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        configured = true;
    }
    const size_t smem = gru_update_smem_bytes(args.Dp, args.A, args.Ap, args.T);
    gru_update_kernel<<<grid, kGruThreads, smem, s>>>(args);
    CPPRL_CUDA_TRY(cudaGetLastError());
    count_launch();
    return 0;
}
} // namespace cpprl_b200
