// gru_update.cu — fused PPO / A2C update (and forward evaluation) for the GRU-fronted MlpBase.
//
// Replaces, for recurrent policies:
//   RecurrentGenerator::next       src/generators/recurrent_generator.cpp:47-98   (one environment per minibatch)
//   NNBase::forward_gru            src/model/nn_base.cpp:47-114   (run splitting at mask zeros, state never carried
//                                                                  across runs: every run restarts from hxs * mask)
//   MlpBase::forward, heads, log-prob / entropy, PPO / A2C loss, autograd backward (BPTT), Adam step, KL stop,
//   statistics                     src/model/mlp_base.cpp:47-64, src/algorithms/ppo.cpp:86-146, a2c.cpp:58-85
//
// The reference's recurrent PPO is a strictly serial chain: N minibatches per epoch, each ONE environment's T steps,
// each followed by an optimiser step.  One CTA therefore walks the whole epoch (no grid barrier, no partial-gradient
// exchange) with every tower / head / GRU-input parameter RESIDENT in shared memory for the whole launch: Adam
// writes the new value to the shared-memory images and to the flat parameter buffer at the same time.  Per minibatch:
//   1. GRU forward over the T steps (W_hh rows in registers, one matrix-vector product per step); h, r, z, n and
//      W_hn h + b_hn go to a tape in L2;
//   2. the T outputs run through both towers, heads, loss and the hand-derived backward in tiles of 32 rows out of
//      shared memory (the building blocks of mlp_update.cu); the 64x64 weight gradients accumulate in registers,
//      dL/dh_t goes to the tape;
//   3. KL stop (ppo.cpp:100-104), Adam on the tower / head parameters straight from the registers;
//   4. back-propagation through time: the tape streams back through shared memory in chunks of 8 steps (cp.async,
//      double buffered), W_hh columns and dW_hh live in registers; Adam on the GRU parameters.
// A2C and forward evaluation (Policy::evaluate_actions / get_values) launch one CTA per environment; A2C sums the
// per-environment gradients in global memory and the stand-alone RMSprop kernel applies them.
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
constexpr int kMat = kHid * kHid;
// per-tower weight image: W1 forward / backward layouts, W2 forward / backward layouts, permuted biases
constexpr int kW1T = 0, kW1N = kMat, kW2T = 2 * kMat, kW2N = 3 * kMat, kB1 = 4 * kMat, kB2 = 4 * kMat + kHid;
constexpr int kTowerImg = 4 * kMat + 2 * kHid;
// "tile" region: tower tiles X, H1[2], H2[2] (5 x 64 x GLD floats) during phase 2; during BPTT two staging buffers
// of kChunk steps x 6 tape rows and the dW_ih accumulator
constexpr int kTileMat = kHid * GLD;
constexpr int kChunk = 8;
constexpr int kChunkFloats = 6 * kChunk * kHid;
constexpr int kTileFloats = 2 * kChunkFloats + kGruGates * kGruMaxObs; // 12288 >= 5 * kTileMat = 11520
static_assert(kTileFloats >= 5 * kTileMat, "tile region");
// small-gradient vector: biases of the towers, then a mirror of the parameter tail [wv, bv, (scale_log), wo, bo]
constexpr int kGsBias = 0, kGsTail = 4 * kHid;

// Padded shared-memory layouts of the per-step vectors: the four lanes of a quad read four different slices with
// 16-byte loads, so consecutive slices start 4 banks past a multiple of 16 (conflict-free)
constexpr int kHcStride = 80;   // hidden state: slices of 16 floats, padded to 20
constexpr int kDghStride = 208; // dL/d(W_hh h + b_hh): slices of 48 floats, padded to 52
__device__ __forceinline__ int hc_pos(int j) { return j + 4 * (j >> 4); }
__device__ __forceinline__ int dgh_pos(int g) { return g + 4 * (g / 48); }

struct GruSmem
{
    int tw, wo, bo, wv, bv, sig;                                        // resident weights
    int wiht, bih, bhh;                                                 // resident GRU input weights [d][gate], biases
    int tile, logit, dlogit, val, dval, ret, oldlp, adv, act, stat, gs; // tower tile, small gradients + statistics
    int gs_stat, tail_len;
    int hcur, dgh, xseq, mseq, cut, total;                              // GRU step state
};

__host__ __device__ inline GruSmem make_gru_smem(int Dp, int A, int Ap, int head, int T)
{
    GruSmem s;
    int o = 0;
    s.tw = o, o += 2 * kTowerImg;
    s.wo = o, o += Ap * kHid;
    s.bo = o, o += Ap;
    s.wv = o, o += kHid;
    s.bv = o, o += 4;
    s.sig = o, o += 2 * Ap;
    s.wiht = o, o += kGruGates * Dp;
    s.bih = o, o += kGruGates;
    s.bhh = o, o += kGruGates;
    s.tile = o, o += kTileFloats;
    s.logit = o, o += Ap * GLD;
    s.dlogit = o, o += Ap * GLD;
    s.val = o, o += GLD;
    s.dval = o, o += GLD;
    s.ret = o, o += GLD;
    s.oldlp = o, o += GLD;
    s.adv = o, o += GLD;
    s.act = o, o += Ap * GLD;
    s.stat = o, o += (5 + Ap) * GLD;
    s.tail_len = kHid + 1 + (head == 1 ? A : 0) + A * kHid + A;
    s.gs = o;
    s.gs_stat = kGsTail + ((s.tail_len + 3) & ~3);
    o += s.gs_stat + 8;
    s.hcur = o, o += 2 * kHcStride; // double buffered state
    s.dgh = o, o += 2 * kDghStride; // double buffered dL/d(W_hh h + b_hh)
    s.xseq = o, o += ((T * Dp) + 3) & ~3;
    s.mseq = o, o += (T + 3) & ~3;
    s.cut = o, o += (T + 4) & ~3;
    s.total = o;
    return s;
}

// Gate non-linearities of the serial recurrence: every instruction here is on the per-step latency chain, so the
// exponential is the two-instruction ex2 form and the quotient the approximate division (2 ulp); the absolute error
// stays below 2e-7, the level of the correctly rounded functions' own rounding.
__device__ __forceinline__ float sigmoidf_(float x) { return __fdividef(1.f, 1.f + __expf(-x)); }
__device__ __forceinline__ float tanhf_(float x)
{
    const float e = __expf(2.f * fminf(fmaxf(x, -15.f), 15.f)); // tanh(+-15) = +-1 in fp32
    return 1.f - __fdividef(2.f, e + 1.f);
}

__device__ __forceinline__ void cp_async16(float *dst_smem, const float *src)
{
    const unsigned d = static_cast<unsigned>(__cvta_generic_to_shared(dst_smem));
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

__global__ void __launch_bounds__(kGruThreads, 1) gru_update_kernel(const __grid_constant__ GruUpdateArgs a)
{
    extern __shared__ __align__(16) float smem[];
    const GruSmem sp = make_gru_smem(a.Dp, a.A, a.Ap, a.head, a.T);
    const GruParamOffsets &po = a.po;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tower = tid >> 7, u = tid & 127, jb = u & 15, rb = u >> 4;
    const int D = a.D, Dp = a.Dp, A = a.A, T = a.T, N = a.N;
    const bool training = a.algo != kAlgoForward;
    // PPO: one environment per minibatch, T rows.  A2C: one batch of all T * N rows, one CTA per environment, the
    // gradient summed over the CTAs in global memory (grad_out) and applied by the RMSprop kernel that follows.
    const bool ppo = a.algo == kAlgoPPO;
    float *GS = smem + sp.gs;
    float *HCUR = smem + sp.hcur, *DGH = smem + sp.dgh;
    float *XSEQ = smem + sp.xseq, *MSEQ = smem + sp.mseq;
    int *CUT = reinterpret_cast<int *>(smem + sp.cut);
    float *WIHT = smem + sp.wiht, *BIH = smem + sp.bih, *BHH = smem + sp.bhh;
    float *TILE = smem + sp.tile;
    float *X = TILE;
    float *h1t = TILE + (1 + tower) * kTileMat, *h2t = TILE + (3 + tower) * kTileMat;
    float *LOGIT = smem + sp.logit, *DLOGIT = smem + sp.dlogit, *VAL = smem + sp.val, *DVAL = smem + sp.dval;
    float *RET = smem + sp.ret, *OLDLP = smem + sp.oldlp, *ADV = smem + sp.adv, *ACT = smem + sp.act, *STAT = smem + sp.stat;
    float *SIG = smem + sp.sig;
    float *tw_img = smem + sp.tw + tower * kTowerImg; // this thread's tower
    float *wo_img = smem + sp.wo, *bo_img = smem + sp.bo, *wv_img = smem + sp.wv, *bv_img = smem + sp.bv;
    // small-gradient vector: offsets of the tail mirror [wv, bv, (scale_log), wo, bo]
    const int gt_wv = kGsTail, gt_bv = kGsTail + kHid, gt_slog = gt_bv + 1, gt_wo = gt_slog + (a.head == 1 ? A : 0),
              gt_bo = gt_wo + A * kHid;
    // forward evaluation keeps only the hidden states (one tape per CTA), training all six
    float *tape = a.tape + static_cast<size_t>(blockIdx.x) * (training ? kGruTapes : 1) * T * kHid;
    // tape rows: h [T][64]; {r, z, n, W_hn h + b_hn} interleaved [T][64][4] (one 16-byte store per step and unit);
    // dL/dh from the towers [T][64]
    float *TH = tape, *TRZNG = tape + T * kHid, *TDX = tape + 5 * T * kHid;

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
    const float inv_b = 1.f / static_cast<float>(ppo ? T : T * N);
    const float ce_b = a.c_entropy * inv_b;
    const int n_seq = ppo ? a.n_mb : 1;

    // ---- (0) once per launch: weight images from the flat parameters ---------------------------------------------
    for (int i = tid; i < 2 * 2 * kMat; i += kGruThreads)
    {
        const int t = i >> 13, layer = (i >> 12) & 1, e = i & 4095, j = e >> 6, k = e & 63;
        const float v = a.params[(layer ? po.w2[t] : po.w1[t]) + e];
        float *img = smem + sp.tw + t * kTowerImg;
        img[(layer ? kW2T : kW1T) + k * kHid + perm16(j)] = v; // forward:  out[j] += W[j][k] in[k]
        img[(layer ? kW2N : kW1N) + j * kHid + perm16(k)] = v; // backward: out[k] += W[j][k] dz[j]
    }
    for (int i = tid; i < 2 * kHid; i += kGruThreads)
    {
        const int t = i >> 6, j = i & 63;
        float *img = smem + sp.tw + t * kTowerImg;
        img[kB1 + perm16(j)] = a.params[po.b1[t] + j];
        img[kB2 + perm16(j)] = a.params[po.b2[t] + j];
    }
    for (int i = tid; i < A * kHid; i += kGruThreads)
        wo_img[i] = a.params[po.wo + i];
    for (int i = tid; i < kHid; i += kGruThreads)
        wv_img[i] = a.params[po.wv + i];
    if (tid < A)
    {
        bo_img[tid] = a.params[po.bo + tid];
        if (a.head == 1)
        {
            const float sigma = expf(a.params[po.slog + tid]);
            SIG[2 * tid] = sigma * sigma;
            SIG[2 * tid + 1] = logf(sigma);
        }
    }
    if (tid == 0)
        bv_img[0] = a.params[po.bv];
    for (int i = tid; i < kGruGates * Dp; i += kGruThreads)
    {
        const int d = i / kGruGates, g = i - d * kGruGates;
        WIHT[i] = d < D ? a.params[po.wih + g * D + d] : 0.f; // [d][gate]: conflict-free for gate-per-thread reads
    }
    for (int i = tid; i < kGruGates; i += kGruThreads)
    {
        BIH[i] = a.params[po.bih + i];
        BHH[i] = a.params[po.bhh + i];
    }
    float norm_clip = 0.f;
    if (a.norm_mean)
        norm_clip = __ldg(a.norm_clip);

#define GRU_MARK(slot)                                                  \
    do                                                                  \
    {                                                                   \
        if (a.dbg_clocks && tid == 0 && mb == a.dbg_step)               \
            a.dbg_clocks[blockIdx.x * 32 + (slot)] = clock64();         \
    } while (0)
    for (int mb = 0; mb < n_seq; ++mb)
    {
        GRU_MARK(0);
        const int agent = ppo ? static_cast<int>(a.perm[mb]) : static_cast<int>(blockIdx.x);
        // ---- stage the sequence: observations (normalised), masks, run boundaries; clear the small gradients -------
        for (int i = tid; i < sp.gs_stat + 8; i += kGruThreads)
            GS[i] = 0.f;
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
                    x = (x - __ldg(a.norm_mean + d)) / sqrtf(__ldg(a.norm_var + d) + 1e-8f);
                    x = fminf(fmaxf(x, -norm_clip), norm_clip);
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
        __syncthreads();

        GRU_MARK(1);
        // ---- (1) GRU forward ----------------------------------------------------------------------------------------
        // Thread (k = tid / 4, p = tid % 4): the four lanes of a quad share hidden unit k and split the 64 recurrent
        // inputs (and the observation columns) four ways; W_hh[{r,z,n} x 64 + k][16p .. 16p+15] lives in registers.
        // Per step: partial dot products (FFMA2), two shuffles, the gate math in every lane, ONE barrier (the state
        // is double buffered).
        const int kk = tid >> 2, pp = tid & 3;
        const float hx0 = __ldg(a.hxs + static_cast<size_t>(agent) * kHid + kk);
        {
            float2 wr[8], wz[8], wn[8];
#pragma unroll
            for (int q = 0; q < 4; ++q)
            {
                const float4 r4 = __ldcg(reinterpret_cast<const float4 *>(a.params + po.whh + (kk) * kHid + 16 * pp) + q);
                const float4 z4 = __ldcg(reinterpret_cast<const float4 *>(a.params + po.whh + (kHid + kk) * kHid + 16 * pp) + q);
                const float4 n4 = __ldcg(reinterpret_cast<const float4 *>(a.params + po.whh + (2 * kHid + kk) * kHid + 16 * pp) + q);
                wr[2 * q] = make_float2(r4.x, r4.y), wr[2 * q + 1] = make_float2(r4.z, r4.w);
                wz[2 * q] = make_float2(z4.x, z4.y), wz[2 * q + 1] = make_float2(z4.z, z4.w);
                wn[2 * q] = make_float2(n4.x, n4.y), wn[2 * q + 1] = make_float2(n4.z, n4.w);
            }
            // input-side partial sums of step t for this lane's observation columns d = p, p+4, ...; lane 0 also carries
            // the biases (b_ir + b_hr, b_iz + b_hz, b_in; b_hn goes with the recurrent part of the n gate)
            const float b_r = pp == 0 ? BIH[kk] + BHH[kk] : 0.f, b_z = pp == 0 ? BIH[kHid + kk] + BHH[kHid + kk] : 0.f;
            const float b_in = pp == 0 ? BIH[2 * kHid + kk] : 0.f, b_hn = pp == 0 ? BHH[2 * kHid + kk] : 0.f;
            float gir, giz, gin;
            auto input_partials = [&](int t)
            {
                gir = b_r, giz = b_z, gin = b_in;
                for (int d = pp; d < Dp; d += 4)
                {
                    const float x = XSEQ[t * Dp + d];
                    gir = fmaf(WIHT[d * kGruGates + kk], x, gir);
                    giz = fmaf(WIHT[d * kGruGates + kHid + kk], x, giz);
                    gin = fmaf(WIHT[d * kGruGates + 2 * kHid + kk], x, gin);
                }
            };
            input_partials(0);
            // every run restarts from the ORIGINAL hxs times the run's first mask (nn_base.cpp:98-102)
            if (pp == 0)
                HCUR[hc_pos(kk)] = hx0 * MSEQ[0];
            __syncthreads();
            for (int t = 0; t < T; ++t)
            {
                const float *hc = HCUR + (t & 1) * kHcStride;
                const float hp = hc[hc_pos(kk)];
                float2 sr = make_float2(gir, 0.f), sz = make_float2(giz, 0.f), sn = make_float2(b_hn, 0.f);
#pragma unroll
                for (int q = 0; q < 4; ++q)
                {
                    const float4 h = *reinterpret_cast<const float4 *>(hc + 20 * pp + 4 * q);
                    const float2 lo = make_float2(h.x, h.y), hi = make_float2(h.z, h.w);
                    sr = __ffma2_rn(wr[2 * q], lo, sr), sr = __ffma2_rn(wr[2 * q + 1], hi, sr);
                    sz = __ffma2_rn(wz[2 * q], lo, sz), sz = __ffma2_rn(wz[2 * q + 1], hi, sz);
                    sn = __ffma2_rn(wn[2 * q], lo, sn), sn = __ffma2_rn(wn[2 * q + 1], hi, sn);
                }
                float pr = sr.x + sr.y, pz = sz.x + sz.y, pn = sn.x + sn.y, pi = gin;
                pr += __shfl_xor_sync(0xffffffffu, pr, 1), pz += __shfl_xor_sync(0xffffffffu, pz, 1);
                pn += __shfl_xor_sync(0xffffffffu, pn, 1), pi += __shfl_xor_sync(0xffffffffu, pi, 1);
                pr += __shfl_xor_sync(0xffffffffu, pr, 2), pz += __shfl_xor_sync(0xffffffffu, pz, 2);
                pn += __shfl_xor_sync(0xffffffffu, pn, 2), pi += __shfl_xor_sync(0xffffffffu, pi, 2);
                const float r = sigmoidf_(pr), z = sigmoidf_(pz);
                const float n = tanhf_(pi + r * pn); // n = tanh(W_in x + b_in + r * (W_hn h + b_hn))
                const float hn = (1.f - z) * n + z * hp;
                const int o = t * kHid + kk;
                if (pp == 0)
                {
                    TH[o] = hn;
                    // state seen by step t+1: a run boundary restarts from hxs
                    HCUR[((t + 1) & 1) * kHcStride + hc_pos(kk)] = (t + 1 < T && CUT[t + 1]) ? hx0 * MSEQ[t + 1] : hn;
                }
                else if (training && pp == 1)
                    *reinterpret_cast<float4 *>(TRZNG + 4 * o) = make_float4(r, z, n, pn);
                if (t + 1 < T)
                    input_partials(t + 1);
                __syncthreads();
            }
            if (!training && a.step_mode && a.out_hxs && pp == 0)
                a.out_hxs[static_cast<size_t>(agent) * kHid + kk] = HCUR[(T & 1) * kHcStride + hc_pos(kk)];
        }

        GRU_MARK(2);
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
                gemm_cf<GR>(tw_img + kW1T, X, kHid, jb, rb, acc);
                store_tanh<GR>(h1t, tw_img + kB1, jb, rb, acc);
            }
            __syncthreads();
            {
                float acc[4][4];
                gemm_cf<GR>(tw_img + kW2T, h1t, kHid, jb, rb, acc);
                store_tanh<GR>(h2t, tw_img + kB2, jb, rb, acc);
            }
            __syncthreads();
            // P3: heads
            for (int e = tid; e < (A + 1) * GR; e += kGruThreads)
            {
                const int o = e / GR, r = e - o * GR;
                const bool is_value = o == A;
                const float *w = is_value ? wv_img : wo_img + o * kHid;
                const float *h = TILE + (is_value ? 4 : 3) * kTileMat + r;
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
                        if (!training && a.out_dist && valid)
                            a.out_dist[(static_cast<size_t>(t0 + r) * N + agent) * A + i] = l;
                    }
                }
                else
                {
                    for (int i = q; i < A; i += 4)
                    {
                        const float var = SIG[2 * i], log_sigma = SIG[2 * i + 1];
                        const float diff = ACT[i * GLD + r] - LOGIT[i * GLD + r];
                        if (!training && a.out_dist && valid)
                            a.out_dist[(static_cast<size_t>(t0 + r) * N + agent) * A + i] = LOGIT[i * GLD + r];
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
                        src = STAT + item * GLD, dst = sp.gs_stat + item;
                    else if (item < 5 + A)
                        src = DLOGIT + (item - 5) * GLD, dst = gt_bo + (item - 5);
                    else if (item == 5 + A)
                        src = DVAL, dst = gt_bv;
                    else
                        src = STAT + (5 + item - (6 + A)) * GLD, dst = gt_slog + (item - (6 + A));
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
                const int gbase = (tower == 0 ? gt_wo : gt_wv) + k;
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
                atomicAdd(&GS[kGsBias + tower * 2 * kHid + kHid + k], bsum);
            }
            __syncthreads();
            // P6: dW2 += dz2 h1^T
            grad_w2_acc<GR>(h2t, h1t, u, accW2);
            __syncthreads();
            // P7: dz1 = (W2^T dz2) * (1 - h1^2), in place over H1; db1
            {
                float acc[4][4];
                gemm_cf<GR>(tw_img + kW2N, h2t, kHid, jb, rb, acc);
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
                    atomicAdd(&GS[kGsBias + tower * 2 * kHid + jb + 16 * q], (dz.x + dz.y) + (dz.z + dz.w));
                }
            }
            __syncthreads();
            // P8: dW1 += dz1 h_gru^T;  P9: dL/dh_gru = W1^T dz1, summed over the two towers, to the tape
            grad_w2_acc<GR>(h1t, X, u, accW1);
            {
                float acc[4][4];
                gemm_cf<GR>(tw_img + kW1N, h1t, kHid, jb, rb, acc);
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
                    TDX[(t0 + r) * kHid + k] = TILE[3 * kTileMat + k * GLD + r] + TILE[4 * kTileMat + k * GLD + r];
            }
            __syncthreads();
        } // tiles
        if (!training)
            break;


        GRU_MARK(3);
        // ---- (3) KL early stop (ppo.cpp:100-104), statistics, Adam on the tower / head parameters -----------------------
        const float sum_v = GS[sp.gs_stat + 0], sum_a = GS[sp.gs_stat + 1], sum_e = GS[sp.gs_stat + 2],
                    sum_kl = GS[sp.gs_stat + 3], sum_clip = GS[sp.gs_stat + 4];
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
        const float step_size = ppo ? static_cast<float>(a.lr / (1.0 - pow(static_cast<double>(a.beta1), tstep))) : 0.f;
        const float bc2_sqrt = ppo ? static_cast<float>(sqrt(1.0 - pow(static_cast<double>(a.beta2), tstep))) : 1.f;
        // one Adam step of a parameter whose value the caller already holds: state loaded by the caller in batches so
        // that the L2 round trips overlap; returns the new value
        auto adam_core = [&](int gi, float g, float p, float m_old, float v_old) -> float
        {
            const float m = m_old * a.beta1 + g * (1.f - a.beta1);
            const float v = v_old * a.beta2 + (g * g) * (1.f - a.beta2);
            a.opt_m[gi] = m;
            a.opt_v[gi] = v;
            const float pn = p - step_size * (m / (sqrtf(v) / bc2_sqrt + a.eps));
            a.params[gi] = pn;
            return pn;
        };
        {
            const int kq = u & 15, jq = u >> 4;
            const int p1 = po.w1[tower], p2 = po.w2[tower];
#pragma unroll
            for (int i = 0; i < 8; ++i)
            {
                const int j = jq + 8 * i;
                if (!ppo)
                {
#pragma unroll
                    for (int jj = 0; jj < 4; ++jj)
                    {
                        const int e = j * kHid + kq + 16 * jj;
                        atomicAdd(a.grad_out + p1 + e, accW1[i][jj].x + accW1[i][jj].y);
                        atomicAdd(a.grad_out + p2 + e, accW2[i][jj].x + accW2[i][jj].y);
                    }
                    continue;
                }
                float m1[4], v1[4], m2[4], v2[4];
#pragma unroll
                for (int jj = 0; jj < 4; ++jj)
                {
                    const int e = j * kHid + kq + 16 * jj;
                    m1[jj] = __ldcg(a.opt_m + p1 + e), v1[jj] = __ldcg(a.opt_v + p1 + e);
                    m2[jj] = __ldcg(a.opt_m + p2 + e), v2[jj] = __ldcg(a.opt_v + p2 + e);
                }
#pragma unroll
                for (int jj = 0; jj < 4; ++jj)
                {
                    const int k = kq + 16 * jj, e = j * kHid + k;
                    const int in = j * kHid + perm16(k), it = k * kHid + perm16(j);
                    const float n1 = adam_core(p1 + e, accW1[i][jj].x + accW1[i][jj].y, tw_img[kW1N + in], m1[jj], v1[jj]);
                    tw_img[kW1N + in] = n1, tw_img[kW1T + it] = n1;
                    const float n2 = adam_core(p2 + e, accW2[i][jj].x + accW2[i][jj].y, tw_img[kW2N + in], m2[jj], v2[jj]);
                    tw_img[kW2N + in] = n2, tw_img[kW2T + it] = n2;
                }
            }
        }
        {
            // tower biases: one element per thread
            const int t = tid >> 7, layer = (tid >> 6) & 1, j = tid & 63;
            const int gi = (layer ? po.b2[t] : po.b1[t]) + j;
            float *slot = smem + sp.tw + t * kTowerImg + (layer ? kB2 : kB1) + perm16(j);
            const float g = GS[kGsBias + tid];
            if (ppo)
                *slot = adam_core(gi, g, *slot, __ldcg(a.opt_m + gi), __ldcg(a.opt_v + gi));
            else
                atomicAdd(a.grad_out + gi, g);
        }
        for (int i = tid; i < sp.tail_len; i += kGruThreads)
        {
            // parameter tail [critic_linear.weight, .bias, (scale_log), head weight, head bias]
            const int gi = po.wv + i;
            const float g = GS[kGsTail + i];
            if (!ppo)
            {
                atomicAdd(a.grad_out + gi, g);
                continue;
            }
            const float pn = adam_core(gi, g, __ldcg(a.params + gi), __ldcg(a.opt_m + gi), __ldcg(a.opt_v + gi));
            if (i < kHid)
                wv_img[i] = pn;
            else if (i == kHid)
                bv_img[0] = pn;
            else
            {
                int r = i - (kHid + 1);
                if (a.head == 1 && r < A)
                {
                    const float sigma = expf(pn);
                    SIG[2 * r] = sigma * sigma;
                    SIG[2 * r + 1] = logf(sigma);
                }
                else
                {
                    r -= a.head == 1 ? A : 0;
                    if (r < A * kHid)
                        wo_img[r] = pn;
                    else
                        bo_img[r - A * kHid] = pn;
                }
            }
        }
        __syncthreads(); // the tile region changes hands

        GRU_MARK(4);
        // ---- (4) back-propagation through time ------------------------------------------------------------------------
        // The tape streams back through shared memory, kChunk steps at a time, one chunk ahead of the step loop.
        // Thread (k = tid / 4, p = tid % 4).  Per step every lane of quad k forms the gate derivatives of unit k from
        // the staged tape (dL/dh_t carried in registers), lanes 0-2 publish dgh = dL/d(W_hh h + b_hh) for the three
        // gates; ONE barrier; then dh_{t-1}[k] = dh_t[k] z[k] + sum_g W_hh[g][k] dgh[g] with gates [48p, 48p + 48)
        // per lane (FFMA2, two shuffles) and dW_hh[g][k] += dgh[g] h_{t-1}[k] in registers.  dgi overwrites the
        // consumed tape rows so that dW_ih[d][g] += dgi_t[g] x_t[d] is done once per chunk by all threads.
        {
            float *GWIH = TILE + 2 * kChunkFloats; // [d][gate]
            for (int i = tid; i < kGruGates * Dp; i += kGruThreads)
                GWIH[i] = 0.f;
            float2 wcol[24], gw[24];
#pragma unroll
            for (int i = 0; i < 24; ++i)
            {
                wcol[i].x = __ldcg(a.params + po.whh + (pp * 48 + 2 * i) * kHid + kk);
                wcol[i].y = __ldcg(a.params + po.whh + (pp * 48 + 2 * i + 1) * kHid + kk);
                gw[i] = make_float2(0.f, 0.f);
            }
            float gbi[3] = {0.f, 0.f, 0.f}, gbh[3] = {0.f, 0.f, 0.f};
            const int n_chunks = (T + kChunk - 1) / kChunk;
            auto stage_chunk = [&](int c)
            {
                const int hi = T - c * kChunk, lo = max(0, hi - kChunk);
                float *buf = TILE + (c & 1) * kChunkFloats;
                // buffer: {r,z,n,g} [s][64][4] | dL/dh [s][64] | h_{t-1} [s][64] (unused at t = 0)
                for (int i = tid; i < 6 * kChunk * 16; i += kGruThreads)
                {
                    const float *src;
                    float *dst;
                    int t;
                    if (i < 4 * kChunk * 16)
                    {
                        const int sidx = i >> 6, c4 = i & 63;
                        t = lo + sidx;
                        src = TRZNG + (static_cast<size_t>(t) * kHid + c4) * 4, dst = buf + sidx * 4 * kHid + 4 * c4;
                    }
                    else
                    {
                        const int j = i - 4 * kChunk * 16, which = j >> 7, sidx = (j >> 4) & (kChunk - 1), c4 = j & 15;
                        t = lo + sidx;
                        src = (which ? TH + static_cast<size_t>(t > 0 ? t - 1 : 0) * kHid : TDX + static_cast<size_t>(t) * kHid) + 4 * c4;
                        dst = buf + (4 + which) * kChunk * kHid + sidx * kHid + 4 * c4;
                    }
                    if (t < hi)
                        cp_async16(dst, src);
                }
                cp_async_commit();
            };
            stage_chunk(0);
            float carry = 0.f; // dL/dh_t arriving from step t+1 (identical in the four lanes of a quad)
            for (int c = 0; c < n_chunks; ++c)
            {
                if (c + 1 < n_chunks)
                    stage_chunk(c + 1);
                else
                    cp_async_commit();
                cp_async_wait<1>();
                __syncthreads();
                const int hi = T - c * kChunk, lo = max(0, hi - kChunk);
                float *buf = TILE + (c & 1) * kChunkFloats;
                for (int t = hi - 1; t >= lo; --t)
                {
                    const bool first = t == 0 || CUT[t]; // step t starts a run: its state came from hxs, not from t-1
                    float4 *rzng = reinterpret_cast<float4 *>(buf + ((t - lo) * kHid + kk) * 4);
                    const float *row = buf + 4 * kChunk * kHid + (t - lo) * kHid + kk;
                    float *dgh = DGH + (t & 1) * kDghStride;
                    const float4 q4 = *rzng;
                    const float c_r = q4.x, c_z = q4.y, c_n = q4.z, c_g = q4.w, c_dx = row[0];
                    const float hp = first ? hx0 * MSEQ[t] : row[kChunk * kHid];
                    const float dh = c_dx + carry;
                    const float dn = dh * (1.f - c_z);
                    const float dz = dh * (hp - c_n);
                    const float dpn = dn * (1.f - c_n * c_n);
                    const float dpr = dpn * c_g * c_r * (1.f - c_r);
                    const float dpz = dz * c_z * (1.f - c_z);
                    gbi[0] += dpr, gbi[1] += dpz, gbi[2] += dpn;
                    gbh[0] += dpr, gbh[1] += dpz, gbh[2] += dpn * c_r;
                    if (pp == 0)
                        dgh[dgh_pos(kk)] = dpr;
                    else if (pp == 1)
                        dgh[dgh_pos(kHid + kk)] = dpz;
                    else if (pp == 2)
                        dgh[dgh_pos(2 * kHid + kk)] = dpn * c_r;
                    else
                        *rzng = make_float4(dpr, dpz, dpn, 0.f); // dgi of the three gates, for dW_ih below
                    __syncthreads();
                    float2 s0 = make_float2(0.f, 0.f), s1 = make_float2(0.f, 0.f);
                    const float2 hp2 = make_float2(hp, hp);
#pragma unroll
                    for (int i = 0; i < 24; i += 2)
                    {
                        const float4 d = *reinterpret_cast<const float4 *>(dgh + pp * 52 + 2 * i);
                        const float2 lo2 = make_float2(d.x, d.y), hi2 = make_float2(d.z, d.w);
                        s0 = __ffma2_rn(wcol[i], lo2, s0);
                        s1 = __ffma2_rn(wcol[i + 1], hi2, s1);
                        gw[i] = __ffma2_rn(lo2, hp2, gw[i]);
                        gw[i + 1] = __ffma2_rn(hi2, hp2, gw[i + 1]);
                    }
                    float sacc = (s0.x + s0.y) + (s1.x + s1.y);
                    sacc += __shfl_xor_sync(0xffffffffu, sacc, 1);
                    sacc += __shfl_xor_sync(0xffffffffu, sacc, 2);
                    carry = first ? 0.f : fmaf(dh, c_z, sacc);
                }
                __syncthreads(); // the chunk's dgi rows are complete
                {
                    const int ns = hi - lo;
                    for (int i = tid; i < kGruGates * Dp; i += kGruThreads)
                    {
                        const int d = i / kGruGates, g = i - d * kGruGates;
                        const float *dgi = buf + (g & 63) * 4 + (g >> 6);
                        float acc = GWIH[i];
                        for (int sidx = 0; sidx < ns; ++sidx)
                            acc = fmaf(dgi[sidx * 4 * kHid], XSEQ[(lo + sidx) * Dp + d], acc);
                        GWIH[i] = acc;
                    }
                }
                __syncthreads(); // before the buffer is staged again
            }
            GRU_MARK(5);
            // Adam on the GRU parameters (W_hh from the registers, the rest resident in shared memory)
#pragma unroll
            for (int i0 = 0; i0 < 24; i0 += 4)
            {
                if (!ppo)
                {
#pragma unroll
                    for (int i = i0; i < i0 + 4; ++i)
                    {
                        atomicAdd(a.grad_out + po.whh + (pp * 48 + 2 * i) * kHid + kk, gw[i].x);
                        atomicAdd(a.grad_out + po.whh + (pp * 48 + 2 * i + 1) * kHid + kk, gw[i].y);
                    }
                    continue;
                }
                float m[8], v[8];
#pragma unroll
                for (int i = 0; i < 8; ++i)
                {
                    const int gi = po.whh + (pp * 48 + 2 * i0 + i) * kHid + kk;
                    m[i] = __ldcg(a.opt_m + gi), v[i] = __ldcg(a.opt_v + gi);
                }
#pragma unroll
                for (int i = 0; i < 8; ++i)
                {
                    const float2 w = wcol[i0 + (i >> 1)], g = gw[i0 + (i >> 1)];
                    adam_core(po.whh + (pp * 48 + 2 * i0 + i) * kHid + kk, (i & 1) ? g.y : g.x, (i & 1) ? w.y : w.x, m[i], v[i]);
                }
            }
            if (pp == 0)
            {
                float m[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f}, v[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
                if (ppo)
                {
#pragma unroll
                    for (int g = 0; g < 3; ++g)
                    {
                        m[g] = __ldcg(a.opt_m + po.bih + g * kHid + kk), v[g] = __ldcg(a.opt_v + po.bih + g * kHid + kk);
                        m[3 + g] = __ldcg(a.opt_m + po.bhh + g * kHid + kk), v[3 + g] = __ldcg(a.opt_v + po.bhh + g * kHid + kk);
                    }
                }
#pragma unroll
                for (int g = 0; g < 3; ++g)
                {
                    const int o = g * kHid + kk;
                    if (ppo)
                    {
                        BIH[o] = adam_core(po.bih + o, gbi[g], BIH[o], m[g], v[g]);
                        BHH[o] = adam_core(po.bhh + o, gbh[g], BHH[o], m[3 + g], v[3 + g]);
                    }
                    else
                    {
                        atomicAdd(a.grad_out + po.bih + o, gbi[g]);
                        atomicAdd(a.grad_out + po.bhh + o, gbh[g]);
                    }
                }
            }
            for (int i = tid; i < kGruGates * D; i += kGruThreads)
            {
                const int g = i / D, d = i - g * D;
                const int gi = po.wih + i, si = d * kGruGates + g;
                if (ppo)
                    WIHT[si] = adam_core(gi, GWIH[si], WIHT[si], __ldcg(a.opt_m + gi), __ldcg(a.opt_v + gi));
                else
                    atomicAdd(a.grad_out + gi, GWIH[si]);
            }
        }
        ++applied;
        GRU_MARK(6);
        __syncthreads(); // parameter stores of this minibatch are visible to the whole CTA before the next one starts
    } // minibatches
}
} // namespace

size_t gru_update_smem_bytes(int Dp, int A, int Ap, int head, int T)
{
    return static_cast<size_t>(make_gru_smem(Dp, A, Ap, head, T).total) * sizeof(float);
}

int launch_gru_update(const GruUpdateArgs &args, int grid, cudaStream_t s)
{
    static std::atomic<uint64_t> configured{0};
    if (!configured_on_device(configured))
    {
        CPPRL_CUDA_TRY(cudaFuncSetAttribute(gru_update_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        mark_configured_on_device(configured);
    }
    const size_t smem = gru_update_smem_bytes(args.Dp, args.A, args.Ap, args.head, args.T);
    gru_update_kernel<<<grid, kGruThreads, smem, s>>>(args);
    CPPRL_CUDA_TRY(cudaGetLastError());
    count_launch();
    return 0;
}
} // namespace cpprl_b200
