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
#include <cstdlib>

#include "common.cuh"
#include "mlp_update.cuh"
#include "mlp_blocks.cuh"
#include "mlp_tc.cuh"

namespace cpprl_b200
{
namespace
{
constexpr int kRedSplit = 8;               // warps splitting the partials of one parameter in phase 2
constexpr int kRedMaxPer = (148 + kRedSplit - 1) / kRedSplit; // partials one warp sums per granule (grid <= 148)

// Shared-memory carve-up (offsets in floats), identical on host and device.
struct SmemPlan
{
    int img, gsm, g2, x, h1[2], h2[2], logit, dlogit, val, dval, ret, oldlp, adv, act, rowid, stat, xn, actn, scn, rowidn, actin, norm, red, own, sig,
        misc, bc, total;
    // tensor-core variant: hi / lo operand planes (see mlp_tc.cuh); offsets in floats like the others
    int xb, xf, hb[2], hf[2];
    int stage; // tensor-core variant: scratch for the heads' partial sums (P3) and the staged dW1 / dW2 of the publish
    int img_base; // first float of the GLOBAL image that the shared-memory image holds (tensor-core variant: its tail)
    int g_base;   // first gradient index the shared-memory accumulator `gsm` holds (tensor-core variant: po.wv)
};

// Gradients whose sums are formed by more than one thread (head weights and b2: the two row halves of P5b; b1: the
// four warps of a tower in P7) would need shared-memory atomics, whose order varies from run to run.  Instead every
// contributing thread but the first owns a slot of a SECOND accumulator, and the slots are added to the gradient in
// a fixed order before it is published: the update gives the same bits every run.
//   G2:  [0, Ap*64) head weights, row half 1 | [.., +64) value weights, half 1
//   the bias slots -- b2[tower] of half 1 (128 floats), b1[tower][warp 1..3] (384) -- overlay the phase-2 reduction
//   scratch RED, which is idle while tiles are being processed (cfg3's 64-row plan has 3 KB of shared memory to spare)
__host__ __device__ inline int g2_floats(int Ap) { return Ap * kHid + kHid; }
constexpr int kG2BiasFloats = 2 * kHid + 2 * 3 * kHid;
constexpr int kTcW1Stage = 2 * kHid * mtc::kMaxDp; // floats reserved for the staged dW1 in front of the staged dW2 (tensor-core publish)
constexpr int kBcSteps = 64;

__host__ __device__ inline SmemPlan make_plan(int R, int img_floats, int Pstride, int Dp, int Ap, int per_cta, bool tc = false,
                                              int img_tail = 0, int g_tail = 0)
{
    const int LD = R + 4;
    SmemPlan p;
    int o = 0;
    // tensor-core variant: the tower matrices live in tensor memory and their gradients in registers; shared memory
    // holds the image's tail (biases, heads) and the gradient accumulators from the value head on (po.wv ..)
    p.img_base = tc ? img_tail : 0;
    p.g_base = tc ? g_tail : 0;
    p.img = o, o += img_floats - p.img_base;
    p.gsm = o, o += ((Pstride + 3) & ~3) - p.g_base;
    p.g2 = o, o += g2_floats(Ap); // second accumulators of the gradients that several threads share (see G2 below)
    p.xb = p.xf = p.hb[0] = p.hb[1] = p.hf[0] = p.hf[1] = p.stage = 0;
    if (tc)
    {
        // operand planes, hi then lo each (bytes / 4); the row-major planes read one non-existent 8-row group past
        // their end (N = 64 instructions over R <= 56 rows): something must follow them, which the order guarantees
        const int lbo_b = (R / 8) * 128 + 16, plane_b = 16 * lbo_b, plane_f = (R / 4) * 1024, plane_xf = (R / 4) * 528;
        for (int t = 0; t < 2; ++t)
            p.hb[t] = o, o += 2 * plane_b / 4;
        p.xf = o, o += 2 * plane_xf / 4;
        for (int t = 0; t < 2; ++t)
            p.hf[t] = o, o += 2 * plane_f / 4;
        // the row-major operand planes are idle when the scratch is in use (between the layer-2 chain and P5b; after
        // the last tile's data-gradient chain) and host it when they are large enough (56- and 64-row tiles)
        const int stage_need = kTcW1Stage + 2 * kHid * 68;
        if (4 * plane_b / 4 >= stage_need)
            p.stage = p.hb[0];
        else
            p.stage = o, o += stage_need;
        p.x = p.h1[0] = p.h1[1] = 0; // (fp32 x / h1 tiles are not kept)
    }
    else
    {
        p.x = o, o += Dp * LD;
        for (int t = 0; t < 2; ++t)
            p.h1[t] = o, o += kHid * LD;
    }
    for (int t = 0; t < 2; ++t)
        p.h2[t] = o, o += kHid * LD;
    // tensor-core variant: the row-major observation planes (layer 1's B operand: at most 2 x 8 x (8 x 128 + 16) bytes) are
    // live from finalize_stage to the end of the layer-1 chain, the fp32 h2 tile (2 x 64 x LD floats) from the layer-2
    // epilogue to P5b -- disjoint, so the planes overlay the tile (what lets the 64-row tile fit in 227 KB)
    if (tc)
        p.xb = p.h2[0];
    p.logit = o, o += Ap * LD;
    p.dlogit = o, o += Ap * LD;
    p.val = o, o += LD;
    p.dval = o, o += LD;
    p.ret = o, o += LD;
    p.oldlp = o, o += LD;
    p.adv = o, o += LD;
    p.act = o, o += Ap * LD;
    p.rowid = o, o += 2 * LD; // long long per row
    p.stat = o, o += (5 + Ap) * LD; // per-row loss terms and scale_log gradients, reduced at the start of P5
    p.xn = o, o += Dp * LD;     // staging buffers of the NEXT work item, filled by cp.async
    p.actn = o, o += Ap * LD;
    p.scn = o, o += 3 * LD;     // returns, old log-probs, old value predictions
    p.rowidn = o, o += 2 * LD;  // long long per row
    p.actin = o, o += 2 * LD;   // int64 actions (Categorical)
    p.norm = o, o += 2 * Dp + 4; // observation normaliser: mean[Dp], sqrt(var + 1e-8)[Dp], clip
    // phase-2 reduction scratch: [split][owned params + stats]; in phase 1 the bias slots of the second accumulators
    {
        const int red = kRedSplit * (per_cta + kNumStats);
        p.red = o, o += (per_cta > 0 && red < kG2BiasFloats) ? kG2BiasFloats : red;
    }
    p.own = o, o += 3 * per_cta;                       // this CTA's slice of params / Adam exp_avg / exp_avg_sq
    p.sig = o, o += 2 * Ap;                            // Normal head: {sigma^2, log sigma} per action
    p.misc = o, o += 48;
    p.bc = o, o += 2 * kBcSteps; // Adam bias corrections of the launch's first kBcSteps optimiser steps
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
            if (il.tc)
                image[il.w1t[t] + ((k >> 2) * kHid + j) * 4 + (k & 3)] = v; // [k / 4][unit][k % 4]: see ImageLayout
            else
                image[il.w1t[t] + k * kHid + perm16(j)] = v;
            return;
        }
        if (i >= po.b1[t] && i < po.b1[t] + kHid)
        {
            image[il.b1[t] + (il.tc ? i - po.b1[t] : perm16(i - po.b1[t]))] = v;
            return;
        }
        if (i >= po.w2[t] && i < po.w2[t] + kHid * kHid)
        {
            const int e = i - po.w2[t], j = e >> 6, k = e & 63;
            if (il.tc)
            {
                image[il.w2t[t] + ((k >> 2) * kHid + j) * 4 + (k & 3)] = v; // row j of W2, 16-byte chunk k / 4
                image[il.w2n[t] + ((j >> 2) * kHid + k) * 4 + (j & 3)] = v; // row k of W2^T, chunk j / 4
            }
            else
            {
                image[il.w2t[t] + k * kHid + perm16(j)] = v;
                image[il.w2n[t] + j * kHid + perm16(k)] = v;
            }
            return;
        }
        if (i >= po.b2[t] && i < po.b2[t] + kHid)
        {
            image[il.b2[t] + (il.tc ? i - po.b2[t] : perm16(i - po.b2[t]))] = v;
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

// ---- TMA bulk copy (cp.async.bulk, SASS UBLKCP) of the weight image global -> shared ------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

__device__ __forceinline__ void mbar_init(void *bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(void *bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(void *bar, uint32_t parity)
{
    uint32_t done = 0;
    const uint32_t addr = smem_u32(bar);
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
__device__ __forceinline__ void bulk_g2s(void *dst_smem, const void *src_gmem, uint32_t bytes, void *bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst_smem)),
                 "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

// ---- grid-wide barrier, split into arrive / wait ------------------------------------------------------
// `bar` is one unsigned counter zeroed by the host before every launch; barrier number k (0-based) of the
// launch is complete when the counter reaches G*(k+1).  Work placed between arrive and wait overlaps the
// round trip of the slowest CTA.
// Ordering: the CTA barrier orders every thread's writes before thread 0's release-increment (release is cumulative
// over what the releasing thread has observed), and thread 0's acquire-load before every thread's later reads through
// the second CTA barrier -- the scheme of CUTLASS's GenericBarrier, without the two full fences (MEMBAR.SC.GPU,
// ~0.5 k cycles each on the critical path of every optimiser step) an earlier version wrapped around it.  Everything a
// CTA reads from another CTA after a barrier bypasses L1 (ld.global.cg, TMA bulk copies).
__device__ __forceinline__ void grid_arrive(unsigned int *bar)
{
    __syncthreads();
    if (threadIdx.x == 0)
        asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(bar) : "memory");
}
__device__ __forceinline__ void grid_wait(unsigned int *bar, unsigned int target)
{
    if (threadIdx.x == 0)
    {
        unsigned int v;
        do
        {
            asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(bar) : "memory");
        } while (v < target);
    }
    __syncthreads();
}

// ---- the kernel -------------------------------------------------------------------------------------
template <int R, bool TC>
__global__ void __launch_bounds__(kUpdThreads, 1) mlp_update_kernel(const MlpUpdateArgs a)
{
    constexpr int LD = R + 4;
    extern __shared__ __align__(16) float smem[];
    const SmemPlan sp = make_plan(R, a.il.floats, a.po.Pstride, a.Dp, a.Ap, a.per_cta, TC, a.il.tail, a.po.wv);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tower = tid >> 7, u = tid & 127; // warps 0-3 run the actor tower, 4-7 the critic tower
    const int jb = u & 15, rb = u >> 4;
    (void)jb, (void)rb;
    // tensor-core variant: thread = (tensor-memory lane F = hidden unit jF of tower tF, column half ch of the tile)
    const int q4 = warp & 3, ch = warp >> 2, F = 32 * q4 + lane, tF = F >> 6, jF = F & 63;
    (void)ch, (void)tF, (void)jF;
    // every pointer is `smem + integer` so the compiler keeps the shared state space (LDS/STS, never generic LD/ST).
    // `img` and `gsm` are indexed with GLOBAL image / gradient offsets: the tensor-core variant keeps only their tails
    // in shared memory (sp.img_base / sp.g_base are zero otherwise)
    float *img = smem + (sp.img - sp.img_base);
    float *gsm = smem + (sp.gsm - sp.g_base);
    float *G2 = smem + sp.g2;
    float *G2_WO = G2, *G2_WV = G2 + a.Ap * kHid;
    float *G2_B2 = smem + sp.red, *G2_B1 = G2_B2 + 2 * kHid; // overlay RED (idle in phase 1)
    float *X = smem + sp.x;
    float *h1t = smem + sp.h1[0] + tower * (kHid * LD); // this thread's tower
    float *h2t = smem + sp.h2[0] + tower * (kHid * LD);
    float *LOGIT = smem + sp.logit, *DLOGIT = smem + sp.dlogit;
    float *VAL = smem + sp.val, *DVAL = smem + sp.dval, *RET = smem + sp.ret, *OLDLP = smem + sp.oldlp,
          *ADV = smem + sp.adv, *ACT = smem + sp.act, *RED = smem + sp.red, *STAT = smem + sp.stat;
    float *OWN_P = smem + sp.own, *OWN_M = OWN_P + a.per_cta, *OWN_V = OWN_M + a.per_cta;
    float *SIG = smem + sp.sig;
    long long *ROWID = reinterpret_cast<long long *>(smem + sp.rowid);
    float *MISC = smem + sp.misc;
    void *mbar = MISC + 24;
    // tensor-core variant: five mbarriers (one per contraction stage of a tile) and the tensor-memory base address
    uint64_t *tcbar = reinterpret_cast<uint64_t *>(MISC + 26);
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(MISC + 38);
    const float *w1t_img = img + (tower ? a.il.w1t[1] : a.il.w1t[0]);
    const float *b1_img = img + (tower ? a.il.b1[1] : a.il.b1[0]);
    const float *w2t_img = img + (tower ? a.il.w2t[1] : a.il.w2t[0]);
    const float *w2n_img = img + (tower ? a.il.w2n[1] : a.il.w2n[0]);
    const float *b2_img = img + (tower ? a.il.b2[1] : a.il.b2[0]);
    const int po_w1 = tower ? a.po.w1[1] : a.po.w1[0], po_b1 = tower ? a.po.b1[1] : a.po.b1[0];
    const int po_w2 = tower ? a.po.w2[1] : a.po.w2[0], po_b2 = tower ? a.po.b2[1] : a.po.b2[0];
    (void)X, (void)h1t, (void)w1t_img, (void)b1_img, (void)w2t_img, (void)w2n_img, (void)b2_img, (void)po_w1, (void)po_b1, (void)po_w2,
        (void)po_b2, (void)G2_B1, (void)tcbar, (void)tmem_slot;

    if (a.dbg_clocks && threadIdx.x == 0)
    {
        a.dbg_clocks[blockIdx.x * 32 + 27] = clock64(); // kernel entry
        if (blockIdx.x == 0)
        {
            unsigned long long ns;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(ns));
            a.dbg_clocks[148 * 32 + 2 * (a.dbg_launch & 63)] = static_cast<long long>(ns);
        }
    }
    const int G = gridDim.x;
    const int P = a.po.P, Pstride = a.po.Pstride, Pstat = Pstride - kNumStats;
    const int D = a.D, Dp = a.Dp, A = a.A;
    const bool training = TC || a.algo != kAlgoForward; // (the tensor-core variant only trains)
    // rows / tiles / publishing CTAs of minibatch `step` (ragged when the rollout is sharded by environment)
    auto rows_of = [&](int step) { return a.step_rows ? __ldg(a.step_rows + step) : a.B; };
    const int world = a.comm.world, my_rank = a.comm.rank;
    // phase-2 ownership: parameters [c0, c0 + n_own) of the flat vector (n_own a multiple of 4, may be 0)
    const int c0 = min(Pstat, static_cast<int>(blockIdx.x) * a.per_cta);
    const int n_own = min(Pstat, c0 + a.per_cta) - c0;
    unsigned int bar_target = 0;

    // ---- prologue: read persistent state, build the weight image from the parameters -----------------
    int stopped = 0, applied = 0;
    long long opt_step0 = 0;
    if (training)
    {
        stopped = a.state->stopped;
        opt_step0 = a.state->opt_step;
        for (int i = tid; i < n_own; i += kUpdThreads)
        {
            const int gi = c0 + i;
            const bool real = gi < P;
            OWN_P[i] = real ? a.params[gi] : 0.f;
            OWN_V[i] = real ? a.opt_v[gi] : 0.f;
            OWN_M[i] = (real && a.opt == kOptAdam) ? a.opt_m[gi] : 0.f;
        }
    }
    // Adam's bias corrections as host-style doubles (libtorch Adam: 1 - pow(beta, t), t = steps applied so far + 1) for
    // every optimiser step of this launch, one step per lane of the last warp: a double pow is ~2 k cycles of dependent
    // software arithmetic, which one thread used to spend between the arrival at the grid barrier and its poll in EVERY
    // step -- 4 k cycles on the critical path of each (the barrier itself had long been satisfied)
    float *BC = smem + sp.bc;
    if (training && a.opt == kOptAdam && tid >= kUpdThreads - 32)
        for (int i = tid - (kUpdThreads - 32); i < min(a.n_steps, kBcSteps); i += 32)
        {
            const double t = static_cast<double>(opt_step0 + i + 1);
            const double bc1 = 1.0 - pow(static_cast<double>(a.beta1), t);
            const double bc2 = 1.0 - pow(static_cast<double>(a.beta2), t);
            BC[2 * i] = static_cast<float>(a.lr / bc1);
            BC[2 * i + 1] = static_cast<float>(sqrt(bc2));
        }
    for (int i = blockIdx.x * kUpdThreads + tid; i < P; i += G * kUpdThreads)
        image_scatter(a, a.image, i, a.params[i]);
    if (tid == 0)
    {
        float mean = 0.f, denom = 1.f;
        if (a.adv_sums)
        {
            double s1 = a.adv_sums[0], s2 = a.adv_sums[1];
            if (world > 1)
            {
                // global moments: every rank publishes its shard's sums once per update (CTA 0), reads all of them in
                // rank order (identical value everywhere)
                if (blockIdx.x == 0)
                    flag_publish(&a.comm.area[my_rank]->adv_flag, a.adv_seq);
                s1 = 0.0, s2 = 0.0;
                for (int r = 0; r < world; ++r)
                {
                    if (r != my_rank && !flag_wait(&a.comm.area[r]->adv_flag, a.adv_seq))
                        a.state->comm_error = 1;
                    s1 += r == my_rank ? a.adv_sums[0] : peer_load(a.comm.area[r]->adv);
                    s2 += r == my_rank ? a.adv_sums[1] : peer_load(a.comm.area[r]->adv + 1);
                }
            }
            const double cnt = static_cast<double>(a.rows_norm);
            const double m = s1 / cnt;
            double var = (s2 - s1 * m) / (cnt - 1.0); // unbiased (ppo.cpp:57)
            if (var < 0.0)
                var = 0.0;
            mean = static_cast<float>(m);
            denom = static_cast<float>(sqrt(var)) + 1e-5f;
        }
        MISC[0] = mean;
        MISC[1] = denom;
        mbar_init(mbar, 1);
        if (TC)
            for (int b = 0; b < 5; ++b)
                mbar_init(&tcbar[b], b == 0 ? 1 : 2); // layer 1 has one issuing thread, the tower-wise stages two
    }
    uint32_t tm = 0;        // tensor-memory base (tensor-core variant)
    uint32_t tc_parity = 0; // parity of the four per-tile mbarriers: tiles processed so far & 1
    (void)tm, (void)tc_parity;
    if (TC)
    {
        if (warp == 0)
        {
            asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(mtc::kTmemCols)
                         : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
        }
        mtc::fence_before();
        __syncthreads();
        mtc::fence_after();
        tm = *tmem_slot;
    }
    const uint32_t lane_base = tm + (static_cast<uint32_t>(32 * q4) << 16); // this warp's 32 tensor-memory lanes
    (void)lane_base;
    // Matrix rows -> tensor memory in two steps, so that the global (L2) loads are in flight while something else runs:
    // tc_fetch32 issues the eight 16-byte loads of this thread's 32 floats, tc_store32 splits them into hi / lo planes
    // and writes columns [col, col + 32) / [col + 64, ..) of the thread's lane (warp-collective tcgen05.st).
    // `mat` is a [16 chunks][64 units][4] matrix of the image: chunk c of unit jF is the 16-byte word c * 64 + jF, so
    // the 32 lanes of a warp read 512 contiguous bytes per load
    auto tc_fetch32 = [&](const float *mat, float4 (&w)[8]) {
        const float4 *src = reinterpret_cast<const float4 *>(mat) + (8 * ch) * kHid + jF;
#pragma unroll
        for (int i = 0; i < 8; ++i)
            w[i] = __ldcg(src + i * kHid);
    };
    auto tc_store32 = [&](uint32_t col, const float4 (&w)[8]) {
        const float4 lo4[4] = {w[0], w[1], w[2], w[3]}, hi4[4] = {w[4], w[5], w[6], w[7]};
        mtc::store_split16(col, 64, lo4);
        mtc::store_split16(col + 16, 64, hi4);
    };
    // every exit path of the tensor-core variant hands its tensor memory back first (CTA-uniform decisions only)
    auto leave = [&]() {
        if (TC)
        {
            mtc::fence_before();
            __syncthreads();
            if (warp == 0)
                asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tm), "r"(mtc::kTmemCols) : "memory");
        }
    };
    const int n_tiles_max = (a.B + R - 1) / R;
    (void)n_tiles_max;
    float *NMEAN = smem + sp.norm, *NSTD = NMEAN + Dp, *NCLIP = NSTD + Dp;
    for (int k = tid; k < Dp; k += kUpdThreads)
        NMEAN[k] = 0.f, NSTD[k] = 1.f;
    __syncthreads();
    if (a.norm_mean)
    {
        // the normaliser is constant during an update (PPO refreshes it after the last epoch, A2C before the launch)
        for (int k = tid; k < D; k += kUpdThreads)
        {
            NMEAN[k] = a.norm_mean[k];
            NSTD[k] = sqrtf(a.norm_var[k] + 1e-8f);
        }
        if (tid == 0)
            NCLIP[0] = a.norm_clip[0];
    }

    // ---- asynchronous gather of the NEXT work item (tile of a minibatch) -------------------------------------
    // The rollout does not depend on the parameters, so the rows of the next tile are fetched with cp.async
    // (LDGSTS) into a second set of staging buffers while the current tile computes:
    //   prefetch_rowid : (top of an item) row id of the next item through the permutation -> register
    //   prefetch_issue : (two phases later) ids -> ROWIDN, then one 4-byte cp.async per observation / action
    //                    element and per-row scalar, global-coalesced along the row
    //   finalize_stage : (top of the next item) wait, then normalise / clamp the observations
    //                    (observation_normalizer.cpp:71-73) while moving them to the feature-major tile X,
    //                    form the normalised advantage (ppo.cpp:57) and narrow the int64 actions.
    float *XN = smem + sp.xn, *ACTN = smem + sp.actn, *SCN = smem + sp.scn;
    long long *ROWIDN = reinterpret_cast<long long *>(smem + sp.rowidn);
    long long *ACTIN = reinterpret_cast<long long *>(smem + sp.actin);
    auto item_rowid = [&](int step, int tile) -> long long {
        long long row = -1;
        if (tid < R)
        {
            const int row0 = tile * R;
            if (row0 + tid < rows_of(step))
            {
                const long long pos = static_cast<long long>(step) * a.perm_stride + row0 + tid;
                row = a.perm ? __ldg(a.perm + pos) : pos;
            }
        }
        return row;
    };
    auto cp_async4 = [](float *dst, const float *src) {
        asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
    };
    // Tensor-core variant: warps 0 and 4 hold the threads that issue the MMA chains and block on the tensor core's queue
    // while a chain is in flight, so work that is meant to run UNDER a chain goes to the other six warps only:
    // wt = this thread's index among them (0..191), -1 for the issuing warps.
    const int wt = !TC ? tid : (q4 == 0 ? -1 : tid - 32 * (1 + ch));
    constexpr int kWorkThreads = TC ? kUpdThreads - 64 : kUpdThreads;
    auto prefetch_issue = [&]() { // ROWIDN must be visible to every thread
        if (wt >= 0)
        {
            constexpr int LPR = kWorkThreads / R; // lanes per row: consecutive lanes fetch consecutive elements of a row
            const int r = wt / LPR, q = wt % LPR;
            const long long row = r < R ? ROWIDN[r] : -1; // (R need not divide the thread count)
            if (row >= 0)
            {
                const float *src = a.obs + row * D;
                for (int k = q; k < D; k += LPR)
                    cp_async4(XN + k * LD + r, src + k);
                if (a.head == 1)
                {
                    const float *asrc = static_cast<const float *>(a.actions) + row * A;
                    for (int k = q; k < A; k += LPR)
                        cp_async4(ACTN + k * LD + r, asrc + k);
                }
            }
        }
        if (wt >= 0 && wt < R)
        {
            const long long row = ROWIDN[wt];
            if (row >= 0)
            {
                if (training)
                {
                    cp_async4(SCN + wt, a.returns + row);
                    cp_async4(SCN + LD + wt, a.old_logp + row);
                    if (a.adv_sums)
                        cp_async4(SCN + 2 * LD + wt, a.value_preds + row);
                }
                if (a.head == 0)
                    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(smem_u32(ACTIN + wt)),
                                 "l"(static_cast<const long long *>(a.actions) + row)
                                 : "memory");
            }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    int mark_step = -1; // the optimiser step in progress (the marks below are also used inside lambdas defined before the loop)
// Phase marks: one opaque predicated store of %clock64 (thread 0 of a CTA, the step being looked at).  A C++ "if (tid == 0
// && ...)" here lets the compiler thread the known tid into the branches that follow, and thread 0 then runs the next phase
// diverged from the rest of its warp -- which showed up as a phase twice as long as it is in an unmarked run.
#define CPPRL_MARK(slot)                                                                                       \
    asm volatile("{\n\t.reg .pred p;\n\t.reg .u64 t;\n\tsetp.ne.u32 p, %0, 0;\n\tmov.u64 t, %%clock64;\n\t"   \
                 "@p st.global.u64 [%1], t;\n\t}" ::"r"(static_cast<unsigned>(a.dbg_clocks && tid == 0 && mark_step == a.dbg_step)), \
                 "l"(a.dbg_clocks + blockIdx.x * 32 + (slot))                                                  \
                 : "memory")
    // The rows of a work item, from the staging buffers the cp.async prefetch filled to the operands of the tile phases.
    // Tensor-core variant, pieces (thread index t of nt threads, so that they can also run on the non-issuing warps under
    // an MMA chain): stage_obs normalises the observations IN PLACE (XN) and writes the row-major planes (layer 1's B
    // operand: they overlay the h2 tile, dead after P5b); stage_rows the actions and per-row terms (dead after P4);
    // stage_xf the column-major planes from the normalised XN (dW1's B operand: in use until the tile's last chain ends).
    bool pre_staged = false; // stage_obs + stage_rows of the staged item already ran (under the last tile's dW chains)
    auto stage_obs = [&](int nrows, int t, int nt) {
        using GeoT = mtc::Geo<TC ? R : 8>;
        const float clip = a.norm_mean ? NCLIP[0] : 0.f;
        const bool norm = a.norm_mean != nullptr;
        unsigned char *xb = reinterpret_cast<unsigned char *>(smem + sp.xb);
        for (int it = t; it < (Dp >> 2) * R; it += nt)
        {
            const int kg = it / R, r = it - kg * R; // R is a compile-time constant
            float x[4], hi[4], lo[4];
#pragma unroll
            for (int j = 0; j < 4; ++j)
                x[j] = XN[(4 * kg + j) * LD + r];
            bool ok[4];
#pragma unroll
            for (int j = 0; j < 4; ++j)
                ok[j] = 4 * kg + j < D && r < nrows;
            if (norm)
            {
                // (padding columns and rows divide 1 by the deviation instead of 0: a zero numerator sends the exact
                // division -- and with it the whole warp -- down its ~500-cycle slow path)
                const float4 mean = *reinterpret_cast<const float4 *>(NMEAN + 4 * kg), sd = *reinterpret_cast<const float4 *>(NSTD + 4 * kg);
                x[0] = (ok[0] ? x[0] - mean.x : 1.f) / sd.x, x[1] = (ok[1] ? x[1] - mean.y : 1.f) / sd.y;
                x[2] = (ok[2] ? x[2] - mean.z : 1.f) / sd.z, x[3] = (ok[3] ? x[3] - mean.w : 1.f) / sd.w;
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    x[j] = fminf(fmaxf(x[j], -clip), clip);
            }
#pragma unroll
            for (int j = 0; j < 4; ++j)
            {
                x[j] = ok[j] ? x[j] : 0.f;
                XN[(4 * kg + j) * LD + r] = x[j];
                hi[j] = mtc::tf32_hi(x[j]), lo[j] = x[j] - hi[j];
            }
            const uint32_t ob = 16u * r + kg * GeoT::kLboB;
            *reinterpret_cast<float4 *>(xb + ob) = make_float4(hi[0], hi[1], hi[2], hi[3]);
            *reinterpret_cast<float4 *>(xb + GeoT::plane_xb(Dp) + ob) = make_float4(lo[0], lo[1], lo[2], lo[3]);
        }
    };
    auto stage_xf = [&](int t, int nt) {
        using GeoT = mtc::Geo<TC ? R : 8>;
        unsigned char *xf = reinterpret_cast<unsigned char *>(smem + sp.xf);
        for (int it = t; it < (Dp >> 2) * R; it += nt)
        {
            const int kg = it / R, r = it - kg * R;
            const uint32_t of = 64u * kg + (r >> 2) * GeoT::kLboXF + (r & 3) * 4u;
#pragma unroll
            for (int j = 0; j < 4; ++j)
            {
                const float x = XN[(4 * kg + j) * LD + r];
                const float hi = mtc::tf32_hi(x);
                *reinterpret_cast<float *>(xf + of + 16u * j) = hi;
                *reinterpret_cast<float *>(xf + GeoT::kPlaneXF + of + 16u * j) = x - hi;
            }
        }
    };
    auto stage_rows = [&](int nrows, int t, int nt) {
        if (a.head == 1)
#pragma unroll 2
            for (int e = t; e < A * R; e += nt)
            {
                const int k = e / R, r = e - k * R;
                ACT[k * LD + r] = r < nrows ? ACTN[k * LD + r] : 0.f;
            }
        if (t < R)
        {
            const bool ok = t < nrows;
            const float ret = ok && training ? SCN[t] : 0.f;
            const float olp = ok && training ? SCN[LD + t] : 0.f;
            const float vold = ok && training && a.adv_sums ? SCN[2 * LD + t] : 0.f;
            RET[t] = ret;
            OLDLP[t] = olp;
            ADV[t] = (ret - vold - MISC[0]) / MISC[1]; // (adv - mean) / (std + 1e-5)              (ppo.cpp:57)
            if (a.head == 0)
                reinterpret_cast<int *>(ACT)[t] = ok ? static_cast<int>(ACTIN[t]) : 0;
        }
    };
    auto finalize_stage = [&](int nrows) { // nrows = valid rows of the item (row r is valid iff r < nrows)
        if (!(TC && pre_staged))
        {
            // (a pre-staged item's rows were waited for, and published by a barrier, before the chains that staged them:
            // its remaining piece starts at once -- while thread 0's warp still sits in the release of the grid arrival)
            asm volatile("cp.async.wait_all;" ::: "memory");
            __syncthreads();
        }
        if (TC)
        {
            if (!pre_staged)
            {
                stage_obs(nrows, tid, kUpdThreads);
                stage_rows(nrows, tid, kUpdThreads);
            }
            if (!pre_staged)
                stage_xf(tid, kUpdThreads); // (same item -> thread mapping as stage_obs: a thread reads back its own values)
            else if (warp > 0)
                stage_xf(tid - 32, kUpdThreads - 32); // (warp 0's stores would queue behind its release fence)
            pre_staged = false;
        }
        else
        {
            const float clip = a.norm_mean ? NCLIP[0] : 0.f;
            const bool norm = a.norm_mean != nullptr;
#pragma unroll 4
            for (int e = tid; e < Dp * R; e += kUpdThreads)
            {
                const int k = e / R, r = e - k * R; // R is a compile-time constant
                const bool ok = k < D && r < nrows;
                float x = XN[k * LD + r];
                if (norm)
                {
                    x = (ok ? x - NMEAN[k] : 1.f) / NSTD[k]; // (never 0 / s: see stage_obs)
                    x = fminf(fmaxf(x, -clip), clip);
                }
                X[k * LD + r] = ok ? x : 0.f;
            }
            stage_rows(nrows, tid, kUpdThreads);
        }
        __syncthreads();
    };

    unsigned int *const gbar = a.barrier + 8 * a.barrier_sel;
    if (blockIdx.x == 0 && tid == 0)
        a.barrier[8 * (a.barrier_sel ^ 1)] = 0u; // the NEXT launch's counter (stream order: the previous launch is complete)
    grid_arrive(gbar); // the image is complete once every CTA has scattered its share of the parameters
    bar_target += G;
    grid_wait(gbar, bar_target);
    if (a.dbg_clocks && tid == 0)
        a.dbg_clocks[blockIdx.x * 32 + 28] = clock64(); // prologue done (image built, first grid barrier passed)
    if (stopped)
    {
        leave();
        return;
    }
    const float inv_b = 1.f / static_cast<float>(a.B_norm);
    uint32_t img_phase = 0;
    bool staged = false; // the first work item of the next step was prefetched during this one
    bool fin_ahead = false; // ... and already finalised (finalize_stage) while waiting at a grid barrier
    // the optimiser state of this CTA's slice lives in shared memory for the whole launch; it goes back to
    // global memory on every exit path
    auto write_back_state = [&]() {
        for (int i = tid; i < n_own; i += kUpdThreads)
        {
            const int gi = c0 + i;
            if (gi < P)
            {
                a.opt_v[gi] = OWN_V[i];
                if (a.opt == kOptAdam)
                    a.opt_m[gi] = OWN_M[i];
            }
        }
    };

    for (int step = 0; step < a.n_steps; ++step)
    {
        mark_step = step;
        const int n_tiles = (rows_of(step) + R - 1) / R;
        const int n_act = min(n_tiles, G);
        const bool active = static_cast<int>(blockIdx.x) < n_act; // owns tiles of this minibatch, publishes a partial
        CPPRL_MARK(0);
        float4 wreg[8], w1reg[4], timg[2]; // matrix rows / image tail on their way to tensor / shared memory (tensor-core variant)
        (void)wreg, (void)w1reg, (void)timg;
        // ---- (re)load the weight image with one TMA bulk copy per 32 KB, clear the gradient accumulator ------
        if (active)
        {
            if (!TC && tid == 0)
            {
                const uint32_t bytes = static_cast<uint32_t>(a.il.floats - sp.img_base) * 4u;
                // the image was written by generic-proxy stores of other CTAs (ordered by the grid barrier); make
                // them visible to the async proxy that performs the bulk copy
                asm volatile("fence.proxy.async;" ::: "memory");
                mbar_expect_tx(mbar, bytes);
                for (uint32_t off = 0; off < bytes; off += 32768u)
                    bulk_g2s(reinterpret_cast<char *>(smem + sp.img) + off,
                             reinterpret_cast<const char *>(a.image + sp.img_base) + off, min(32768u, bytes - off), mbar);
            }
            if (TC)
            {
                // this lane's row of W2 (the 32 columns of this thread's half) and of W1 (16 columns): the global image is
                // complete (grid barrier); the loads fly while the shared-memory image arrives
                tc_fetch32(a.image + (tF ? a.il.w2t[1] : a.il.w2t[0]), wreg);
                const float4 *w1src = reinterpret_cast<const float4 *>(a.image + (tF ? a.il.w1t[1] : a.il.w1t[0])) + (4 * ch) * kHid + jF;
#pragma unroll
                for (int i = 0; i < 4; ++i)
                    w1reg[i] = 16 * ch + 4 * i < Dp ? __ldcg(w1src + i * kHid) : make_float4(0.f, 0.f, 0.f, 0.f);
                // the image's tail (biases, heads: at most 345 16-byte words) with the same loads -- one L2 round trip
                // for everything instead of a bulk copy and its mbarrier
                const float4 *tsrc = reinterpret_cast<const float4 *>(a.image + sp.img_base);
#pragma unroll
                for (int i = 0; i < 2; ++i)
                    if (tid + i * kUpdThreads < (a.il.floats - sp.img_base) / 4)
                        timg[i] = __ldcg(tsrc + tid + i * kUpdThreads);
            }
            for (int i = sp.g_base + tid; i < Pstride; i += kUpdThreads)
                gsm[i] = 0.f;
            for (int i = tid; i < g2_floats(a.Ap); i += kUpdThreads)
                G2[i] = 0.f;
            if (training) // (forward launches carry no reduction scratch)
                for (int i = tid; i < kG2BiasFloats; i += kUpdThreads)
                    G2_B2[i] = 0.f;
            if (!staged)
            {
                const long long row = item_rowid(step, blockIdx.x);
                if (tid < R)
                    ROWIDN[tid] = row;
                __syncthreads();
                prefetch_issue();
            }
            staged = false;
            if (TC)
            {
#pragma unroll
                for (int i = 0; i < 2; ++i)
                    if (tid + i * kUpdThreads < (a.il.floats - sp.img_base) / 4)
                        reinterpret_cast<float4 *>(smem + sp.img)[tid + i * kUpdThreads] = timg[i];
            }
            else
            {
                mbar_wait(mbar, img_phase);
                img_phase ^= 1u;
            }
            if (!TC && a.head == 1 && tid < A)
            {
                const float sigma = expf(img[a.il.slog + tid]); // Normal(loc, exp(scale_log))  (output_layers.cpp:54-59)
                SIG[2 * tid] = sigma * sigma;
                SIG[2 * tid + 1] = logf(sigma);
            }
            CPPRL_MARK(26);
            if (TC)
            {
                tc_store32(lane_base + mtc::kColW + 32 * ch, wreg);
                mtc::store_split16(lane_base + mtc::kColX + 16 * ch, 32, w1reg);
                mtc::wait_st();
                mtc::fence_before();
            }
        }
        else if (training && !staged && step + 1 < a.n_steps && static_cast<long long>(blockIdx.x) * R < rows_of(step + 1))
        {
            // No tile in this minibatch but one in the next (env-sharded updates: a rank's share of a minibatch is ragged,
            // the CTAs around the tile count drop in and out): fetch that tile now and lay it out while waiting at the grid
            // barrier below, like the CTAs that prefetched under their last tile -- otherwise this CTA would start the next
            // step with a cold fetch + the full staging on the critical path, and every other CTA would wait for its partial.
            const long long row = item_rowid(step + 1, blockIdx.x);
            if (tid < R)
                ROWIDN[tid] = row;
            __syncthreads();
            prefetch_issue();
            staged = true;
        }
        // gradients of the tower matrices of this minibatch: registers of the tensor-core variant
        float acc_w2[TC ? 32 : 1], acc_w1[TC ? 16 : 1], acc_b1 = 0.f, acc_b2 = 0.f;
#pragma unroll
        for (int i = 0; i < (TC ? 32 : 1); ++i)
            acc_w2[i] = 0.f;
#pragma unroll
        for (int i = 0; i < (TC ? 16 : 1); ++i)
            acc_w1[i] = 0.f;
        (void)acc_w2, (void)acc_w1, (void)acc_b1, (void)acc_b2;
        __syncthreads();
        if (TC && active && a.head == 1 && tid < A)
        {
            // (the tail was stored by all threads just above; P4, the reader, is several barriers away)
            const float sigma = expf(img[a.il.slog + tid]);
            SIG[2 * tid] = sigma * sigma;
            SIG[2 * tid + 1] = logf(sigma);
        }
        const float ce_b = a.c_entropy * inv_b;
        CPPRL_MARK(1);

        for (int tile = blockIdx.x; active && tile < n_tiles; tile += G)
        {
            const int row0 = tile * R;
            const int nrows = min(R, rows_of(step) - row0);
            // next work item of this CTA: the following tile of this minibatch, else its first tile of the next one
            int nstep = step, ntile = tile + G;
            if (ntile >= n_tiles)
                nstep = step + 1, ntile = blockIdx.x;
            const bool has_next = nstep < a.n_steps && (nstep == step || ntile * R < rows_of(nstep));
            if (fin_ahead)
                fin_ahead = false; // (done while the previous minibatch waited at its grid barrier)
            else
                finalize_stage(nrows); // the rows of THIS item are in place (waits for their cp.async group)
            const long long next_row = has_next ? item_rowid(nstep, ntile) : -1;
            if (TC)
            {
                using GeoT = mtc::Geo<TC ? R : 8>;
                constexpr int CH = GeoT::CH;
                const uint32_t xb_hi = smem_u32(smem + sp.xb), xb_lo = xb_hi + GeoT::plane_xb(Dp);
                unsigned char *hb_hi = reinterpret_cast<unsigned char *>(smem + sp.hb[tF]), *hb_lo = hb_hi + GeoT::kPlaneB;
                unsigned char *hf_hi = reinterpret_cast<unsigned char *>(smem + sp.hf[tF]), *hf_lo = hf_hi + GeoT::kPlaneF;
                CPPRL_MARK(20);
                // ---- T1: layer 1 of both towers: D[F][r] = sum_k W1[F][k] x[r][k]  (A = W1 in tensor memory) --------
                mtc::fence_async_smem(); // the operand planes finalize_stage wrote -> visible to the tensor core
                mtc::fence_before();
                if (tid < R)
                    ROWIDN[tid] = next_row; // read by prefetch_issue after the next barrier
                __syncthreads();
                if (tid == 0)
                {
                    mtc::fence_after();
                    const uint32_t d = tm + mtc::kColA, ah = tm + mtc::kColX, al = ah + 32, id = mtc::idesc_n(GeoT::kN);
                    switch (Dp / 8) // (k-steps are compile-time constants: every descriptor is base + immediate)
                    {
                    case 1: mtc::issue_3xtf32<0, 1, GeoT::kLboB>(d, ah, al, xb_hi, xb_lo, id); break;
                    case 2: mtc::issue_3xtf32<0, 2, GeoT::kLboB>(d, ah, al, xb_hi, xb_lo, id); break;
                    case 3: mtc::issue_3xtf32<0, 3, GeoT::kLboB>(d, ah, al, xb_hi, xb_lo, id); break;
                    default: mtc::issue_3xtf32<0, 4, GeoT::kLboB>(d, ah, al, xb_hi, xb_lo, id); break;
                    }
                    mtc::commit(&tcbar[0]);
                }
                mbar_wait(&tcbar[0], tc_parity);
                mtc::fence_after();
                CPPRL_MARK(2);
                // h1 = tanh(. + b1) as the two operands it feeds: row-major [row][unit] (layer 2) and unit-major
                // [unit][row] (dW2), hi / lo planes each
                {
                    const float bias = img[(tF ? a.il.b1[1] : a.il.b1[0]) + jF];
                    const uint32_t ob = (jF >> 2) * GeoT::kLboB + (jF & 3) * 4u;
                    // (rolled, four columns per tensor-memory load: the phase is bound by its 70 shared-memory stores and the
                    // issue slots of 28 activations per thread; one wide load + straight-line code measured 15 % slower here,
                    // while the store-free dz1 epilogue below gained 25 % from it)
#pragma unroll 1
                    for (int c = 0; c < CH; c += 4)
                    {
                        float v[4], hi[4], lo[4];
                        mtc::ld4(lane_base + mtc::kColA + ch * CH + c, v);
                        const int r0 = ch * CH + c;
#pragma unroll
                        for (int i = 0; i < 4; ++i)
                        {
                            const float h = mtc::tanh_sfu(v[i] + bias);
                            hi[i] = mtc::tf32_hi(h), lo[i] = h - hi[i];
                            *reinterpret_cast<float *>(hb_hi + 16u * (r0 + i) + ob) = hi[i];
                            *reinterpret_cast<float *>(hb_lo + 16u * (r0 + i) + ob) = lo[i];
                        }
                        const uint32_t of = 16u * jF + (r0 >> 2) * GeoT::kLboF;
                        *reinterpret_cast<float4 *>(hf_hi + of) = make_float4(hi[0], hi[1], hi[2], hi[3]);
                        *reinterpret_cast<float4 *>(hf_lo + of) = make_float4(lo[0], lo[1], lo[2], lo[3]);
                    }
                }
                mtc::fence_async_smem();
                mtc::fence_before();
                __syncthreads();
                CPPRL_MARK(21);
                // ---- T2: layer 2, one instruction chain per tower (lanes 0..63 of the first, 64..127 of the second) ----
                if ((tid & 127) == 0) // one issuing thread per tower (warps 0 and 4): the chains are independent
                {
                    const int t = tid >> 7;
                    mtc::fence_after();
                    const uint32_t b_hi = smem_u32(smem + sp.hb[t]);
                    mtc::issue_3xtf32<0, kHid / 8, GeoT::kLboB>(tm + mtc::kColA + 64 * t, tm + mtc::kColW, tm + mtc::kColW + 64, b_hi,
                                                                b_hi + GeoT::kPlaneB, mtc::idesc_n(GeoT::kN));
                    mtc::commit(&tcbar[1]);
                }
                // W2^T (this lane's row of the transposed image) for the data gradient: in flight during the chain and E2
                tc_fetch32(a.image + (tF ? a.il.w2n[1] : a.il.w2n[0]), wreg);
                if (has_next)
                {
                    // the next work item's rows: issued under the layer-2 chain (48 instructions, ~1.5 k cycles of tensor
                    // pipe), they land while this tile computes (and during phase 2 when they belong to the next minibatch)
                    prefetch_issue();
                    staged = nstep != step;
                }
                mbar_wait(&tcbar[1], tc_parity);
                mtc::fence_after();
                CPPRL_MARK(22);
                {
                    const float bias = img[(tF ? a.il.b2[1] : a.il.b2[0]) + jF];
                    float *h2row = smem + sp.h2[0] + F * LD + ch * CH; // (the two towers' h2 tiles are contiguous)
#pragma unroll 1
                    for (int c = 0; c < CH; c += 4)
                    {
                        float v[4];
                        mtc::ld4(lane_base + mtc::kColA + 64 * tF + ch * CH + c, v);
                        *reinterpret_cast<float4 *>(h2row + c) =
                            make_float4(mtc::tanh_sfu(v[0] + bias), mtc::tanh_sfu(v[1] + bias), mtc::tanh_sfu(v[2] + bias), mtc::tanh_sfu(v[3] + bias));
                    }
                }
                CPPRL_MARK(23);
                // layer 2 has consumed W2: its columns take W2^T
                tc_store32(lane_base + mtc::kColW + 32 * ch, wreg);
                mtc::wait_st();
                mtc::fence_before();
                __syncthreads();
            }
            else
            {
            // ---- P1: layer 1, both towers ------------------------------------------------------------
            {
                float acc[4][Tile<TC ? 32 : R>::NG * 4];
                gemm_cf<TC ? 32 : R>(w1t_img, X, Dp, jb, rb, acc);
                store_tanh<TC ? 32 : R>(h1t, b1_img, jb, rb, acc);
            }
            if (tid < R)
                ROWIDN[tid] = next_row; // read by prefetch_issue after the next barrier
            __syncthreads();
            CPPRL_MARK(2);
            // ---- P2: layer 2 ---------------------------------------------------------------------------
            {
                float acc[4][Tile<TC ? 32 : R>::NG * 4];
                gemm_cf<TC ? 32 : R>(w2t_img, h1t, kHid, jb, rb, acc);
                store_tanh<TC ? 32 : R>(h2t, b2_img, jb, rb, acc);
            }
            __syncthreads();
            }
            if (!TC && has_next)
            {
                prefetch_issue(); // lands during P3..P8 (and phase 2 when it belongs to the next minibatch)
                staged = nstep != step;
            }
            CPPRL_MARK(3);
            // ---- P3: heads: logits / loc [A][R] from the actor tower, value [R] from the critic tower ----
            if (TC)
            {
                // Thread = (row r, quarter kq of the 64 hidden units): its 16 activations of each tower stay in registers
                // while the head rows stream past as broadcast 16-byte loads (60 shared-memory loads per thread for 7
                // outputs instead of 80 per output); the four quarters meet through shared memory -- the layer-2 operand
                // planes, idle between the layer-2 chain and P5b.
                float *PART = smem + sp.stage; // [kq][A + 1][64]: at most 4 x 17 x 64 floats
                const int r = (warp & 1) * 32 + lane, kq = warp >> 1;
                if (r < R)
                {
                    const float *h2a = smem + sp.h2[0] + (16 * kq) * LD + r, *h2c = smem + sp.h2[1] + (16 * kq) * LD + r;
                    float ha[16], hc[16];
#pragma unroll
                    for (int j = 0; j < 16; ++j)
                        ha[j] = h2a[j * LD], hc[j] = h2c[j * LD];
                    auto dot16 = [](const float *w, const float (&h)[16]) {
                        float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
#pragma unroll
                        for (int j = 0; j < 16; j += 4)
                        {
                            const float4 wq = *reinterpret_cast<const float4 *>(w + j);
                            s0 = fmaf(wq.x, h[j], s0);
                            s1 = fmaf(wq.y, h[j + 1], s1);
                            s2 = fmaf(wq.z, h[j + 2], s2);
                            s3 = fmaf(wq.w, h[j + 3], s3);
                        }
                        return (s0 + s1) + (s2 + s3);
                    };
                    for (int o = 0; o < A; ++o)
                        PART[(kq * (A + 1) + o) * 64 + r] = dot16(img + a.il.wo + o * kHid + 16 * kq, ha);
                    PART[(kq * (A + 1) + A) * 64 + r] = dot16(img + a.il.wv + 16 * kq, hc);
                }
                __syncthreads();
                for (int e = tid; e < (A + 1) * R; e += kUpdThreads)
                {
                    const int o = e / R, r2 = e - o * R;
                    const float *pp = PART + o * 64 + r2;
                    const int q1 = (A + 1) * 64;
                    const float sum = ((pp[0] + pp[q1]) + (pp[2 * q1] + pp[3 * q1])) + img[o == A ? a.il.bv : a.il.bo + o];
                    if (o == A)
                        VAL[r2] = sum;
                    else
                        LOGIT[o * LD + r2] = sum;
                }
            }
            else
            for (int e = tid; e < (A + 1) * R; e += kUpdThreads)
            {
                const int o = e / R, r = e - o * R;
                const bool is_value = o == A;
                const float *w = img + (is_value ? a.il.wv : a.il.wo + o * kHid);
                const float *h = smem + (is_value ? sp.h2[1] : sp.h2[0]) + r;
                float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
#pragma unroll 4
                for (int k = 0; k < kHid; k += 4)
                {
                    const float4 wq = *reinterpret_cast<const float4 *>(w + k);
                    s0 = fmaf(wq.x, h[k * LD], s0);
                    s1 = fmaf(wq.y, h[(k + 1) * LD], s1);
                    s2 = fmaf(wq.z, h[(k + 2) * LD], s2);
                    s3 = fmaf(wq.w, h[(k + 3) * LD], s3);
                }
                const float s = (s0 + s1) + (s2 + s3) + img[is_value ? a.il.bv : a.il.bo + o];
                if (is_value)
                    VAL[r] = s;
                else
                    LOGIT[o * LD + r] = s;
            }
            __syncthreads();
            CPPRL_MARK(4);
            // ---- P4: log-prob, entropy, loss and its gradient w.r.t. the head outputs.  Four lanes per row split
            // the actions (i = q, q+4, ...) and combine with two shuffles; per-action scratch lives in the
            // LOGIT / DLOGIT / STAT columns of the row.
            if (tid < 4 * R)
            {
                const int r = tid >> 2, q = tid & 3;
                const bool valid = r < nrows;
                const bool live = training && valid;
                float newlp = 0.f, ent = 0.f;
                const float value = VAL[r];
                int act = 0;
                if (a.head == 0)
                {
                    // Categorical(logits): l = z - logsumexp(z); p = softmax(l)             (categorical.cpp:35-36)
                    act = reinterpret_cast<const int *>(ACT)[r];
                    float m = -INFINITY;
                    for (int i = q; i < A; i += 4)
                        m = fmaxf(m, LOGIT[i * LD + r]);
                    m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 1));
                    m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 2));
                    float se = 0.f;
                    for (int i = q; i < A; i += 4)
                        se += expf(LOGIT[i * LD + r] - m);
                    se += __shfl_xor_sync(0xffffffffu, se, 1);
                    se += __shfl_xor_sync(0xffffffffu, se, 2);
                    const float lse = m + logf(se);
                    for (int i = q; i < A; i += 4)
                    {
                        const float l = LOGIT[i * LD + r] - lse; // normalised log-probability
                        const float pr = expf(l);                // probability
                        ent -= l * pr;                           // entropy = -sum p log p        (categorical.cpp:48-52)
                        newlp += i == act ? l : 0.f;             // log_prob = gather(l, action)  (categorical.cpp:54-61)
                        LOGIT[i * LD + r] = l;
                        DLOGIT[i * LD + r] = pr;
                        if (!training && a.out_dist && valid)
                            a.out_dist[static_cast<size_t>(row0 + r) * A + i] = l;
                    }
                }
                else
                {
                    // Normal(loc, sigma)                                                          (normal.cpp:24-37)
                    for (int i = q; i < A; i += 4)
                    {
                        const float var = SIG[2 * i], log_sigma = SIG[2 * i + 1];
                        const float diff = ACT[i * LD + r] - LOGIT[i * LD + r];
                        if (!training && a.out_dist && valid)
                            a.out_dist[static_cast<size_t>(row0 + r) * A + i] = LOGIT[i * LD + r];
                        // (one division per action: the three quotients by var share it)
                        const float dq = diff / var;
                        newlp += -0.5f * (diff * dq) - log_sigma - 0.91893853320467274178f;
                        ent += 0.5f + 0.5f * 1.8378770664093454836f + log_sigma;
                        LOGIT[i * LD + r] = dq;                    // d logp / d loc
                        STAT[(5 + i) * LD + r] = diff * dq - 1.f;  // d logp / d scale_log
                    }
                }
                newlp += __shfl_xor_sync(0xffffffffu, newlp, 1);
                newlp += __shfl_xor_sync(0xffffffffu, newlp, 2);
                ent += __shfl_xor_sync(0xffffffffu, ent, 1);
                ent += __shfl_xor_sync(0xffffffffu, ent, 2);
                float g_lp = 0.f, dval = 0.f, st_v = 0.f, st_a = 0.f, st_kl = 0.f, st_clip = 0.f;
                if (live)
                {
                    const float ret = RET[r];
                    if (a.algo == kAlgoPPO)
                    {
                        const float olp = OLDLP[r], adv = ADV[r];
                        st_kl = olp - newlp;                                     // ppo.cpp:96
                        const float ratio = expf(newlp - olp);                   // ppo.cpp:107
                        const float s1 = ratio * adv;
                        const float s2 = fminf(fmaxf(ratio, a.clip_lo), a.clip_hi) * adv;
                        st_clip = fabsf(ratio - 1.f) > a.clip_param ? 1.f : 0.f; // ppo.cpp:116-122
                        st_a = fminf(s1, s2);                                    // action_loss = -mean(min)
                        const bool in_range = ratio >= a.clip_lo && ratio <= a.clip_hi;
                        g_lp = (in_range || s1 < s2) ? -a.c_actor * inv_b * adv * ratio : 0.f;
                        const float dv = ret - value;
                        st_v = dv * dv;                                          // value_loss = 0.5*mean
                        dval = -a.c_value * inv_b * dv;
                    }
                    else
                    {
                        const float adv = ret - value;                           // a2c.cpp:69
                        st_v = adv * adv;                                        // a2c.cpp:72
                        st_a = adv * newlp;                                      // a2c.cpp:75
                        g_lp = -inv_b * adv;
                        dval = -2.f * a.c_value * inv_b * adv;
                    }
                }
                if (a.head == 0)
                {
                    for (int i = q; i < A; i += 4)
                    {
                        const float l = LOGIT[i * LD + r], pr = DLOGIT[i * LD + r];
                        const float d = g_lp * ((i == act ? 1.f : 0.f) - pr) + ce_b * pr * (l + ent);
                        DLOGIT[i * LD + r] = live ? d : 0.f;
                    }
                }
                else
                {
                    for (int i = q; i < A; i += 4)
                    {
                        DLOGIT[i * LD + r] = live ? g_lp * LOGIT[i * LD + r] : 0.f;
                        STAT[(5 + i) * LD + r] = live ? g_lp * STAT[(5 + i) * LD + r] - ce_b : 0.f;
                    }
                }
                if (q == 0)
                {
                    DVAL[r] = dval;
                    STAT[0 * LD + r] = st_v;
                    STAT[1 * LD + r] = st_a;
                    STAT[2 * LD + r] = live ? ent : 0.f;
                    STAT[3 * LD + r] = st_kl;
                    STAT[4 * LD + r] = st_clip;
                }
                if (!training)
                {
                    if (valid && q == 0)
                    {
                        a.out_value[row0 + r] = value;
                        a.out_logp[row0 + r] = newlp;
                    }
                    const float e = warp_sum(valid && q == 0 ? ent : 0.f);
                    if (lane == 0)
                        atomicAdd(a.out_entropy_sum, e);
                }
            }
            __syncthreads();
            if (!training)
                continue;
            CPPRL_MARK(5);
            // ---- P5a: per-row terms -> gradient vector: loss sums, head biases, scale_log (one item per warp) ----
            // (tensor-core variant: runs under the data-gradient chain below; nothing before the next tile's P4 touches
            // its sources, and its destinations in the accumulator have no other writer)
            auto p5a = [&]() {
                const int n_items = 5 + A + 1 + (a.head == 1 ? A : 0);
                for (int item = wt >> 5; wt >= 0 && item < n_items; item += kWorkThreads / 32)
                {
                    const float *src;
                    int dst;
                    if (item < 5)
                        src = STAT + item * LD, dst = Pstat + item;
                    else if (item < 5 + A)
                        src = DLOGIT + (item - 5) * LD, dst = a.po.bo + (item - 5);
                    else if (item == 5 + A)
                        src = DVAL, dst = a.po.bv;
                    else
                        src = STAT + (5 + item - (6 + A)) * LD, dst = a.po.slog + (item - (6 + A));
                    float sum = 0.f;
#pragma unroll
                    for (int r = lane; r < R; r += 32)
                        sum += src[r];
                    sum = warp_sum(sum);
                    if (lane == 0)
                        gsm[dst] += sum; // each destination has exactly one owner in this phase
                }
            };
            if (!TC)
                p5a();
            // ---- P5b: head weight gradients and dz2 = (head^T dout) * (1 - h2^2), in place over H2.
            // Thread = (feature k, half of the rows); dh for its rows stays in registers while the loop over the
            // runtime number of head outputs streams dout rows (broadcast LDS) against the cached h2 row.
            // Tensor-core variant: (tower, k) is the thread's tensor-memory lane, and dz2 leaves as the two operands it
            // feeds -- the A operand of dW2 in tensor memory and row-major [row][unit] planes for the data gradient.
            {
                constexpr int NQ = R / 8; // float4 groups in this thread's half of the rows
                const int tw = TC ? tF : tower, k = TC ? jF : (u & 63), half = TC ? ch : (u >> 6);
                const int nout = tw == 0 ? A : 1;
                const float *dout = (tw == 0 ? DLOGIT : DVAL) + half * (R / 2);
                const float *wh = img + (tw == 0 ? a.il.wo : a.il.wv) + k;
                float *h2 = smem + sp.h2[0] + (tw * kHid + k) * LD + half * (R / 2);
                const int wbase = (tw == 0 ? a.po.wo : a.po.wv) + k;
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
                        const float4 d = *reinterpret_cast<const float4 *>(dout + o * LD + 4 * qd);
                        g0 = fmaf(d.x, h[qd].x, g0);
                        g1 = fmaf(d.y, h[qd].y, g1);
                        g0 = fmaf(d.z, h[qd].z, g0);
                        g1 = fmaf(d.w, h[qd].w, g1);
                        dh[qd].x = fmaf(d.x, wk, dh[qd].x);
                        dh[qd].y = fmaf(d.y, wk, dh[qd].y);
                        dh[qd].z = fmaf(d.z, wk, dh[qd].z);
                        dh[qd].w = fmaf(d.w, wk, dh[qd].w);
                    }
                    // row half 0 owns the gradient slot, half 1 its twin in G2 (no atomics: deterministic)
                    if (half == 0)
                        gsm[wbase + o * kHid] += g0 + g1;
                    else
                        (tw == 0 ? G2_WO : G2_WV)[o * kHid + k] += g0 + g1;
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
                    if (TC)
                    {
                        using GeoT = mtc::Geo<TC ? R : 8>;
                        unsigned char *db_hi = reinterpret_cast<unsigned char *>(smem + sp.hb[tw]), *db_lo = db_hi + GeoT::kPlaneB;
                        const float z[4] = {dz.x, dz.y, dz.z, dz.w};
                        float hi[4], lo[4];
                        const int r0 = half * (R / 2) + 4 * qd;
                        const uint32_t ob = (k >> 2) * GeoT::kLboB + (k & 3) * 4u;
#pragma unroll
                        for (int i = 0; i < 4; ++i)
                        {
                            hi[i] = mtc::tf32_hi(z[i]), lo[i] = z[i] - hi[i];
                            *reinterpret_cast<float *>(db_hi + 16u * (r0 + i) + ob) = hi[i]; // (layer 2 has consumed h1's planes)
                            *reinterpret_cast<float *>(db_lo + 16u * (r0 + i) + ob) = lo[i];
                        }
                        mtc::st4(lane_base + mtc::kColA + r0, hi[0], hi[1], hi[2], hi[3]);
                        mtc::st4(lane_base + mtc::kColA + 64 + r0, lo[0], lo[1], lo[2], lo[3]);
                    }
                    else
                        *reinterpret_cast<float4 *>(h2 + 4 * qd) = dz;
                }
                if (TC)
                    acc_b2 += bsum; // (a register per lane; the two halves meet when the partial is published)
                else if (half == 0)
                    gsm[po_b2 + k] += bsum;
                else
                    G2_B2[tower * kHid + k] += bsum;
            }
            if (TC)
            {
                mtc::wait_st();
                mtc::fence_async_smem();
                mtc::fence_before();
            }
            __syncthreads();
            CPPRL_MARK(6);
            if (TC)
            {
                using GeoT = mtc::Geo<TC ? R : 8>;
                constexpr int CH = GeoT::CH;
                unsigned char *hf_hi = reinterpret_cast<unsigned char *>(smem + sp.hf[tF]), *hf_lo = hf_hi + GeoT::kPlaneF;
                // ---- T3: data gradient dh1[F][r] = sum_j W2^T[F][j] dz2[r][j], one chain per tower -------------------
                if ((tid & 127) == 0)
                {
                    const int t = tid >> 7;
                    mtc::fence_after();
                    const uint32_t b_hi = smem_u32(smem + sp.hb[t]);
                    mtc::issue_3xtf32<0, kHid / 8, GeoT::kLboB>(tm + mtc::kColG + 64 * t, tm + mtc::kColW, tm + mtc::kColW + 64, b_hi,
                                                                b_hi + GeoT::kPlaneB, mtc::idesc_n(GeoT::kN));
                    mtc::commit(&tcbar[2]);
                }
                p5a();
                mbar_wait(&tcbar[2], tc_parity);
                mtc::fence_after();
                CPPRL_MARK(7);
                // dz1 = dh1 * (1 - h1^2) -> tensor memory (the A operand of dW1, over W2^T which the chain has consumed); db1
                {
                    float bsum = 0.f;
                    float va[CH];
                    mtc::ld_cols<CH>(lane_base + mtc::kColG + 64 * tF + ch * CH, va);
#pragma unroll
                    for (int c = 0; c < CH; c += 4)
                    {
                        const float v[4] = {va[c], va[c + 1], va[c + 2], va[c + 3]};
                        const int r0 = ch * CH + c;
                        const uint32_t of = 16u * jF + (r0 >> 2) * GeoT::kLboF;
                        const float4 hh = *reinterpret_cast<const float4 *>(hf_hi + of), hl = *reinterpret_cast<const float4 *>(hf_lo + of);
                        const float h1v[4] = {hh.x + hl.x, hh.y + hl.y, hh.z + hl.z, hh.w + hl.w}; // hi + lo == h1 exactly
                        float hi[4], lo[4];
#pragma unroll
                        for (int i = 0; i < 4; ++i)
                        {
                            const float dz = v[i] * (1.f - h1v[i] * h1v[i]);
                            bsum += dz;
                            hi[i] = mtc::tf32_hi(dz), lo[i] = dz - hi[i];
                        }
                        mtc::st4(lane_base + mtc::kColW + r0, hi[0], hi[1], hi[2], hi[3]);
                        mtc::st4(lane_base + mtc::kColW + 64 + r0, lo[0], lo[1], lo[2], lo[3]);
                    }
                    acc_b1 += bsum;
                }
                mtc::wait_st();
                mtc::fence_before();
                // last tile of this minibatch with the next minibatch's first tile prefetched: its rows are staged by the
                // non-issuing warps under the two weight-gradient chains below (the barrier publishes every thread's copies)
                const bool stage_ahead = staged && tile + G >= n_tiles && !a.no_stage_ahead;
                if (stage_ahead)
                    asm volatile("cp.async.wait_all;" ::: "memory");
                __syncthreads();
                CPPRL_MARK(8);
                // ---- T4: weight gradients of the tile: dW2[F][k] = sum_r dz2[F][r] h1[k][r] (one chain per tower's h1),
                //          dW1[F][k] = sum_r dz1[F][r] x[k][r].  These sums are the parameters' gradients, where
                //          cancellation makes the tensor core's truncating accumulation visible (a2c cfg2: 2e-4 of a
                //          parameter through RMSprop's normalisation), so the rows run as TWO chains of at most four
                //          full-size instructions each, drained into fp32 registers with round-to-nearest adds --------
#pragma unroll
                for (int pass = 0; pass < 2; ++pass)
                {
                    constexpr int KS = (TC ? R : 8) / 8, K0 = (KS + 1) / 2;
                    const int k_begin = pass ? K0 : 0, k_end = pass ? KS : K0;
                    if ((tid & 127) == 0 && k_begin < k_end)
                    {
                        const int t = tid >> 7;
                        mtc::fence_after();
                        const uint32_t b_hi = smem_u32(smem + sp.hf[t]), xf_hi = smem_u32(smem + sp.xf);
                        const uint32_t d2 = tm + mtc::kColG + 64 * t, a2 = tm + mtc::kColA, d1 = tm + mtc::kColD1, a1 = tm + mtc::kColW;
                        if (pass == 0)
                        {
                            mtc::issue_3xtf32<0, K0, GeoT::kLboF>(d2, a2, a2 + 64, b_hi, b_hi + GeoT::kPlaneF, mtc::idesc_n(64));
                            if (t == 0)
                                mtc::issue_3xtf32<0, K0, GeoT::kLboXF>(d1, a1, a1 + 64, xf_hi, xf_hi + GeoT::kPlaneXF, mtc::idesc_n(32));
                        }
                        else
                        {
                            mtc::issue_3xtf32<K0, KS, GeoT::kLboF>(d2, a2, a2 + 64, b_hi, b_hi + GeoT::kPlaneF, mtc::idesc_n(64));
                            if (t == 0)
                                mtc::issue_3xtf32<K0, KS, GeoT::kLboXF>(d1, a1, a1 + 64, xf_hi, xf_hi + GeoT::kPlaneXF, mtc::idesc_n(32));
                        }
                        mtc::commit(&tcbar[3 + pass]);
                    }
                    if (pass == 0 && tile + G >= n_tiles)
                    {
                        // last tile of this CTA in this minibatch, first chain in flight: the parts of the gradient partial that
                        // are already final -- the bias sums of column half 1 are handed to half 0 through shared memory, the
                        // head gradients of row half 1 are folded into the accumulator
                        if (ch == 1)
                            G2_B2[F] = acc_b2, G2_B2[2 * kHid + F] = acc_b1;
                        for (int i = wt; wt >= 0 && i < (A + 1) * kHid; i += kWorkThreads)
                        {
                            const int o = i >> 6, k = i & 63;
                            if (o < A)
                                gsm[a.po.wo + o * kHid + k] += G2_WO[o * kHid + k];
                            else
                                gsm[a.po.wv + k] += G2_WV[k];
                        }
                    }
                    if (stage_ahead && wt >= 0)
                    {
                        const int nr = min(R, rows_of(step + 1) - static_cast<int>(blockIdx.x) * R);
                        if (pass == 0)
                            stage_obs(nr, wt, kWorkThreads);
                        else
                            stage_rows(nr, wt, kWorkThreads);
                    }
                    if (k_begin < k_end)
                    {
                        mbar_wait(&tcbar[3 + pass], tc_parity);
                        mtc::fence_after();
#pragma unroll
                        for (int c = 0; c < 32; c += 4)
                        {
                            float v[4];
                            mtc::ld4(lane_base + mtc::kColG + 64 * tF + 32 * ch + c, v);
#pragma unroll
                            for (int i = 0; i < 4; ++i)
                                acc_w2[c + i] += v[i];
                        }
#pragma unroll
                        for (int c = 0; c < 16; c += 4)
                        {
                            float v[4];
                            mtc::ld4(lane_base + mtc::kColD1 + 16 * ch + c, v);
#pragma unroll
                            for (int i = 0; i < 4; ++i)
                                acc_w1[c + i] += v[i];
                        }
                        if (pass == 0)
                        {
                            mtc::fence_before();
                            __syncthreads(); // every lane has read the first chain's sums before the second overwrites them
                            CPPRL_MARK(24);
                        }
                    }
                }
                if (stage_ahead)
                    pre_staged = true;
                tc_parity ^= 1u;
                CPPRL_MARK(25);
                if (tile + G >= n_tiles)
                {
                    // last tile of this CTA in this minibatch: everything of the gradient partial that needs no barrier of
                    // its own leaves NOW, under the tile's closing barrier -- this lane's 32 columns of dW2 straight from
                    // registers, dW1 staged in the flat parameter order [tower][unit * D + k] (rows are D floats, not a
                    // multiple of 4; the data-gradient operand planes are idle)
                    float *part = a.gpart + static_cast<size_t>(blockIdx.x) * Pstride;
                    (void)part;
                    // dW2 [lane][64] staged with a row stride of 68 floats (conflict-free 16-byte stores; a lane's 128 bytes
                    // written straight to global memory would touch 32 different lines per warp instruction)
                    float *w2s = smem + sp.stage + kTcW1Stage + F * 68 + 32 * ch;
#pragma unroll
                    for (int i = 0; i < 8; ++i)
                        *reinterpret_cast<float4 *>(w2s + 4 * i) =
                            make_float4(acc_w2[TC ? 4 * i : 0], acc_w2[TC ? 4 * i + 1 : 0], acc_w2[TC ? 4 * i + 2 : 0], acc_w2[TC ? 4 * i + 3 : 0]);
                    float *w1s = smem + sp.stage;
#pragma unroll
                    for (int i = 0; i < 16; ++i)
                        if (16 * ch + i < D)
                            w1s[F * D + 16 * ch + i] = acc_w1[TC ? i : 0];
                }
                if (tile + G < n_tiles)
                {
                    // another tile of this minibatch follows: its layer 2 needs W2 back in the columns dz1 now occupies
                    tc_fetch32(a.image + (tF ? a.il.w2t[1] : a.il.w2t[0]), wreg);
                    tc_store32(lane_base + mtc::kColW + 32 * ch, wreg);
                    mtc::wait_st();
                }
                mtc::fence_before();
                __syncthreads(); // the next tile's contractions reuse every accumulator and operand plane
            }
            else
            {
            // ---- P6: dW2 += dz2 h1^T ------------------------------------------------------------------------
            grad_w2<TC ? 32 : R>(h2t, h1t, u, gsm + po_w2);
            __syncthreads();
            CPPRL_MARK(7);
            // ---- P7: dz1 = (W2^T dz2) * (1 - h1^2), in place over H1; db1 -----------------------------------
            {
                float acc[4][Tile<TC ? 32 : R>::NG * 4];
                gemm_cf<TC ? 32 : R>(w2n_img, h2t, kHid, jb, rb, acc);
#pragma unroll
                for (int q = 0; q < 4; ++q)
                {
                    float bsum = 0.f;
#pragma unroll
                    for (int g = 0; g < Tile<TC ? 32 : R>::NG; ++g)
                    {
                        float *hp = h1t + (jb + 16 * q) * LD + 32 * g + 4 * rb;
                        const float4 h = *reinterpret_cast<const float4 *>(hp);
                        float4 dz;
                        dz.x = acc[q][g * 4 + 0] * (1.f - h.x * h.x);
                        dz.y = acc[q][g * 4 + 1] * (1.f - h.y * h.y);
                        dz.z = acc[q][g * 4 + 2] * (1.f - h.z * h.z);
                        dz.w = acc[q][g * 4 + 3] * (1.f - h.w * h.w);
                        bsum += (dz.x + dz.y) + (dz.z + dz.w);
                        *reinterpret_cast<float4 *>(hp) = dz;
                    }
                    // the two row groups of this warp (lanes jb and jb + 16) first, then one slot per warp of the tower
                    bsum += __shfl_xor_sync(0xffffffffu, bsum, 16);
                    if (lane < 16)
                    {
                        const int wt = warp & 3; // warp inside the tower
                        if (wt == 0)
                            gsm[po_b1 + jb + 16 * q] += bsum;
                        else
                            G2_B1[(tower * 3 + wt - 1) * kHid + jb + 16 * q] += bsum;
                    }
                }
            }
            __syncthreads();
            CPPRL_MARK(8);
            // ---- P8: dW1 += dz1 x^T ---------------------------------------------------------------------------
            grad_w1<TC ? 32 : R>(h1t, X, u, D, Dp, gsm + po_w1);
            __syncthreads();
            }
        } // tiles

        if (!training)
            break;

        CPPRL_MARK(9);
        // ---- fold the second accumulators into the gradient in a fixed order, then publish this CTA's partial ----
        if (active)
        {
            float *part = a.gpart + static_cast<size_t>(blockIdx.x) * Pstride;
            if (TC)
            {
                // (the register-held gradients, the bias hand-over and the fold were done at the end of the last tile,
                // ahead of its closing barrier: see tc_publish_early)
                if (ch == 0)
                {
                    __stcg(part + (tF ? a.po.b2[1] : a.po.b2[0]) + jF, acc_b2 + G2_B2[F]);
                    __stcg(part + (tF ? a.po.b1[1] : a.po.b1[0]) + jF, acc_b1 + G2_B2[2 * kHid + F]);
                }
                const float *w1s = smem + sp.stage, *w2s = w1s + kTcW1Stage;
                for (int e = tid; e < 2 * kHid * kHid / 4; e += kUpdThreads) // consecutive threads, consecutive 16-byte words
                {
                    const int t = e >> 10, rem = e & 1023, j = rem >> 4, c4 = rem & 15;
                    __stcg(reinterpret_cast<float4 *>(part + (t ? a.po.w2[1] : a.po.w2[0])) + rem,
                           *reinterpret_cast<const float4 *>(w2s + (t * kHid + j) * 68 + 4 * c4));
                }
                for (int e = tid; e < 2 * kHid * D / 4; e += kUpdThreads)
                {
                    const int t = e >= kHid * D / 4, rem = e - t * (kHid * D / 4);
                    __stcg(reinterpret_cast<float4 *>(part + (t ? a.po.w1[1] : a.po.w1[0])) + rem, *reinterpret_cast<const float4 *>(w1s + 4 * e));
                }
                // value head, scale_log, action head, statistics: the shared-memory accumulators (po.wv is a multiple of 128)
                float4 *dst = reinterpret_cast<float4 *>(part + sp.g_base);
                const float4 *src = reinterpret_cast<const float4 *>(gsm + sp.g_base);
                for (int i = tid; i < (Pstride - sp.g_base) / 4; i += kUpdThreads)
                    __stcg(dst + i, src[i]);
            }
            else
            {
            for (int i = tid; i < (A + 1) * kHid; i += kUpdThreads)
            {
                const int o = i >> 6, k = i & 63;
                if (o < A)
                    gsm[a.po.wo + o * kHid + k] += G2_WO[o * kHid + k];
                else
                    gsm[a.po.wv + k] += G2_WV[k];
            }
            for (int i = tid; i < 2 * kHid; i += kUpdThreads)
            {
                const int t = i >> 6, k = i & 63;
                gsm[(t ? a.po.b2[1] : a.po.b2[0]) + k] += G2_B2[i];
                gsm[(t ? a.po.b1[1] : a.po.b1[0]) + k] += (G2_B1[(t * 3 + 0) * kHid + k] + G2_B1[(t * 3 + 1) * kHid + k]) + G2_B1[(t * 3 + 2) * kHid + k];
            }
            __syncthreads();
            float4 *dst = reinterpret_cast<float4 *>(part);
            const float4 *src = reinterpret_cast<const float4 *>(gsm);
            for (int i = tid; i < Pstride / 4; i += kUpdThreads)
                __stcg(dst + i, src[i]);
            }
        }
        CPPRL_MARK(10);
        grid_arrive(gbar);
        CPPRL_MARK(30);
        if (tid == 0 && a.opt == kOptAdam && applied >= kBcSteps)
        {
            // (launches of more than kBcSteps optimiser steps: the later steps' corrections are formed here)
            const double t = static_cast<double>(opt_step0 + applied + 1);
            const double bc1 = 1.0 - pow(static_cast<double>(a.beta1), t);
            const double bc2 = 1.0 - pow(static_cast<double>(a.beta2), t);
            MISC[16] = static_cast<float>(a.lr / bc1);
            MISC[17] = static_cast<float>(sqrt(bc2));
        }
        // The rollout does not depend on the parameters: while the slowest CTA is still on its way to the barrier, the
        // rows of this CTA's first tile of the NEXT minibatch (prefetched during this one) are normalised and laid out
        // as operands -- work that would otherwise sit on the critical path right after the optimiser step.
        if (staged && step + 1 < a.n_steps && static_cast<long long>(blockIdx.x) * R < rows_of(step + 1))
        {
            finalize_stage(min(R, rows_of(step + 1) - static_cast<int>(blockIdx.x) * R));
            fin_ahead = true;
        }
        CPPRL_MARK(31);
        bar_target += G;
        grid_wait(gbar, bar_target);
        CPPRL_MARK(11);

        // ---- phase 2: fixed-order reduction of the partials for this CTA's slice + the statistics slots ---------
        // Granule = 4 consecutive floats.  Warp w sums the partials of CTAs w, w+8, ...; all of a lane's loads are
        // issued before the first add, so the phase costs one L2 round trip.
        const int red_ld = a.per_cta + kNumStats;
        // (multi-GPU: the error word the exchange below looks at is fetched here, so that its L2 round trip runs under the
        // local reduction instead of between the push and the first poll of every step)
        int comm_error_seen = 0;
        if (world > 1)
            comm_error_seen = *reinterpret_cast<volatile int *>(&a.state->comm_error);
        {
            const int own_gran = n_own >> 2;
            const int n_gran = own_gran + (kNumStats >> 2);
            for (int gi = lane; gi < n_gran; gi += 32)
            {
                const int off = gi < own_gran ? c0 + 4 * gi : Pstat + 4 * (gi - own_gran);
                const float4 *src = reinterpret_cast<const float4 *>(a.gpart + off);
                float4 v[kRedMaxPer];
#pragma unroll
                for (int q = 0; q < kRedMaxPer; ++q)
                {
                    const int c = warp + kRedSplit * q;
                    v[q] = c < n_act ? __ldcg(src + static_cast<size_t>(c) * (Pstride >> 2)) : make_float4(0.f, 0.f, 0.f, 0.f);
                }
                float4 s = v[0];
#pragma unroll
                for (int q = 1; q < kRedMaxPer; ++q)
                {
                    s.x += v[q].x;
                    s.y += v[q].y;
                    s.z += v[q].z;
                    s.w += v[q].w;
                }
                *reinterpret_cast<float4 *>(RED + warp * red_ld + 4 * gi) = s;
            }
        }
        __syncthreads();
        // fold the split rows into row 0: RED[i] = gradient of owned parameter i (i < n_own) / statistic i - n_own
        for (int i = tid; i < n_own + kNumStats; i += kUpdThreads)
        {
            float g = 0.f;
#pragma unroll
            for (int w = 0; w < kRedSplit; ++w)
                g += RED[w * red_ld + i];
            RED[i] = g;
        }
        __syncthreads();
        if (world > 1)
        {
            // ---- gradient all-reduce across the GPUs of the box, inside the kernel, over NVLink peer memory ------
            // CTA b of every rank owns the same slice of the parameter vector: it pushes its rank-local sums into the
            // exchange area of every peer as 8-byte {value, sequence} words (no fence, no flag), polls its own area for
            // the peers' words and adds the R records in rank order, so all ranks obtain bit-identical gradients and
            // take the same optimiser step (and the same KL decision).
            const unsigned int seq = a.xseq0 + static_cast<unsigned int>(step) + 1u;
            // word i of CTA b's record from sender r lives at xbuf[parity][r][b * (per_cta + stats) + i] of the
            // RECEIVER's area: senders push with one-way NVLink stores, receivers poll their own HBM
            const int cta_off = static_cast<int>(blockIdx.x) * (a.per_cta + kNumStats);
            const size_t par_off = static_cast<size_t>(seq & 1u) * world * a.comm.x_stride;
            CPPRL_MARK(16);
            for (int i = tid; i < n_own + kNumStats; i += kUpdThreads)
            {
                const float mine = RED[i];
#pragma unroll
                for (int r = 0; r < kMaxRanks; ++r)
                    if (r < world && r != my_rank)
                        ll_store(exchange_xbuf(a.comm.area[r]) + par_off + static_cast<size_t>(my_rank) * a.comm.x_stride + cta_off + i,
                                 mine, seq);
            }
            CPPRL_MARK(17);
            const bool comm_ok = comm_error_seen == 0;
            for (int i = tid; i < n_own + kNumStats; i += kUpdThreads)
            {
                // the words of ALL peers are polled together (independent loads in flight: one L2 round trip per sweep
                // instead of one per peer -- 7 dependent round trips per step at 8 GPUs), then added in rank order
                const LLWord *base = exchange_xbuf(a.comm.area[my_rank]) + par_off + cta_off + i;
                float v[kMaxRanks];
                unsigned int pending = 0;
#pragma unroll
                for (int r = 0; r < kMaxRanks; ++r)
                {
                    v[r] = 0.f;
                    if (r < world && r != my_rank && comm_ok)
                        pending |= 1u << r;
                }
                for (long long spins = 0; pending && spins < (1ll << 22); ++spins)
                {
                    unsigned int w[kMaxRanks], sq[kMaxRanks];
#pragma unroll
                    for (int r = 0; r < kMaxRanks; ++r)
                        if (pending >> r & 1u)
                            asm volatile("ld.relaxed.sys.global.v2.b32 {%0, %1}, [%2];"
                                         : "=r"(w[r]), "=r"(sq[r])
                                         : "l"(base + static_cast<size_t>(r) * a.comm.x_stride)
                                         : "memory");
#pragma unroll
                    for (int r = 0; r < kMaxRanks; ++r)
                        if ((pending >> r & 1u) && sq[r] == seq)
                        {
                            v[r] = __uint_as_float(w[r]);
                            pending &= ~(1u << r);
                        }
                    if (pending && spins > 8192)
                        __nanosleep(200);
                }
                if (pending)
                    a.state->comm_error = 1; // a peer never published its record (bounded wait expired)
                float g = 0.f;
#pragma unroll
                for (int r = 0; r < kMaxRanks; ++r) // rank order: bit-identical sums on every rank
                    if (r < world)
                        g += r == my_rank ? RED[i] : v[r];
                RED[i] = g;
            }
            CPPRL_MARK(18);
            __syncthreads();
            CPPRL_MARK(19);
        }
        float stat_sum[5];
#pragma unroll
        for (int s = 0; s < 5; ++s)
            stat_sum[s] = RED[n_own + s];
        const float sum_v = stat_sum[0], sum_a = stat_sum[1], sum_e = stat_sum[2], sum_kl = stat_sum[3],
                    sum_clip = stat_sum[4];
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
            {
                asm volatile("cp.async.wait_all;" ::: "memory");
                write_back_state();
                leave();
                return; // identical decision in every CTA (same data, same summation order)
            }
        }
        else if (blockIdx.x == 0 && tid == 0)
        {
            UpdateState *st = a.state;
            st->a2c_value_loss = sum_v * inv_b;
            st->a2c_action_loss = -(sum_a * inv_b);
            st->a2c_entropy = sum_e * inv_b;
            st->num_updates += 1;
        }
        CPPRL_MARK(12);
        {
            const float step_size = applied < kBcSteps ? BC[2 * applied] : MISC[16], bc2_sqrt = applied < kBcSteps ? BC[2 * applied + 1] : MISC[17];
            const float lr = static_cast<float>(a.lr);
            for (int i = tid; i < n_own; i += kUpdThreads)
            {
                const int gi = c0 + i;
                if (gi >= P)
                    continue;
                const float g = RED[i];
                float p = OWN_P[i];
                if (a.opt == kOptAdam)
                {
                    const float m = OWN_M[i] * a.beta1 + g * (1.f - a.beta1);
                    const float v = OWN_V[i] * a.beta2 + (g * g) * (1.f - a.beta2);
                    OWN_M[i] = m;
                    OWN_V[i] = v;
                    p -= step_size * (m / (sqrtf(v) / bc2_sqrt + a.eps));
                }
                else
                {
                    const float s = OWN_V[i] * a.alpha + (g * g) * (1.f - a.alpha);
                    OWN_V[i] = s;
                    p -= lr * (g / (sqrtf(s) + a.eps));
                }
                OWN_P[i] = p;
                a.params[gi] = p;
                image_scatter(a, a.image, gi, p);
            }
        }
        ++applied;
        CPPRL_MARK(13);
        if (step + 1 < a.n_steps)
        {
            grid_arrive(gbar);
            CPPRL_MARK(14);
            bar_target += G;
            grid_wait(gbar, bar_target);
            CPPRL_MARK(15);
        }
    } // steps
    if (training)
        write_back_state();
    leave();
    if (a.dbg_clocks && tid == 0)
    {
        a.dbg_clocks[blockIdx.x * 32 + 29] = clock64(); // kernel exit
        if (blockIdx.x == 0)
        {
            unsigned long long ns;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(ns));
            a.dbg_clocks[148 * 32 + 2 * (a.dbg_launch & 63) + 1] = static_cast<long long>(ns);
        }
    }
}

template <int R, bool TC>
int launch_update(const MlpUpdateArgs &args, int grid, size_t smem_bytes, cudaStream_t s)
{
    static std::atomic<uint64_t> configured{0};
    if (!configured_on_device(configured))
    {
        CPPRL_CUDA_TRY(cudaFuncSetAttribute(mlp_update_kernel<R, TC>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        mark_configured_on_device(configured);
    }
    // (the grid barrier counts arrivals from zero in every launch: the previous launch zeroed this launch's counter)
    void *kargs[] = {const_cast<MlpUpdateArgs *>(&args)};
    CPPRL_CUDA_TRY(cudaLaunchCooperativeKernel(reinterpret_cast<const void *>(mlp_update_kernel<R, TC>), dim3(grid),
                                               dim3(kUpdThreads), kargs, smem_bytes, s));
    count_launch();
    return 0;
}
} // namespace

// ---- host-side helpers used by cabi_update.cu ------------------------------------------------------------
size_t mlp_update_smem_bytes(int R, int img_floats, int Pstride, int Dp, int Ap, int per_cta, bool tc, int img_tail, int g_tail)
{
    return sizeof(float) * static_cast<size_t>(make_plan(R, img_floats, Pstride, Dp, Ap, per_cta, tc, img_tail, g_tail).total);
}

int mlp_update_launch(int R, const MlpUpdateArgs &args, int grid, cudaStream_t s)
{
    const bool tc = args.tc != 0;
    const size_t smem = mlp_update_smem_bytes(R, args.il.floats, args.po.Pstride, args.Dp, args.Ap, args.per_cta, tc, args.il.tail, args.po.wv);
    if (smem > 227 * 1024)
        return fail(CPPRL_B200_EUNSUPPORTED, "mlp_update: shared-memory plan exceeds 227 KB for this obs/action size");
    if (tc != (args.il.tc != 0))
        return fail(CPPRL_B200_EINVAL, "mlp_update: image layout does not match the kernel variant");
    if (tc)
    {
        if ((R != kTcRows && R != kTcRowsWide && R != kTcRowsSmall && R != kTcRowsMid) || args.Dp > mtc::kMaxDp || args.algo == kAlgoForward)
            return fail(CPPRL_B200_EINVAL, "mlp_update: unsupported tensor-core plan");
        if (R == kTcRowsSmall)
            return launch_update<kTcRowsSmall, true>(args, grid, smem, s);
        if (R == kTcRowsMid)
            return launch_update<kTcRowsMid, true>(args, grid, smem, s);
        return R == kTcRows ? launch_update<kTcRows, true>(args, grid, smem, s) : launch_update<kTcRowsWide, true>(args, grid, smem, s);
    }
    if (R == 64)
        return launch_update<64, false>(args, grid, smem, s);
    if (R == 32)
        return launch_update<32, false>(args, grid, smem, s);
    return fail(CPPRL_B200_EINVAL, "mlp_update: unsupported tile size");
}
} // namespace cpprl_b200
