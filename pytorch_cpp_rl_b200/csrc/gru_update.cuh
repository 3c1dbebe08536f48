// gru_update.cuh — argument block of the fused update / forward kernel for the GRU-fronted MlpBase.
#pragma once
#include <stdint.h>

#include "mlp_update.cuh"

namespace cpprl_b200
{
constexpr int kGruThreads = 256;
constexpr int kGruGates = 3 * kHid; // r, z, n rows of weight_ih_l0 / weight_hh_l0
constexpr int kGruMaxObs = 32;      // observation columns the kernel stages per step
constexpr int kGruTapes = 6;        // per sequence scratch: h, r, z, n, W_hn h + b_hn, dL/dh (from the towers)

struct GruParamOffsets
{
    int wih, whh, bih, bhh;          // gru.weight_ih_l0 [3H, D], weight_hh_l0 [3H, H], bias_ih_l0, bias_hh_l0
    int w1[2], b1[2], w2[2], b2[2];  // towers ([0] actor, [1] critic); first layers take H inputs
    int wv, bv, slog, wo, bo;        // critic_linear, Normal scale_log (-1 for Categorical), head linear
    int P;
};

struct GruUpdateArgs
{
    float *params;
    GruParamOffsets po;
    int D, Dp, A, Ap, head;
    const float *norm_mean, *norm_var, *norm_clip; // nullable (no observation normaliser)
    // sequences: row (t, agent) of every per-row tensor is t * N + agent
    const float *obs;        // [T(+1), N, D]
    const float *hxs;        // [N, H]: initial hidden state of each agent (hidden_states[0] in training)
    const float *masks;      // [T(+1), N]
    const void *actions;     // int64 [T, N] or fp32 [T, N, A]
    const float *value_preds, *returns, *old_logp;
    int T, N;
    int cut_over_all_agents; // 1: a run ends where ANY agent's mask is 0 (forward_gru sequence mode, nn_base.cpp:72-78);
                             // 0: only this sequence's own mask (one environment per minibatch)
    const long long *perm;   // training: environment of minibatch k (recurrent_generator.cpp:38,58)
    int n_mb;                // training: minibatches in this launch; forward: unused (one CTA per agent)
    const double *adv_sums;
    long long rows;          // T * N
    int algo;                // kAlgoPPO | kAlgoA2C | kAlgoForward
    float *grad_out;         // A2C: [P] gradient accumulator (zeroed by the host), summed over the CTAs with atomics
    float clip_param, clip_lo, clip_hi, c_actor, c_value, c_entropy;
    double kl_limit;
    float *opt_m, *opt_v;
    double lr;
    float beta1, beta2, eps;
    float *tape;             // [grid][kGruTapes][T][H]
    UpdateState *state;
    float *out_value, *out_logp, *out_entropy_sum, *out_hxs;
    float *out_dist; // nullable [rows][A]: normalised log-probabilities / means (see mlp_update.cuh)
    long long *dbg_clocks;   // nullable: clock64() phase markers of minibatch dbg_step (profiling hook)
    int dbg_step;
    int step_mode;           // forward: write the new hidden state to out_hxs (x rows == hxs rows)
};
size_t gru_update_smem_bytes(int Dp, int A, int Ap, int head, int T);
// grid: PPO 1 (the epoch's minibatches are a serial chain), A2C / forward one CTA per environment
int launch_gru_update(const GruUpdateArgs &args, int grid, cudaStream_t s);
} // namespace cpprl_b200
