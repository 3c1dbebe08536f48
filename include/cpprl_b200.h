/* cpprl_b200.h — C ABI of the B200-native on-policy update hot path.
 *
 * This is the drop-in boundary.  Plain pointers and sizes only: no torch types,
 * no C++ types, no exceptions.  Every pointer named `d_*` or documented "device" is a CUDA device
 * pointer on the current device; `stream` is a `cudaStream_t` passed as `void *` (NULL = legacy
 * default stream).  Every function returns 0 on success or a CPPRL_B200_E* code and leaves a
 * message retrievable with cpprl_b200_last_error() (thread-local).  There is NO CPU fallback:
 * with no usable sm_100 device every compute entry point returns CPPRL_B200_ECUDA.
 *
 * The C++ host shim (include/cpprl/ *.h, namespace cpprl) keeps the reference's class API and
 * forwards here; the "replaces" notes cite the reference implementation (paths relative to the
 * reference repository root) that each entry point stands in for.
 *
 * Layout conventions (identical to the reference's RolloutStorage, src/storage.cpp:23-48):
 * all rollout tensors are time-major contiguous [T or T+1, N, ...] fp32, Discrete actions int64.
 * A "row" is one transition; its flat id is t*N + env (src/generators/feed_forward_generator.cpp:51-76).
 */
#ifndef CPPRL_B200_H
#define CPPRL_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CPPRL_B200_ABI_VERSION 1

enum cpprl_b200_status
{
    CPPRL_B200_OK = 0,
    CPPRL_B200_EINVAL = 1,       /* bad argument (null pointer, non-positive size, ...) */
    CPPRL_B200_ECUDA = 2,        /* CUDA runtime error / no sm_100 device */
    CPPRL_B200_EUNSUPPORTED = 3, /* configuration not implemented on the GPU path (fails loudly) */
    CPPRL_B200_ENCCL = 4,        /* communicator error */
    CPPRL_B200_ESTATE = 5        /* call sequence error (e.g. epoch() before begin()) */
};

enum cpprl_b200_base_kind
{
    CPPRL_B200_BASE_MLP = 0,     /* MlpBase, non recurrent  (src/model/mlp_base.cpp:9-64) */
    CPPRL_B200_BASE_GRU_MLP = 1, /* MlpBase with GRU front  (src/model/nn_base.cpp:47-114) */
    CPPRL_B200_BASE_CNN = 2      /* CnnBase                 (src/model/cnn_base.cpp:9-47) */
};

enum cpprl_b200_head_kind
{
    CPPRL_B200_HEAD_CATEGORICAL = 0, /* CategoricalOutput + Categorical (src/model/output_layers.cpp:31-43) */
    CPPRL_B200_HEAD_NORMAL = 1       /* NormalOutput + Normal           (src/model/output_layers.cpp:45-59) */
};

/* ---- library / error ------------------------------------------------------------------- */

int cpprl_b200_abi_version(void);
/* Thread-local text of the last failure on this thread ("" if none). */
const char *cpprl_b200_last_error(void);
/* Number of kernels this library has launched in this process (all threads). */
uint64_t cpprl_b200_launch_count(void);
/* 0 if the current device is usable (compute capability 10.x); fills sm_count / name if non-NULL. */
int cpprl_b200_device_info(int *sm_count, int *cc_major, int *cc_minor, char *name, int name_len);

/* ---- (a) returns / GAE ------------------------------------------------------------------
 * replaces RolloutStorage::compute_returns, src/storage.cpp:113-144.
 *   use_gae != 0: value_preds[T] <- next_value; backward scan
 *        delta = r[t] + gamma*V[t+1]*m[t+1] - V[t];  gae = delta + (gamma*tau)*m[t+1]*gae;
 *        returns[t] = gae + V[t]            (returns[T] is left untouched, as the reference does)
 *   use_gae == 0: returns[T] <- next_value; returns[t] = returns[t+1]*gamma*m[t+1] + r[t]
 * rewards [T,N], value_preds [T+1,N] (row T written in GAE mode), masks [T+1,N], next_value [N],
 * returns [T+1,N].  One thread per environment (4 with 16-byte accesses when N % 4 == 0),
 * backward in time with the loads of several steps in flight; no communication between envs. */
int cpprl_b200_compute_returns(const float *d_rewards, float *d_value_preds, const float *d_masks,
                               const float *d_next_value, float *d_returns, int64_t T, int64_t N,
                               int use_gae, float gamma, float tau, void *stream);

/* ---- advantage = returns[:T] - value_preds[:T], normalised -------------------------------
 * replaces the prologue of PPO::update, src/algorithms/ppo.cpp:51-57:
 *   adv = (adv - mean(adv)) / (std(adv) + 1e-5), std with Bessel's correction over all T*N values.
 * d_moments (device, 4 doubles, zeroed by the call) receives {sum, sum of squares, mean, std}.
 * d_advantages [T,N] receives the normalised advantages (normalize != 0) or the raw ones. */
int cpprl_b200_advantages(const float *d_returns, const float *d_value_preds, int64_t T, int64_t N,
                          int normalize, float *d_advantages, double *d_moments, void *stream);

/* The same computation WITHOUT any allocation (cpprl_b200_advantages takes a stream-ordered temporary for its per-CTA
 * partial sums).  No counterpart in the reference: added for callers that cannot
 * call cudaMalloc -- code under CUDA graph capture, tight serving loops, harnesses that forbid it.  The scratch is a
 * file-scope device array of the library (32 KB, held by every GPU that loads the library until the process exits).
 * Limits and rules: T * N <= CPPRL_B200_ADVANTAGES_NOALLOC_MAX_ROWS, else CPPRL_B200_EUNSUPPORTED; all calls share the
 * one array, so calls that overlap on a GPU (different streams) race -- serialise them; results are bit-identical to
 * cpprl_b200_advantages (same kernels, same launch shape). */
#define CPPRL_B200_ADVANTAGES_NOALLOC_MAX_ROWS (1ll << 22)
int cpprl_b200_advantages_noalloc(const float *d_returns, const float *d_value_preds, int64_t T, int64_t N,
                                  int normalize, float *d_advantages, double *d_moments, void *stream);

/* ---- (b) minibatch gather ----------------------------------------------------------------
 * replaces the eight `.index(indices[k])` gathers of FeedForwardGenerator::next,
 * src/generators/feed_forward_generator.cpp:42-80, for ONE field:
 *   dst[i, :] = src[idx[i], :]   for i < n_idx, rows of row_bytes bytes (any dtype; bit-exact copy).
 * idx is int64 (device), values in [0, n_src_rows).  row_bytes must be a multiple of 4. */
int cpprl_b200_gather_rows(const void *d_src, int64_t n_src_rows, int64_t row_bytes, const int64_t *d_idx,
                           int64_t n_idx, void *d_dst, void *stream);

/* ---- (e) observation normaliser ------------------------------------------------------------
 * replaces RunningMeanStdImpl::update / update_from_moments, src/running_mean_std.cpp:20-43:
 * batch mean and POPULATION variance of x [M,D] per column, then the Chan merge into
 * (count[1], mean[D], var[D]) (all fp32 device buffers, updated in place).
 * d_workspace: device scratch of cpprl_b200_rms_workspace_bytes(M, D) bytes. */
int64_t cpprl_b200_rms_workspace_bytes(int64_t M, int64_t D);
int cpprl_b200_rms_update(const float *d_x, int64_t M, int64_t D, float *d_count, float *d_mean,
                          float *d_var, void *d_workspace, void *stream);
/* replaces ObservationNormalizerImpl::process_observation, src/observation_normalizer.cpp:69-74:
 *   out = clamp((x - mean) / sqrt(var + 1e-8), -clip, +clip),  x/out [M,D]; clip read from device. */
int cpprl_b200_normalize_observations(const float *d_x, int64_t M, int64_t D, const float *d_mean,
                                      const float *d_var, const float *d_clip, float *d_out, void *stream);

/* ---- (d) optimiser steps on a flat fp32 buffer ----------------------------------------------
 * replaces torch::optim::Adam::step as constructed at src/algorithms/ppo.cpp:39-42 (betas 0.9/0.999,
 * no weight decay, no amsgrad):  m = b1*m + (1-b1)*g;  v = b2*v + (1-b2)*g*g;
 *   p -= (lr / (1-b1^t)) * m / (sqrt(v)/sqrt(1-b2^t) + eps),   t = step (1-based, host supplied). */
int cpprl_b200_adam_step(float *d_params, const float *d_grads, float *d_exp_avg, float *d_exp_avg_sq,
                         int64_t n, int64_t step, float lr, float beta1, float beta2, float eps, void *stream);
/* replaces torch::optim::RMSprop::step as constructed at src/algorithms/a2c.cpp:30-34 (momentum 0, not
 * centered):  s = alpha*s + (1-alpha)*g*g;  p -= lr * g / (sqrt(s) + eps). */
int cpprl_b200_rmsprop_step(float *d_params, const float *d_grads, float *d_square_avg, int64_t n, float lr,
                            float alpha, float eps, void *stream);

/* ---- policy description -------------------------------------------------------------------
 * Parameters live in ONE flat fp32 device buffer in the reference's module registration order
 * (checked against the oracle's named_parameters() in tests/):
 *   MLP      : actor.0.{weight[H,D],bias[H]} actor.2.{weight[H,H],bias[H]}
 *              critic.0.{weight[H,D],bias[H]} critic.2.{weight[H,H],bias[H]}
 *              critic_linear.{weight[1,H],bias[1]}  then the head
 *   GRU_MLP  : gru.{weight_ih_l0[3H,D],weight_hh_l0[3H,H],bias_ih_l0[3H],bias_hh_l0[3H]} first,
 *              towers then take H inputs
 *   head     : Categorical  linear.{weight[A,H],bias[A]}
 *              Normal       scale_log[A], linear_loc.{weight[A,H],bias[A]}
 * The host shim makes the torch modules' parameter tensors views into this buffer, so the kernels
 * update them in place. */
typedef struct cpprl_b200_policy
{
    int32_t base_kind;    /* cpprl_b200_base_kind */
    int32_t head_kind;    /* cpprl_b200_head_kind */
    int64_t obs_dim;      /* D (MLP bases) or C*H*W (CNN) */
    int64_t obs_chw[3];   /* CNN only: C, H, W */
    int64_t hidden;       /* H: 64 (MLP bases) / 512 (CNN) */
    int64_t num_actions;  /* A */
    float *d_params;      /* flat parameters, n_params floats */
    int64_t n_params;
    /* observation normaliser buffers (src/observation_normalizer.cpp:9-11, src/running_mean_std.cpp:8-11);
     * all NULL when the policy was built with normalize_observations = false */
    float *d_norm_clip;   /* [1] */
    float *d_norm_count;  /* [1] */
    float *d_norm_mean;   /* [D] */
    float *d_norm_var;    /* [D] */
} cpprl_b200_policy;

/* Number of parameters the layout above implies (0 if the description is unsupported). */
int64_t cpprl_b200_policy_num_params(const cpprl_b200_policy *policy);

/* ---- rollout description -------------------------------------------------------------------
 * Device views of the reference's RolloutStorage buffers (src/storage.cpp:15-49). */
typedef struct cpprl_b200_rollout
{
    int64_t T, N;
    int64_t hidden_dim;              /* Hs of the storage (row width of d_hidden_states) */
    int64_t action_cols;             /* 1 for Discrete, A for Box */
    const float *d_observations;     /* [T+1,N,D] */
    const float *d_hidden_states;    /* [T+1,N,Hs]   (read by recurrent policies only) */
    const void *d_actions;           /* int64 [T,N,1] (Categorical) or fp32 [T,N,A] (Normal) */
    const float *d_value_preds;      /* [T+1,N] */
    const float *d_returns;          /* [T+1,N] */
    const float *d_masks;            /* [T+1,N] */
    const float *d_action_log_probs; /* [T,N] */
} cpprl_b200_rollout;

/* ---- (c) forward evaluation ------------------------------------------------------------------
 * replaces PolicyImpl::evaluate_actions, src/model/policy.cpp:83-115 (forward only), on B explicit rows:
 * obs [B,D], masks [B] and hxs [n_hxs,H] (recurrent only; n_hxs == B selects step mode, otherwise
 * sequence mode with n_hxs agents, src/model/nn_base.cpp:51-112), actions int64 [B] or fp32 [B,A].
 * Outputs: value [B], logp [B], entropy [1] (mean over rows), hxs_out [n_hxs,H] (recurrent; may be NULL). */
int cpprl_b200_evaluate_actions(const cpprl_b200_policy *policy, const float *d_obs, const float *d_hxs,
                                int64_t n_hxs, const float *d_masks, const void *d_actions, int64_t B,
                                float *d_value, float *d_logp, float *d_entropy, float *d_hxs_out,
                                void *stream);

/* replaces PolicyImpl::act, src/model/policy.cpp:52-81: forward + Distribution::sample + log_prob on B rows (arguments
 * as evaluate_actions).  Categorical::sample (src/distributions/categorical.cpp:63-80, torch::multinomial) becomes
 * the inverse CDF of one Philox uniform per row, Normal::sample (src/distributions/normal.cpp:39-44) loc + scale * z;
 * the generator is Philox4x32-10 keyed by (seed, row) at `offset`, so (seed, offset) reproduces the actions.
 * Outputs: value [B], actions int64 [B] or fp32 [B,A], logp [B] (summed over A for Normal), hxs_out (recurrent).
 * Parity with the reference is distributional (it consumes the CPU mt19937 stream); value and logp agree with
 * evaluate_actions on the sampled actions. */
int cpprl_b200_act(const cpprl_b200_policy *policy, const float *d_obs, const float *d_hxs, int64_t n_hxs,
                   const float *d_masks, int64_t B, uint64_t seed, uint64_t offset, float *d_value, void *d_actions,
                   float *d_logp, float *d_hxs_out, void *stream);

/* ---- (c)+(d) PPO update -----------------------------------------------------------------------
 * replaces PPO::update, src/algorithms/ppo.cpp:44-178, including the generator's gathers
 * (feed_forward_generator.cpp:42-80 / recurrent_generator.cpp:47-98), Policy::evaluate_actions,
 * the loss, autograd backward, torch::optim::Adam::step and the final normaliser update.
 * The handle owns the Adam state (exp_avg, exp_avg_sq, step), which persists across updates exactly
 * as the reference's optimizer member does (ppo.h:22). */
typedef struct cpprl_b200_ppo_options
{
    float clip_param;      /* ppo.cpp:47 (multiplied by decay_level by the caller) */
    int32_t num_epoch;
    int32_t num_mini_batch;
    float actor_loss_coef, value_loss_coef, entropy_coef;
    float learning_rate;   /* already multiplied by decay_level, ppo.cpp:48 */
    float epsilon;         /* Adam eps */
    float kl_target;       /* stop before applying a minibatch whose KL > 1.5*kl_target, ppo.cpp:96-104 */
} cpprl_b200_ppo_options;

/* Statistics in the order PPO::update returns them, ppo.cpp:156-177. */
typedef struct cpprl_b200_ppo_stats
{
    float value_loss, action_loss, clip_fraction, entropy, kl_divergence;
    float kl_early_stop_update; /* -1 if the update ran to the end */
    int32_t num_updates;        /* optimizer steps applied */
} cpprl_b200_ppo_stats;

typedef struct cpprl_b200_ppo cpprl_b200_ppo; /* opaque */

int cpprl_b200_ppo_create(cpprl_b200_ppo **out, int64_t n_params);
void cpprl_b200_ppo_destroy(cpprl_b200_ppo *h);
/* Device pointers to the optimizer state (n_params floats each) and the step counter, for checkpointing. */
int cpprl_b200_ppo_state(cpprl_b200_ppo *h, float **d_exp_avg, float **d_exp_avg_sq, int64_t *step);
int cpprl_b200_ppo_set_step(cpprl_b200_ppo *h, int64_t step);

/* Whole update in one call.  h_perms: HOST int64 permutations, one per epoch, drawn by the caller exactly
 * as the reference's generators draw them (torch::randperm(T*N) per epoch for feed-forward policies,
 * feed_forward_generator.cpp:31-34; torch::randperm(N) per epoch for recurrent ones,
 * recurrent_generator.cpp:38) laid out [num_epoch, perm_len].  Synchronous at return (stats are host
 * values), one stream synchronisation in total.
 * Feed-forward minibatch k of an epoch is rows perm[k*mbs .. (k+1)*mbs), mbs = (T*N)/num_mini_batch;
 * (T*N) % mbs != 0 is EINVAL (the reference's .view throws).  Recurrent: num_mini_batch must equal N
 * (reference quirk): minibatch k is environment perm[k], all T steps. */
int cpprl_b200_ppo_update(cpprl_b200_ppo *h, const cpprl_b200_policy *policy, const cpprl_b200_rollout *rollout,
                          const cpprl_b200_ppo_options *opt, const int64_t *h_perms, int64_t perm_len,
                          cpprl_b200_ppo_stats *stats, void *stream);

/* The same update with the permutations drawn HERE from the reference's own random stream: `rng_state` is the state
 * of torch's default CPU generator (at::CPUGeneratorImpl::get_state / torch.get_rng_state(),
 * cpprl_b200_rng_state_bytes() = 5056 bytes) on entry and the state to put back (set_state) on return — exactly
 * where the reference's torch::randperm calls (feed_forward_generator.cpp:31-34, recurrent_generator.cpp:38; one per
 * epoch, none after the KL early stop of ppo.cpp:100-104) would have left it, with bit-identical permutations.  The
 * epochs' Fisher-Yates walks run on background host threads while the device works on the previous epoch. */
int cpprl_b200_ppo_update_rng(cpprl_b200_ppo *h, const cpprl_b200_policy *policy, const cpprl_b200_rollout *rollout,
                              const cpprl_b200_ppo_options *opt, void *rng_state, cpprl_b200_ppo_stats *stats, void *stream);
/* Host only: `count` back-to-back torch::randperm(n, kLong) results [count][n] from the generator state blob
 * (updated in place).  n must be below 2^32 / 20 (randperm_cpu's 32-bit branch). */
int cpprl_b200_host_randperm(void *rng_state, int64_t n, int64_t count, int64_t *out, int max_threads);
int64_t cpprl_b200_rng_state_bytes(void);

/* The same update in stages, so the host can draw epoch e+1's permutation while epoch e runs and can
 * honour the KL early stop without drawing permutations the reference would not draw:
 *   begin (advantage moments) -> epoch x num_epoch (each: async H2D of the permutation + one persistent
 *   kernel over the epoch's minibatches) -> [stopped() between epochs when kl_target can trigger] -> end
 *   (normaliser update, statistics read-back, the only mandatory synchronisation). */
int cpprl_b200_ppo_begin(cpprl_b200_ppo *h, const cpprl_b200_policy *policy, const cpprl_b200_rollout *rollout,
                         const cpprl_b200_ppo_options *opt, void *stream);
int cpprl_b200_ppo_epoch(cpprl_b200_ppo *h, const int64_t *h_perm, int64_t perm_len);
/* Same as ppo_epoch with the permutation already on the device (int64). */
int cpprl_b200_ppo_epoch_device(cpprl_b200_ppo *h, const int64_t *d_perm, int64_t perm_len);
int cpprl_b200_ppo_stopped(cpprl_b200_ppo *h, int *stopped); /* synchronises */
int cpprl_b200_ppo_end(cpprl_b200_ppo *h, cpprl_b200_ppo_stats *stats);
/* Like ppo_end but leaves the statistics on the device and does not synchronise (for CUDA-event timing);
 * fetch later with ppo_fetch_stats. */
int cpprl_b200_ppo_end_async(cpprl_b200_ppo *h);
int cpprl_b200_ppo_fetch_stats(cpprl_b200_ppo *h, cpprl_b200_ppo_stats *stats);

/* ---- A2C update --------------------------------------------------------------------------------
 * replaces A2C::update, src/algorithms/a2c.cpp:36-90 (normaliser update FIRST, one forward/backward over
 * all T*N rows, value loss mean(adv^2) without 0.5, actor_loss_coef ignored, RMSprop step). */
typedef struct cpprl_b200_a2c_options
{
    float value_loss_coef, entropy_coef;
    float learning_rate; /* already multiplied by decay_level, a2c.cpp:39 */
    float epsilon, alpha;
} cpprl_b200_a2c_options;

typedef struct cpprl_b200_a2c_stats
{
    float value_loss, action_loss, entropy; /* a2c.cpp:87-89 */
} cpprl_b200_a2c_stats;

typedef struct cpprl_b200_a2c cpprl_b200_a2c; /* opaque */

int cpprl_b200_a2c_create(cpprl_b200_a2c **out, int64_t n_params);
void cpprl_b200_a2c_destroy(cpprl_b200_a2c *h);
int cpprl_b200_a2c_state(cpprl_b200_a2c *h, float **d_square_avg);
int cpprl_b200_a2c_update(cpprl_b200_a2c *h, const cpprl_b200_policy *policy, const cpprl_b200_rollout *rollout,
                          const cpprl_b200_a2c_options *opt, cpprl_b200_a2c_stats *stats, void *stream);
int cpprl_b200_a2c_update_async(cpprl_b200_a2c *h, const cpprl_b200_policy *policy,
                                const cpprl_b200_rollout *rollout, const cpprl_b200_a2c_options *opt,
                                void *stream);
int cpprl_b200_a2c_fetch_stats(cpprl_b200_a2c *h, cpprl_b200_a2c_stats *stats);

/* ---- (e) multi-GPU: rollouts sharded by environment over the GPUs of one box -----------------------------
 * The reference has no distributed path; this is new.  One process per GPU; rank r
 * holds environments [env_offset, env_offset + N_local) of a GLOBAL rollout of global_envs environments, all T steps.
 * Semantics are those of ONE reference process over the global rollout: minibatch k of an epoch is rows
 * perm[k*mbs .. (k+1)*mbs) of the GLOBAL permutation over flat ids t*global_envs + env (every rank passes the same
 * permutation, length T*global_envs); each rank processes the ids whose environment it owns, every mean divides by
 * the global count, and the per-minibatch gradient (with the loss / KL partial sums appended) is all-reduced INSIDE
 * the fused update kernel over NVLink peer memory: CTA b of every rank publishes its slice in an exchange area
 * mapped into its peers (CUDA IPC), waits for the b-th flag of each peer and adds the records in rank order — all
 * ranks take bit-identical optimiser steps and the same KL-stop decision.  The advantage moments (ppo.cpp:57) and the
 * RunningMeanStd batch moments (running_mean_std.cpp:20-43) are merged the same way.
 * Setup: comm_create on every rank -> comm_export (64-byte handle) -> exchange the handles with any host transport
 * (torch.distributed all_gather in the Python mirror) -> comm_connect(world x 64 bytes, rank order). */
typedef struct cpprl_b200_comm cpprl_b200_comm; /* opaque */
int cpprl_b200_comm_create(cpprl_b200_comm **out, int rank, int world, int64_t max_params);
int cpprl_b200_comm_export(cpprl_b200_comm *c, void *handle_out_64_bytes);
int cpprl_b200_comm_connect(cpprl_b200_comm *c, const void *handles_world_x_64_bytes);
int cpprl_b200_comm_info(const cpprl_b200_comm *c, int *rank, int *world);
void cpprl_b200_comm_destroy(cpprl_b200_comm *c);
/* comm == NULL switches back to the single-GPU update. */
int cpprl_b200_ppo_set_comm(cpprl_b200_ppo *h, cpprl_b200_comm *comm, int64_t env_offset, int64_t global_envs);
int cpprl_b200_a2c_set_comm(cpprl_b200_a2c *h, cpprl_b200_comm *comm, int64_t global_envs);
/* RunningMeanStd update whose batch is the union of every rank's rows x [M_local, D]. */
int cpprl_b200_rms_update_sharded(cpprl_b200_comm *comm, const float *d_x, int64_t M, int64_t D, float *d_count,
                                  float *d_mean, float *d_var, void *d_workspace, void *stream);
/* Synchronises `stream` and returns CPPRL_B200_ENCCL if a sharded normaliser update on this communicator gave up
 * waiting for a peer (the statistics were left untouched in that case), 0 otherwise.  The PPO / A2C updates report
 * the same condition through their fetch_stats. */
int cpprl_b200_comm_check(cpprl_b200_comm *comm, void *stream);

/* ---- (f2) the write side of the rollout buffer ------------------------------------------------------
 * Writable description of a RolloutStorage (src/storage.cpp:15-49), every buffer time-major contiguous. */
typedef struct cpprl_b200_storage
{
    int64_t T, N;
    int64_t obs_dim;            /* product of the observation shape */
    int64_t hidden_dim;         /* Hs */
    int64_t action_cols;        /* 1 for Discrete, A for Box */
    int32_t action_elem_bytes;  /* 8 (int64, Discrete) or 4 (fp32, Box) */
    float *d_observations;      /* [T+1,N,obs_dim] */
    float *d_hidden_states;     /* [T+1,N,Hs] */
    float *d_rewards;           /* [T,N] */
    float *d_value_preds;       /* [T+1,N] */
    float *d_returns;           /* [T+1,N] */
    float *d_action_log_probs;  /* [T,N] */
    void *d_actions;            /* [T,N,action_cols] */
    float *d_masks;             /* [T+1,N] */
} cpprl_b200_storage;

/* replaces RolloutStorage::insert, src/storage.cpp:176-193 (seven copy_ calls) with ONE launch: observation [N,obs_dim],
 * hidden_state [N,Hs] and mask [N] go to time row step+1; action [N,action_cols], action_log_prob [N],
 * value_prediction [N] and reward [N] to time row step.  The caller advances step = (step + 1) % T as the
 * reference does (storage.cpp:192).  Inputs must already have the storage's dtypes (the reference's copy_ converts;
 * the host mirrors convert before calling). */
int cpprl_b200_storage_insert(const cpprl_b200_storage *st, int64_t step, const float *d_observation,
                              const float *d_hidden_state, const void *d_action, const float *d_action_log_prob,
                              const float *d_value_prediction, const float *d_reward, const float *d_mask,
                              void *stream);
/* replaces RolloutStorage::after_update, src/storage.cpp:106-111: time row 0 <- time row T of observations,
 * hidden states and masks, one launch. */
int cpprl_b200_storage_after_update(const cpprl_b200_storage *st, void *stream);
/* replaces RolloutStorage::set_first_observation, src/storage.cpp:220-223. */
int cpprl_b200_storage_set_first_observation(const cpprl_b200_storage *st, const float *d_observation, void *stream);
/* replaces the body of the concatenating constructor, src/storage.cpp:51-104 (torch::cat(parts, 1) of the eight
 * fields): writes `part` into environments [env_offset, env_offset + part->N) of `dst`, all eight fields in one
 * launch; call once per part.  Also the scatter of a global rollout into per-GPU shards, read the other way. */
int cpprl_b200_storage_concat(const cpprl_b200_storage *dst, const cpprl_b200_storage *part, int64_t env_offset,
                              void *stream);

/* ---- (f3) bootstrap value + returns ---------------------------------------------------------------------
 * replaces PolicyImpl::get_values, src/model/policy.cpp:132-144: the value head of the fused forward on B rows
 * (arguments as evaluate_actions; no actions needed).  value [B]. */
int cpprl_b200_get_values(const cpprl_b200_policy *policy, const float *d_obs, const float *d_hxs, int64_t n_hxs,
                          const float *d_masks, int64_t B, float *d_value, void *stream);
/* replaces the bootstrap sequence of the example driver, example/gym_client.cpp:266-275:
 *     next_value = policy->get_values(observations[T], hidden_states[T], masks[T]);
 *     storage.compute_returns(next_value, use_gae, gamma, tau);
 * as one call with no temporary and no host round trip: the value forward of time row T writes straight to where
 * compute_returns stores next_value (value_preds[T] with GAE, returns[T] otherwise) and the scan follows on the
 * same stream.  Results are those of the two separate calls, bit for bit. */
int cpprl_b200_bootstrap_returns(const cpprl_b200_policy *policy, const cpprl_b200_storage *st, int use_gae, float gamma,
                                 float tau, void *stream);

/* ---- timing hooks ---------------------------------------------------------------------------------
 * The fused kernels record CUDA events on the caller's stream around each launch; these return the
 * accumulated device time and launch count of the named kernel family since the last reset
 * ("update", "returns", "advantages", "rms", "gather").  Synchronises the events it reads. */
int cpprl_b200_kernel_time_ms(const char *family, double *total_ms, int64_t *launches);
void cpprl_b200_kernel_time_reset(void);
void cpprl_b200_kernel_time_enable(int on);

/* Test hook for the tensor-core (tcgen05, 3xTF32) GEMM behind the CnnBase contractions: C[M][N] = A * B with plain
 * row-major device operands A [M][K] (variant bit 0: given transposed, [K][M]) and B [K][N]; variant bits 4..: tile
 * columns (32, 64, 128).  M, N, K multiples of 4. */
int cpprl_b200_debug_tc_gemm(int variant, const float *d_a, const float *d_b, float *d_c, int M, int N, int K, void *stream);
/* Test hook for the permutation stream behind cpprl_b200_ppo_update_rng (csrc/host_rng.cu): `updates` consecutive updates
 * on ONE stream with a ring of 2 * count permutations; update u takes used[u] <= count permutations (fewer = an early
 * stop) and hands the state after its last one to the next; reseed_before[u] != 0 (may be NULL) first replaces the state
 * with reseed_state.  out receives the sum(used) permutations in order, *restarts how many updates had to restart the
 * stream instead of continuing with the permutations drawn ahead.  Host code only. */
int cpprl_b200_debug_perm_stream(void *rng_state, const void *reseed_state, int64_t n, int count, int updates, const int *used,
                                 const int *reseed_before, int64_t *out, long long *restarts);
/* Profiling hook: when d_buffer (device, grid*32 int64) is non-NULL, thread 0 of every CTA of the fused PPO
 * update kernel stores clock64() at its phase boundaries during minibatch `step` of each launch. */
void cpprl_b200_debug_phase_clocks(long long *d_buffer, int step);
/* Which variant of the fused MlpBase update a minibatch of this shape trains on: > 0 = tensor cores (tcgen05 kind::tf32 as
 * 3xTF32, at most 32 observation columns: minibatches of more than 32 rows per SM on 56- or 64-row tiles, smaller ones on
 * 32- or 16-row tiles; the value is the tile's row count), 0 = fp32 CUDA cores (more than 32 observation columns).
 * CPPRL_B200_MLP_TC=0 in the environment forces 0 (A/B runs). */
int cpprl_b200_mlp_update_variant(int64_t minibatch_rows, int64_t obs_dim, int64_t num_actions);

#ifdef __cplusplus
}
#endif
#endif /* CPPRL_B200_H */
