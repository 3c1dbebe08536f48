// mlp_update.cuh — argument block shared by the fused MLP update kernel and its host launcher.
#pragma once
#include <stdint.h>

#include "comm.cuh"

namespace cpprl_b200
{
constexpr int kHid = 64;           // MlpBase hidden size this kernel is specialised for
constexpr int kUpdThreads = 256;   // 8 warps: warps 0-3 run the actor tower, 4-7 the critic tower
constexpr int kNumStats = 8;       // slots appended to the gradient vector (5 used)
constexpr int kMaxActions = 16;
constexpr int kTcRows = 56;        // tile rows of the tensor-core variant: 148 CTAs x 56 rows cover an 8192-row minibatch in one wave
constexpr int kTcRowsMid = 32;     // ... for minibatches of 16 to 32 rows per SM
constexpr int kTcRowsSmall = 16;   // ... and for minibatches that do not fill the chip (cfg1: 256 rows = 16 tiles instead of 8 CUDA-core tiles of 32)
constexpr int kTcRowsWide = 64;    // ... and the tile for minibatches that need fewer waves with it (ragged env-sharded minibatches: 8192 +- 4 sigma rows)

enum { kAlgoPPO = 0, kAlgoA2C = 1, kAlgoForward = 2 };
enum { kOptAdam = 0, kOptRmsprop = 1 };

// Offsets (in floats) of each tensor inside the flat parameter buffer (reference registration order).
struct ParamOffsets
{
    int w1[2], b1[2], w2[2], b2[2]; // [0] actor, [1] critic
    int wv, bv;                     // critic_linear
    int slog;                       // Normal scale_log (-1 for Categorical)
    int wo, bo;                     // head linear
    int P;                          // total parameter count
    int Pstride;                    // P rounded up to 4, + kNumStats
};

// Offsets (in floats) inside the shared-memory weight image (also kept as a global scratch copy).
struct ImageLayout
{
    int w1t[2]; // [Dp][64]  contraction-major, output index permuted
    int b1[2];  // [64] permuted
    int w2t[2]; // [64][64]  [k][perm(j)]   forward
    int w2n[2]; // [64][64]  [j][perm(k)]   backward (dH1)
    int b2[2];  // [64] permuted
    int wo;     // [Ap][64]
    int bo;     // [Ap]
    int wv;     // [64]
    int bv;     // [4]
    int slog;   // [Ap]
    int floats; // total, multiple of 4
    // Tensor-core variant (tc != 0): every matrix row goes to ONE tensor-memory lane, read by one thread with 16-byte
    // loads; the image therefore interleaves the rows chunk-wise, [k / 4][unit][k % 4], so that the 32 lanes of a warp
    // load 512 contiguous bytes -- w1t[t] = W1 (Dp / 4 chunks), w2t[t] = W2 (row = unit j), w2n[t] = W2^T (row = input
    // k) -- the biases are unpermuted, and everything from b1[0] on (biases, heads: `tail`) is what the kernel keeps
    // in shared memory.
    int tc;
    int tail;
};

// Device-resident running state of one update (reset by begin, read back by end).
struct UpdateState
{
    int stopped;          // KL early stop triggered (ppo.cpp:100-104)
    int comm_error;       // a peer never published its exchange record (bounded wait expired)
    int num_updates;      // optimizer steps applied in this update
    float total_value_loss, total_action_loss, total_entropy, total_clip_fraction;
    float last_kl;
    float kl_stop_update; // num_updates at the stop, -1 otherwise
    long long opt_step;   // Adam step count (persists across updates)
    float a2c_value_loss, a2c_action_loss, a2c_entropy;
};

struct MlpUpdateArgs
{
    // policy
    float *params;        // flat parameters (updated in place)
    float *image;         // global copy of the shared-memory weight image
    ParamOffsets po;
    ImageLayout il;
    int D, Dp, A, Ap;     // obs dim, padded to 8; actions, padded to 4
    int head;             // 0 categorical, 1 normal
    const float *norm_mean, *norm_var, *norm_clip; // nullable (no observation normaliser)
    // rollout (row id = t*N + env)
    const float *obs;     // [rows(+N), D]
    const void *actions;  // int64 [rows] or fp32 [rows, A]
    const float *value_preds, *returns, *old_logp;
    long long rows;       // T*N
    // minibatches
    const long long *perm; // row ids for this launch, nullable = identity
    int B;                // rows per minibatch
    int n_steps;          // minibatches in this launch
    const int *step_rows; // nullable: rows of each minibatch owned by THIS rank (env-sharded update); else B for all
    long long perm_stride; // distance between the row lists of consecutive minibatches inside `perm`
    int B_norm;           // rows of the GLOBAL minibatch: every mean of the loss divides by this
    long long rows_norm;  // rows of the GLOBAL rollout (advantage normalisation)
    CommView comm;        // world == 1: single GPU
    unsigned int xseq0;   // gradient-exchange sequence number before the first minibatch of this launch
    unsigned int adv_seq; // advantage-moment exchange sequence number of this update
    int n_active;         // CTAs that own tiles and publish a gradient partial (= min(#tiles, grid))
    int per_cta;          // parameters owned by each CTA in phase 2 (multiple of 4; grid * per_cta >= P rounded to 4)
    int tc;               // tensor-core variant of the kernel (large minibatches, obs <= 32 columns): see mlp_tc.cuh
    const double *adv_sums; // {sum, sumsq} of raw advantages over all rows (PPO), nullable
    // loss
    int algo;
    float clip_param, clip_lo, clip_hi;
    float c_actor, c_value, c_entropy;
    double kl_limit;      // 1.5 * kl_target
    // optimiser
    int opt;
    float *opt_m, *opt_v; // Adam exp_avg / exp_avg_sq, RMSprop square_avg in opt_v
    double lr;
    float beta1, beta2, eps, alpha;
    // scratch / state
    float *gpart;         // [grid][Pstride] per-CTA gradient partials
    unsigned int *barrier; // two grid-barrier counters, 8 words apart: a launch counts on [8 * barrier_sel] and zeroes the other one
    int barrier_sel;       // for the NEXT launch (no memset between the cooperative launches of an update)
    UpdateState *state;
    long long *dbg_clocks; // nullable: [grid][32] clock64() phase markers of step dbg_step (profiling hook)
    int dbg_step;
    int no_stage_ahead;    // A/B switch (CPPRL_B200_MLP_STAGE_AHEAD=0): the next minibatch's first tile is staged at the grid barrier only
    int dbg_launch;        // launch counter: with dbg_clocks, CTA 0 stores %globaltimer at entry / exit in [148 * 32 + 2 * (dbg_launch % 64)]
    // forward-only outputs (kAlgoForward)
    float *out_value, *out_logp, *out_entropy_sum;
    float *out_dist; // nullable [B][A]: normalised log-probabilities (Categorical) / means (Normal): what act() samples from
};
} // namespace cpprl_b200
