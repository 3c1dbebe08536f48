// cabi_update.cu — C-ABI entry points for evaluate_actions, PPO update and A2C update
// (host-side orchestration: validation, scratch, permutation staging, launches, statistics read-back).
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <utility>
#include <vector>

#include <curand_kernel.h>

#include "common.cuh"
#include "host_rng.h"
#include "cnn_update.cuh"
#include "gru_update.cuh"
#include "mlp_update.cuh"

struct cpprl_b200_comm;

namespace cpprl_b200
{
// from mlp_update.cu / returns_scan.cu
size_t mlp_update_smem_bytes(int R, int img_floats, int Pstride, int Dp, int Ap, int per_cta, bool tc, int img_tail, int g_tail);
int mlp_update_launch(int R, const MlpUpdateArgs &args, int grid, cudaStream_t s);
int launch_advantage_moments(const float *d_returns, const float *d_value_preds, int64_t n, double *d_sums, void *d_ws,
                             cudaStream_t s);
size_t advantage_workspace_bytes();
int launch_rms_update(const float *d_x, int64_t M, int64_t D, float *d_count, float *d_mean, float *d_var,
                      void *d_workspace, cpprl_b200_comm *comm, cudaStream_t s, int *d_comm_error);
int launch_shard_permutation(const int64_t *d_perm, int64_t n_mb, int64_t B, int64_t n_global, int64_t env0,
                             int64_t n_local, int64_t *d_out, int *d_counts, cudaStream_t s);
CommView comm_view(const cpprl_b200_comm *c);
unsigned int *comm_step_seq(cpprl_b200_comm *c);
unsigned int *comm_update_seq(cpprl_b200_comm *c);
double *comm_local_adv(cpprl_b200_comm *c);

namespace
{
inline int round_up(int v, int m) { return (v + m - 1) / m * m; }

// Parameter offsets of the non-recurrent MLP policy in the reference's registration order.
bool mlp_param_offsets(const cpprl_b200_policy &p, ParamOffsets &po)
{
    if (p.base_kind != CPPRL_B200_BASE_MLP || p.hidden != kHid)
        return false;
    const int D = static_cast<int>(p.obs_dim), A = static_cast<int>(p.num_actions);
    int o = 0;
    po.w1[0] = o, o += kHid * D;
    po.b1[0] = o, o += kHid;
    po.w2[0] = o, o += kHid * kHid;
    po.b2[0] = o, o += kHid;
    po.w1[1] = o, o += kHid * D;
    po.b1[1] = o, o += kHid;
    po.w2[1] = o, o += kHid * kHid;
    po.b2[1] = o, o += kHid;
    po.wv = o, o += kHid;
    po.bv = o, o += 1;
    if (p.head_kind == CPPRL_B200_HEAD_NORMAL)
        po.slog = o, o += A;
    else
        po.slog = -1;
    po.wo = o, o += A * kHid;
    po.bo = o, o += A;
    po.P = o;
    po.Pstride = round_up(o, 4) + kNumStats;
    return true;
}

// Parameter offsets of the GRU-fronted MlpBase: the GRU's four tensors are registered first (nn_base.cpp:13-21), the
// towers then take the H GRU outputs as inputs (mlp_base.cpp:14-17).
bool gru_param_offsets(const cpprl_b200_policy &p, GruParamOffsets &po)
{
    if (p.base_kind != CPPRL_B200_BASE_GRU_MLP || p.hidden != kHid)
        return false;
    const int D = static_cast<int>(p.obs_dim), A = static_cast<int>(p.num_actions);
    int o = 0;
    po.wih = o, o += kGruGates * D;
    po.whh = o, o += kGruGates * kHid;
    po.bih = o, o += kGruGates;
    po.bhh = o, o += kGruGates;
    for (int t = 0; t < 2; ++t)
    {
        po.w1[t] = o, o += kHid * kHid;
        po.b1[t] = o, o += kHid;
        po.w2[t] = o, o += kHid * kHid;
        po.b2[t] = o, o += kHid;
    }
    po.wv = o, o += kHid;
    po.bv = o, o += 1;
    if (p.head_kind == CPPRL_B200_HEAD_NORMAL)
        po.slog = o, o += A;
    else
        po.slog = -1;
    po.wo = o, o += A * kHid;
    po.bo = o, o += A;
    po.P = o;
    return true;
}

ImageLayout make_image_layout(int Dp, int Ap, bool tc = false)
{
    ImageLayout il;
    int o = 0;
    il.tc = tc ? 1 : 0;
    if (tc)
    {
        // matrices first (one row per tensor-memory lane, loaded straight from global memory), then the tail the
        // kernel keeps in shared memory
        for (int t = 0; t < 2; ++t)
            il.w1t[t] = o, o += kHid * Dp;
        for (int t = 0; t < 2; ++t)
            il.w2t[t] = o, o += kHid * kHid;
        for (int t = 0; t < 2; ++t)
            il.w2n[t] = o, o += kHid * kHid;
        il.tail = o;
        for (int t = 0; t < 2; ++t)
            il.b1[t] = o, o += kHid;
        for (int t = 0; t < 2; ++t)
            il.b2[t] = o, o += kHid;
    }
    else
    {
        il.tail = 0;
        for (int t = 0; t < 2; ++t)
        {
            il.w1t[t] = o, o += Dp * kHid;
            il.b1[t] = o, o += kHid;
            il.w2t[t] = o, o += kHid * kHid;
            il.w2n[t] = o, o += kHid * kHid;
            il.b2[t] = o, o += kHid;
        }
    }
    il.wo = o, o += Ap * kHid;
    il.bo = o, o += Ap;
    il.wv = o, o += kHid;
    il.bv = o, o += 4;
    il.slog = o, o += Ap;
    il.floats = round_up(o, 4);
    return il;
}

int validate_policy(const cpprl_b200_policy *p, const char *who)
{
    if (!p || !p->d_params)
        return fail(CPPRL_B200_EINVAL, std::string(who) + ": null policy / parameter buffer");
    if (p->base_kind == CPPRL_B200_BASE_CNN)
    {
        // CnnBase (cnn_base.cpp:9-31): the reference hard-codes the 32*7*7 flatten, i.e. 84x84 frames
        if (p->head_kind != CPPRL_B200_HEAD_CATEGORICAL && p->head_kind != CPPRL_B200_HEAD_NORMAL)
            return fail(CPPRL_B200_EUNSUPPORTED, std::string(who) + ": only Categorical and Normal heads are implemented (Bernoulli is out of scope)");
        if (p->obs_chw[1] != kCnnIn || p->obs_chw[2] != kCnnIn || p->obs_chw[0] < 1 || p->obs_chw[0] > 16)
            return fail(CPPRL_B200_EUNSUPPORTED, std::string(who) + ": CnnBase takes [C <= 16, 84, 84] observations");
        if (p->obs_dim != p->obs_chw[0] * kCnnIn * kCnnIn)
            return fail(CPPRL_B200_EINVAL, std::string(who) + ": obs_dim does not match obs_chw");
        if (p->hidden != kCnnHid)
            return fail(CPPRL_B200_EUNSUPPORTED, std::string(who) + ": CnnBase hidden size must be 512 on the GPU path");
        if (p->num_actions < 1 || p->num_actions > kMaxActions)
            return fail(CPPRL_B200_EUNSUPPORTED, std::string(who) + ": num_actions must be in [1, 16]");
        if (p->d_norm_clip || p->d_norm_count || p->d_norm_mean || p->d_norm_var)
            return fail(CPPRL_B200_EUNSUPPORTED, std::string(who) + ": observation normalisation of image observations is not implemented on the GPU path");
        const CnnParamOffsets cpo = cnn_param_offsets(static_cast<int>(p->obs_chw[0]), static_cast<int>(p->num_actions),
                                                      p->head_kind == CPPRL_B200_HEAD_NORMAL ? 1 : 0);
        if (p->n_params != cpo.P)
            return fail(CPPRL_B200_EINVAL, std::string(who) + ": n_params does not match the policy layout (expected " +
                                               std::to_string(cpo.P) + ", got " + std::to_string(p->n_params) + ")");
        return 0;
    }
    const bool gru = p->base_kind == CPPRL_B200_BASE_GRU_MLP;
    if (p->base_kind != CPPRL_B200_BASE_MLP && !gru)
        return fail(CPPRL_B200_EINVAL, std::string(who) + ": unknown base kind");
    if (p->head_kind != CPPRL_B200_HEAD_CATEGORICAL && p->head_kind != CPPRL_B200_HEAD_NORMAL)
        return fail(CPPRL_B200_EUNSUPPORTED, std::string(who) + ": only Categorical and Normal heads are implemented (Bernoulli is out of scope)");
    if (p->hidden != kHid)
        return fail(CPPRL_B200_EUNSUPPORTED, std::string(who) + ": MlpBase hidden size must be 64 on the GPU path");
    if (p->obs_dim < 1 || p->obs_dim > 256)
        return fail(CPPRL_B200_EUNSUPPORTED, std::string(who) + ": obs_dim must be in [1, 256]");
    if (gru && p->obs_dim > kGruMaxObs)
        return fail(CPPRL_B200_EUNSUPPORTED, std::string(who) + ": the recurrent (GRU) path takes at most 32 observation columns");
    if (p->num_actions < 1 || p->num_actions > kMaxActions)
        return fail(CPPRL_B200_EUNSUPPORTED, std::string(who) + ": num_actions must be in [1, 16]");
    int expect_p = 0;
    if (gru)
    {
        GruParamOffsets gpo;
        gru_param_offsets(*p, gpo);
        expect_p = gpo.P;
    }
    else
    {
        ParamOffsets po;
        mlp_param_offsets(*p, po);
        expect_p = po.P;
    }
    if (p->n_params != expect_p)
        return fail(CPPRL_B200_EINVAL, std::string(who) + ": n_params does not match the policy layout (expected " +
                                           std::to_string(expect_p) + ", got " + std::to_string(p->n_params) + ")");
    const bool any = p->d_norm_clip || p->d_norm_count || p->d_norm_mean || p->d_norm_var;
    const bool all = p->d_norm_clip && p->d_norm_count && p->d_norm_mean && p->d_norm_var;
    if (any && !all)
        return fail(CPPRL_B200_EINVAL, std::string(who) + ": normaliser buffers must be all set or all NULL");
    return 0;
}

// Scratch shared by every update flavour.
struct Scratch
{
    float *image = nullptr;
    size_t image_floats = 0;
    float *gpart = nullptr;
    size_t gpart_floats = 0;
    unsigned int *barrier = nullptr; // two counters (see MlpUpdateArgs::barrier)
    unsigned int barrier_launches = 0;
    UpdateState *state = nullptr;
    double *adv_sums = nullptr;
    void *adv_ws = nullptr; // per-CTA partials + arrival counter of the deterministic advantage moments
    void *rms_ws = nullptr;
    size_t rms_ws_bytes = 0;
    float *tape = nullptr; // GRU path: per-CTA activation tapes
    size_t tape_floats = 0;

    int ensure_tape(size_t n)
    {
        if (n > tape_floats)
        {
            if (tape)
                cudaFree(tape);
            tape = nullptr, tape_floats = 0;
            CPPRL_CUDA_TRY(cudaMalloc(&tape, n * sizeof(float)));
            tape_floats = n;
        }
        return 0;
    }

    int ensure(size_t img, size_t gp, size_t rms)
    {
        if (!barrier)
        {
            CPPRL_CUDA_TRY(cudaMalloc(&barrier, 64));
            CPPRL_CUDA_TRY(cudaMemset(barrier, 0, 64));
            CPPRL_CUDA_TRY(cudaMalloc(&state, sizeof(UpdateState)));
            CPPRL_CUDA_TRY(cudaMemset(state, 0, sizeof(UpdateState)));
            CPPRL_CUDA_TRY(cudaMalloc(&adv_sums, 4 * sizeof(double)));
            CPPRL_CUDA_TRY(cudaMalloc(&adv_ws, advantage_workspace_bytes()));
            CPPRL_CUDA_TRY(cudaMemset(adv_ws, 0, advantage_workspace_bytes()));
        }
        if (img > image_floats)
        {
            if (image)
                cudaFree(image);
            CPPRL_CUDA_TRY(cudaMalloc(&image, img * sizeof(float)));
            image_floats = img;
        }
        if (gp > gpart_floats)
        {
            if (gpart)
                cudaFree(gpart);
            CPPRL_CUDA_TRY(cudaMalloc(&gpart, gp * sizeof(float)));
            gpart_floats = gp;
        }
        if (rms > rms_ws_bytes)
        {
            if (rms_ws)
                cudaFree(rms_ws);
            CPPRL_CUDA_TRY(cudaMalloc(&rms_ws, rms));
            rms_ws_bytes = rms;
        }
        return 0;
    }
    void release()
    {
        cudaFree(image), cudaFree(gpart), cudaFree(barrier), cudaFree(state), cudaFree(adv_sums), cudaFree(adv_ws), cudaFree(rms_ws);
        cudaFree(tape);
        image = gpart = tape = nullptr;
        barrier = nullptr;
        state = nullptr;
        adv_sums = nullptr;
        adv_ws = nullptr;
        rms_ws = nullptr;
    }
};

struct Plan
{
    int R = 0, grid = 0, n_active = 0, per_cta = 0;
    bool tc = false; // tensor-core variant of the fused kernel (mlp_tc.cuh)
    ImageLayout il;
};

// CPPRL_B200_MLP_STAGE_AHEAD=0: A/B switch of the tensor-core kernel (MlpUpdateArgs::no_stage_ahead)
static int mlp_no_stage_ahead()
{
    const char *e = std::getenv("CPPRL_B200_MLP_STAGE_AHEAD");
    return (e && e[0] == '0') ? 1 : 0;
}

// CPPRL_B200_MLP_TC=0 keeps every update on the fp32 CUDA-core kernel (A/B runs)
bool mlp_tc_enabled()
{
    const char *e = std::getenv("CPPRL_B200_MLP_TC");
    return !(e && e[0] == '0');
}

// Tile rows, grid and kernel variant for a minibatch of B rows: the smallest tile that still gives every SM at most one
// tile, falling back to the smaller tile when the shared-memory plan of the larger one does not fit.  Minibatches that
// fill the chip with 64-row tiles (more than 32 rows per SM) and have at most 32 observation columns train on the
// tensor-core variant: 56-row tiles, so that 148 CTAs cover an 8192-row minibatch in one wave; minibatches of at most 16
// rows per SM train on its 16-row tile, those up to 32 rows per SM on its 32-row tile.
// Training launches keep at least 16 CTAs so that the phase-2 slice of the parameter vector each CTA reduces,
// steps and keeps resident in shared memory stays small; CTAs beyond the tile count only take part in phase 2.
// Training launches keep at least this many CTAs: CTAs beyond the tile count only take part in phase 2, where each
// reduces, steps and keeps resident in shared memory 1 / grid of the parameter vector (CPPRL_B200_MIN_GRID for A/B runs).
static int min_training_grid(int sms)
{
    static const int forced = [] {
        const char *e = std::getenv("CPPRL_B200_MIN_GRID");
        return e ? std::atoi(e) : 0;
    }();
    return std::min(forced > 0 ? forced : sms, sms); // every SM: measured on cfg1 (8 tiles per minibatch) 16 -> 148 CTAs: 0.467 -> 0.416 ms per update
}

// B_max >= B: the largest minibatch the launch may meet (env-sharded updates: minibatches are ragged around B).
int choose_plan(int64_t B, const ParamOffsets &po, int Dp, int Ap, bool training, Plan &plan, int64_t B_max = 0)
{
    const int sms = sm_count();
    B_max = std::max(B, B_max);
    auto shape = [&](int r, bool tc, Plan &pl) {
        const int64_t tiles = (B + r - 1) / r;
        pl.R = r;
        pl.tc = tc;
        pl.il = make_image_layout(Dp, Ap, tc);
        pl.n_active = static_cast<int>(tiles < sms ? tiles : sms);
        pl.grid = training ? std::max(pl.n_active, min_training_grid(sms)) : pl.n_active;
        const int p4 = po.Pstride - kNumStats; // parameter count rounded up to 4
        pl.per_cta = training ? round_up((p4 + pl.grid - 1) / pl.grid, 4) : 0;
        return mlp_update_smem_bytes(r, pl.il.floats, po.Pstride, Dp, Ap, pl.per_cta, tc, pl.il.tail, po.wv);
    };
    if (training && B > 32ll * sms && Dp <= 32 && mlp_tc_enabled())
    {
        // 56-row tiles unless 64-row tiles take fewer waves over the SMs (a tile costs about the same either way)
        auto waves = [&](int r) { return ((B_max + r - 1) / r + sms - 1) / sms; };
        const int r = waves(kTcRowsWide) < waves(kTcRows) ? kTcRowsWide : kTcRows;
        if (shape(r, true, plan) <= 227 * 1024)
            return 0;
        if (r != kTcRows && shape(kTcRows, true, plan) <= 227 * 1024)
            return 0;
    }
    // minibatches that do not fill the chip with 32-row CUDA-core tiles: 16-row tensor-core tiles put twice as many SMs to
    // work on a tile that is itself shorter (cfg1: 256 rows, 16 tiles: 0.426 -> 0.369 ms per update, 2.40 -> 2.78 M transitions/s)
    if (training && B_max <= 16ll * sms && Dp <= 32 && mlp_tc_enabled() && shape(kTcRowsSmall, true, plan) <= 227 * 1024)
        return 0;
    // ... and the sizes in between (16 to 32 rows per SM) on 32-row tensor-core tiles: one tile per SM either way, the tile
    // itself a third shorter than the CUDA-core one
    if (training && B_max <= 32ll * sms && Dp <= 32 && mlp_tc_enabled() && shape(kTcRowsMid, true, plan) <= 227 * 1024)
        return 0;
    int R = (B <= 32ll * sms) ? 32 : 64;
    if (R == 64 && shape(64, false, plan) > 227 * 1024)
        R = 32;
    if (shape(R, false, plan) > 227 * 1024)
        return fail(CPPRL_B200_EUNSUPPORTED, "update: observation / action size needs more than 227 KB of shared memory");
    return 0;
}

void fill_policy_args(MlpUpdateArgs &a, const cpprl_b200_policy &p, const ParamOffsets &po, const ImageLayout &il,
                      int Dp, int Ap)
{
    a.params = p.d_params;
    a.po = po;
    a.il = il;
    a.D = static_cast<int>(p.obs_dim);
    a.Dp = Dp;
    a.A = static_cast<int>(p.num_actions);
    a.Ap = Ap;
    a.head = p.head_kind == CPPRL_B200_HEAD_NORMAL ? 1 : 0;
    a.norm_mean = p.d_norm_mean;
    a.norm_var = p.d_norm_var;
    a.norm_clip = p.d_norm_clip;
}

int fill_gru_policy_args(GruUpdateArgs &a, const cpprl_b200_policy &p, int64_t T, const char *who)
{
    std::memset(&a, 0, sizeof(a));
    gru_param_offsets(p, a.po);
    a.params = p.d_params;
    a.D = static_cast<int>(p.obs_dim);
    a.Dp = round_up(a.D, 8);
    a.A = static_cast<int>(p.num_actions);
    a.Ap = round_up(a.A, 4);
    a.head = p.head_kind == CPPRL_B200_HEAD_NORMAL ? 1 : 0;
    a.norm_mean = p.d_norm_mean;
    a.norm_var = p.d_norm_var;
    a.norm_clip = p.d_norm_clip;
    if (T > 0x7fffffffll / (kGruGates * 8) || gru_update_smem_bytes(a.Dp, a.A, a.Ap, a.head, static_cast<int>(T)) > 227 * 1024)
        return fail(CPPRL_B200_EUNSUPPORTED, std::string(who) + ": sequence length x observation size needs more than 227 KB of shared memory on the recurrent path");
    return 0;
}

void fill_cnn_policy_args(CnnUpdateArgs &a, const cpprl_b200_policy &p)
{
    std::memset(&a, 0, sizeof(a));
    a.C = static_cast<int>(p.obs_chw[0]);
    a.A = static_cast<int>(p.num_actions);
    a.head = p.head_kind == CPPRL_B200_HEAD_NORMAL ? 1 : 0;
    a.po = cnn_param_offsets(a.C, a.A, a.head);
    a.il = cnn_image_layout(a.C);
    a.params = p.d_params;
    a.comm.world = 1;
}

// rows per pass of the CNN path: activations + their gradients take 160 KB per row
// (CPPRL_B200_CNN_CHUNK overrides the pass size: the tests use it to exercise multi-pass accumulation on small rollouts)
inline int cnn_chunk_rows(int64_t B)
{
    int64_t chunk = 4096;
    if (const char *e = std::getenv("CPPRL_B200_CNN_CHUNK"))
        chunk = std::max<int64_t>(1, std::atoll(e));
    return static_cast<int>(std::min<int64_t>(B, chunk));
}

__global__ void reset_state_kernel(UpdateState *st)
{
    const long long keep = st->opt_step;
    UpdateState z = {};
    z.opt_step = keep;
    z.kl_stop_update = -1.f;
    *st = z;
}

__global__ void set_opt_step_kernel(UpdateState *st, long long step) { st->opt_step = step; }

__global__ void finish_entropy_kernel(float *entropy, float inv_b) { entropy[0] *= inv_b; }
} // namespace
} // namespace cpprl_b200

using namespace cpprl_b200;

// =====================================================================================================
// PPO
// =====================================================================================================
struct cpprl_b200_ppo
{
    int64_t n_params = 0;
    float *d_m = nullptr, *d_v = nullptr;
    Scratch scratch;
    // permutation staging (double buffered)
    int64_t *d_perm[2] = {nullptr, nullptr};
    int64_t *h_perm[2] = {nullptr, nullptr};
    size_t perm_cap = 0;
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t copied[2] = {nullptr, nullptr};   // H2D of buffer b finished
    cudaEvent_t consumed[2] = {nullptr, nullptr}; // kernel reading buffer b finished
    bool consumed_valid[2] = {false, false};
    UpdateState *h_state = nullptr; // pinned
    // per-update context
    bool active = false;
    cudaStream_t stream = nullptr;
    cpprl_b200_policy policy;
    cpprl_b200_rollout rollout;
    cpprl_b200_ppo_options opt;
    MlpUpdateArgs args;
    GruUpdateArgs gargs; // recurrent policies
    bool gru = false;
    CnnUpdateArgs cargs; // CnnBase
    CnnWorkspace cws;
    bool cnn = false;
    Plan plan;
    int64_t mbs = 0, n_mb = 0;
    int epoch = 0;
    int64_t host_step = 0;
    // env-sharded (multi-GPU) update: this rank holds envs [env_offset, env_offset + rollout.N) of global_envs
    cpprl_b200_comm *comm = nullptr;
    int64_t env_offset = 0, global_envs = 0;
    int64_t *d_lperm = nullptr; // the global permutation compacted to this rank's rows, [n_mb][mbs]
    int *d_counts = nullptr;    // rows of each minibatch owned by this rank
    size_t lperm_cap = 0, counts_cap = 0;
    // cpprl_b200_ppo_update_rng: every epoch's permutation of the reference's CPU random stream, drawn by background
    // threads straight into pinned memory (host_rng.cu)
    int64_t *h_rng_perms = nullptr; // pinned [num_epoch][perm_len]
    size_t rng_perms_cap = 0;
    PermStream *perm_stream = nullptr;
};

extern "C" int cpprl_b200_ppo_create(cpprl_b200_ppo **out, int64_t n_params)
{
    if (!out || n_params <= 0)
        return fail(CPPRL_B200_EINVAL, "ppo_create: bad arguments");
    if (int rc = require_device())
        return rc;
    auto *h = new cpprl_b200_ppo();
    h->n_params = n_params;
    int rc = 0;
    auto attempt = [&]() -> int {
        CPPRL_CUDA_TRY(cudaMalloc(&h->d_m, n_params * sizeof(float)));
        CPPRL_CUDA_TRY(cudaMalloc(&h->d_v, n_params * sizeof(float)));
        CPPRL_CUDA_TRY(cudaMemset(h->d_m, 0, n_params * sizeof(float)));
        CPPRL_CUDA_TRY(cudaMemset(h->d_v, 0, n_params * sizeof(float)));
        CPPRL_CUDA_TRY(cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking));
        for (int b = 0; b < 2; ++b)
        {
            CPPRL_CUDA_TRY(cudaEventCreateWithFlags(&h->copied[b], cudaEventDisableTiming));
            CPPRL_CUDA_TRY(cudaEventCreateWithFlags(&h->consumed[b], cudaEventDisableTiming));
        }
        CPPRL_CUDA_TRY(cudaMallocHost(&h->h_state, sizeof(UpdateState)));
        return h->scratch.ensure(0, 0, 0);
    };
    rc = attempt();
    if (rc != 0)
    {
        cpprl_b200_ppo_destroy(h);
        return rc;
    }
    *out = h;
    return 0;
}

extern "C" void cpprl_b200_ppo_destroy(cpprl_b200_ppo *h)
{
    if (!h)
        return;
    cudaFree(h->d_m), cudaFree(h->d_v);
    cudaFree(h->d_lperm), cudaFree(h->d_counts);
    for (int b = 0; b < 2; ++b)
    {
        cudaFree(h->d_perm[b]);
        if (h->h_perm[b])
            cudaFreeHost(h->h_perm[b]);
        if (h->copied[b])
            cudaEventDestroy(h->copied[b]);
        if (h->consumed[b])
            cudaEventDestroy(h->consumed[b]);
    }
    if (h->copy_stream)
        cudaStreamDestroy(h->copy_stream);
    if (h->h_state)
        cudaFreeHost(h->h_state);
    delete h->perm_stream; // joins its threads before the pinned buffer they write goes away
    if (h->h_rng_perms)
        cudaFreeHost(h->h_rng_perms);
    h->scratch.release();
    delete h;
}

extern "C" int cpprl_b200_ppo_state(cpprl_b200_ppo *h, float **d_exp_avg, float **d_exp_avg_sq, int64_t *step)
{
    if (!h)
        return fail(CPPRL_B200_EINVAL, "ppo_state: null handle");
    if (d_exp_avg)
        *d_exp_avg = h->d_m;
    if (d_exp_avg_sq)
        *d_exp_avg_sq = h->d_v;
    if (step)
        *step = h->host_step;
    return 0;
}

extern "C" int cpprl_b200_ppo_set_comm(cpprl_b200_ppo *h, cpprl_b200_comm *comm, int64_t env_offset, int64_t global_envs)
{
    if (!h)
        return fail(CPPRL_B200_EINVAL, "ppo_set_comm: null handle");
    if (h->active)
        return fail(CPPRL_B200_ESTATE, "ppo_set_comm called inside an update");
    if (comm && (env_offset < 0 || global_envs <= 0))
        return fail(CPPRL_B200_EINVAL, "ppo_set_comm: bad environment range");
    h->comm = comm;
    h->env_offset = env_offset;
    h->global_envs = global_envs;
    return 0;
}

extern "C" int cpprl_b200_ppo_set_step(cpprl_b200_ppo *h, int64_t step)
{
    if (!h || step < 0)
        return fail(CPPRL_B200_EINVAL, "ppo_set_step: bad arguments");
    set_opt_step_kernel<<<1, 1>>>(h->scratch.state, step);
    CPPRL_LAUNCH_CHECK("set_opt_step_kernel");
    CPPRL_CUDA_TRY(cudaDeviceSynchronize());
    h->host_step = step;
    return 0;
}

// Recurrent policies: RolloutStorage::recurrent_generator (storage.cpp:195-218) + RecurrentGenerator
// (recurrent_generator.cpp:30-98).  The generator always walks all N entries of randperm(N), each minibatch being
// num_processes / num_mini_batch environments starting at the drawn index; only one environment per minibatch
// survives the reference's own flatten, so that is the only shape accepted here.
static int ppo_begin_gru(cpprl_b200_ppo *h, const cpprl_b200_policy *policy, const cpprl_b200_rollout *ro,
                         const cpprl_b200_ppo_options *opt, void *stream)
{
    if (h->comm)
        return fail(CPPRL_B200_EUNSUPPORTED, "ppo_update: the recurrent path is single-GPU (its minibatch chain is serial)");
    if (!ro->d_hidden_states || !ro->d_masks)
        return fail(CPPRL_B200_EINVAL, "ppo_update: recurrent policy needs the rollout's hidden_states and masks");
    if (ro->hidden_dim != kHid)
        return fail(CPPRL_B200_EINVAL, "ppo_update: rollout hidden_dim must equal the GRU hidden size (64)");
    if (ro->N < opt->num_mini_batch)
        return fail(CPPRL_B200_EINVAL, "PPO needs the number of processes (" + std::to_string(ro->N) +
                                           ") to be greater than or equal to the number of minibatches (" +
                                           std::to_string(opt->num_mini_batch) + ")");
    if (ro->N / opt->num_mini_batch != 1)
        return fail(CPPRL_B200_EUNSUPPORTED, "ppo_update: the recurrent generator only works with one environment per "
                                             "minibatch (num_mini_batch must equal the number of processes)");
    const int64_t expect_cols = policy->head_kind == CPPRL_B200_HEAD_CATEGORICAL ? 1 : policy->num_actions;
    if (ro->action_cols != expect_cols)
        return fail(CPPRL_B200_EINVAL, "ppo_update: rollout action_cols does not match the policy head");
    if (ro->T * ro->N > 0x7fffffffll)
        return fail(CPPRL_B200_EUNSUPPORTED, "ppo_update: more than 2^31 rows");
    h->policy = *policy;
    h->rollout = *ro;
    h->opt = *opt;
    h->stream = static_cast<cudaStream_t>(stream);
    h->mbs = 1; // perm_len = mbs * n_mb = N
    h->n_mb = ro->N;
    h->epoch = 0;
    GruUpdateArgs &a = h->gargs;
    if (int rc = fill_gru_policy_args(a, *policy, ro->T, "ppo_update"))
        return rc;
    const size_t rms_ws = policy->d_norm_mean ? cpprl_b200_rms_workspace_bytes((ro->T + 1) * ro->N, policy->obs_dim) : 0;
    if (int rc = h->scratch.ensure(0, 0, rms_ws))
        return rc;
    if (int rc = h->scratch.ensure_tape(static_cast<size_t>(kGruTapes) * ro->T * kHid))
        return rc;
    a.obs = ro->d_observations;
    a.hxs = ro->d_hidden_states; // hidden_states[0] (recurrent_generator.cpp:62-64)
    a.masks = ro->d_masks;
    a.actions = ro->d_actions;
    a.value_preds = ro->d_value_preds;
    a.returns = ro->d_returns;
    a.old_logp = ro->d_action_log_probs;
    a.T = static_cast<int>(ro->T);
    a.N = static_cast<int>(ro->N);
    a.cut_over_all_agents = 0;
    a.n_mb = static_cast<int>(ro->N);
    a.adv_sums = h->scratch.adv_sums;
    a.rows = ro->T * ro->N;
    a.algo = kAlgoPPO;
    a.clip_param = opt->clip_param;
    a.clip_lo = static_cast<float>(1.0 - static_cast<double>(opt->clip_param));
    a.clip_hi = static_cast<float>(1.0 + static_cast<double>(opt->clip_param));
    a.c_actor = opt->actor_loss_coef;
    a.c_value = opt->value_loss_coef;
    a.c_entropy = opt->entropy_coef;
    a.kl_limit = static_cast<double>(opt->kl_target) * 1.5;
    a.opt_m = h->d_m;
    a.opt_v = h->d_v;
    a.lr = static_cast<double>(opt->learning_rate);
    a.beta1 = 0.9f;
    a.beta2 = 0.999f;
    a.eps = opt->epsilon;
    a.tape = h->scratch.tape;
    a.state = h->scratch.state;
    cudaStream_t s = h->stream;
    KernelTimer timer("prologue", s);
    reset_state_kernel<<<1, 1, 0, s>>>(h->scratch.state);
    CPPRL_LAUNCH_CHECK("reset_state_kernel");
    CPPRL_CUDA_TRY(cudaMemsetAsync(h->scratch.adv_sums, 0, 2 * sizeof(double), s));
    if (int rc = launch_advantage_moments(ro->d_returns, ro->d_value_preds, ro->T * ro->N, h->scratch.adv_sums, h->scratch.adv_ws, s))
        return rc;
    h->active = true;
    return 0;
}

extern "C" int cpprl_b200_ppo_begin(cpprl_b200_ppo *h, const cpprl_b200_policy *policy, const cpprl_b200_rollout *ro,
                                    const cpprl_b200_ppo_options *opt, void *stream)
{
    if (!h || !ro || !opt)
        return fail(CPPRL_B200_EINVAL, "ppo_begin: null argument");
    if (int rc = require_device())
        return rc;
    if (int rc = validate_policy(policy, "ppo_update"))
        return rc;
    if (policy->n_params != h->n_params)
        return fail(CPPRL_B200_EINVAL, "ppo_update: handle was created for a different parameter count");
    if (!ro->d_observations || !ro->d_actions || !ro->d_value_preds || !ro->d_returns || !ro->d_action_log_probs)
        return fail(CPPRL_B200_EINVAL, "ppo_update: null rollout buffer");
    if (ro->T <= 0 || ro->N <= 0)
        return fail(CPPRL_B200_EINVAL, "ppo_update: non-positive rollout size");
    if (opt->num_epoch < 0 || opt->num_mini_batch <= 0)
        return fail(CPPRL_B200_EINVAL, "ppo_update: bad epoch / minibatch count");
    const bool sharded = h->comm != nullptr;
    h->gru = policy->base_kind == CPPRL_B200_BASE_GRU_MLP;
    h->cnn = policy->base_kind == CPPRL_B200_BASE_CNN;
    if (h->gru)
        return ppo_begin_gru(h, policy, ro, opt, stream);
    const int64_t n_envs = sharded ? h->global_envs : ro->N; // minibatches are defined on the GLOBAL rollout
    if (sharded && (h->env_offset < 0 || h->env_offset + ro->N > h->global_envs))
        return fail(CPPRL_B200_EINVAL, "ppo_update: this rank's environment range lies outside the global rollout");
    const int64_t rows = ro->T * n_envs;
    const int64_t local_rows = ro->T * ro->N;
    if (rows < opt->num_mini_batch)
        return fail(CPPRL_B200_EINVAL, "PPO needs the number of processes (" + std::to_string(n_envs) +
                                           ") * the number of steps (" + std::to_string(ro->T) + ") = " +
                                           std::to_string(rows) +
                                           " to be greater than or equal to the number of minibatches (" +
                                           std::to_string(opt->num_mini_batch) + ")");
    const int64_t mbs = rows / opt->num_mini_batch; // storage.cpp:163
    if (rows % mbs != 0)
        return fail(CPPRL_B200_EINVAL, "ppo_update: T*N is not divisible by the minibatch size (the reference's view throws)");
    const int64_t expect_cols = policy->head_kind == CPPRL_B200_HEAD_CATEGORICAL ? 1 : policy->num_actions;
    if (ro->action_cols != expect_cols)
        return fail(CPPRL_B200_EINVAL, "ppo_update: rollout action_cols does not match the policy head");

    h->policy = *policy;
    h->rollout = *ro;
    h->opt = *opt;
    h->stream = static_cast<cudaStream_t>(stream);
    h->mbs = mbs;
    h->n_mb = rows / mbs;
    h->epoch = 0;
    if (h->cnn)
    {
        if (mbs > 0x7fffffffll / kCnnP1 / 4)
            return fail(CPPRL_B200_EUNSUPPORTED, "ppo_update: minibatch too large");
        CnnUpdateArgs &c = h->cargs;
        fill_cnn_policy_args(c, *policy);
        if (sharded)
        {
            c.comm = comm_view(h->comm);
            if (cnn_exchange_floats(c.C, c.A) > 2ll * c.comm.x_stride)
                return fail(CPPRL_B200_EINVAL, "ppo_update: the communicator was created for fewer parameters than this policy has");
        }
        const int chunk = cnn_chunk_rows(mbs);
        if (int rc = h->scratch.ensure(0, 0, 0))
            return rc;
        if (int rc = h->scratch.ensure_tape(cnn_workspace_floats(c.C, c.A, chunk, true)))
            return rc;
        cnn_workspace_carve(h->cws, h->scratch.tape, c.C, c.A, chunk, true);
        c.obs = ro->d_observations;
        // one pass over the frames this update reads (0.6 ms for cfg5's 3.7 GB): raw pixels sit on the TF32 grid and
        // conv1's two GEMMs then run the instantiation with one MMA per k-step instead of two, same bits out (one stream
        // synchronisation brings the answer to the host)
        if (int rc = cnn_scan_observations(c.obs, local_rows * policy->obs_dim, cnn_inexact_flag(h->cws), &c.obs_exact, h->stream))
            return rc;
        c.actions = ro->d_actions;
        c.value_preds = ro->d_value_preds;
        c.returns = ro->d_returns;
        c.old_logp = ro->d_action_log_probs;
        c.rows = rows; // box-wide: the advantage moments and every mean are over the GLOBAL rollout / minibatch
        c.B = static_cast<int>(mbs);
        c.B_norm = static_cast<int>(mbs);
        c.adv_sums = h->scratch.adv_sums;
        c.algo = kAlgoPPO;
        c.clip_param = opt->clip_param;
        c.clip_lo = static_cast<float>(1.0 - static_cast<double>(opt->clip_param));
        c.clip_hi = static_cast<float>(1.0 + static_cast<double>(opt->clip_param));
        c.c_actor = opt->actor_loss_coef;
        c.c_value = opt->value_loss_coef;
        c.c_entropy = opt->entropy_coef;
        c.kl_limit = static_cast<double>(opt->kl_target) * 1.5;
        c.opt = kOptAdam;
        c.opt_m = h->d_m;
        c.opt_v = h->d_v;
        c.lr = static_cast<double>(opt->learning_rate);
        c.beta1 = 0.9f;
        c.beta2 = 0.999f;
        c.eps = opt->epsilon;
        c.state = h->scratch.state;
        cudaStream_t cs = h->stream;
        KernelTimer timer("prologue", cs);
        reset_state_kernel<<<1, 1, 0, cs>>>(h->scratch.state);
        CPPRL_LAUNCH_CHECK("reset_state_kernel");
        double *moments = sharded ? comm_local_adv(h->comm) : h->scratch.adv_sums;
        CPPRL_CUDA_TRY(cudaMemsetAsync(moments, 0, 2 * sizeof(double), cs));
        if (int rc = launch_advantage_moments(ro->d_returns, ro->d_value_preds, local_rows, moments, h->scratch.adv_ws, cs))
            return rc;
        if (sharded)
            if (int rc = cnn_adv_exchange(c.comm, ++*comm_update_seq(h->comm), h->scratch.adv_sums, h->scratch.state, cs))
                return rc;
        if (int rc = cnn_build_images(c, h->cws, cs))
            return rc;
        h->active = true;
        return 0;
    }

    ParamOffsets po;
    mlp_param_offsets(*policy, po);
    const int Dp = round_up(static_cast<int>(policy->obs_dim), 8);
    const int Ap = round_up(static_cast<int>(policy->num_actions), 4);
    const CommView view = comm_view(h->comm);
    // sharded: tiles are sized for the expected local share of a minibatch; the grid (identical on every rank, CTA b
    // pairs with CTA b of the peers) covers the whole box-wide slice of parameters, and ragged minibatches that
    // exceed it simply take more than one tile per CTA
    const int64_t plan_rows = sharded ? (mbs + view.world - 1) / view.world : mbs;
    // (a rank's share of a minibatch is binomial around plan_rows: four standard deviations of headroom)
    const int64_t plan_max = sharded ? plan_rows + static_cast<int64_t>(4.0 * std::sqrt(static_cast<double>(plan_rows))) + 1 : plan_rows;
    if (int rc = choose_plan(plan_rows, po, Dp, Ap, true, h->plan, plan_max))
        return rc;
    const ImageLayout il = h->plan.il;
    if (sharded)
    {
        const int sms = sm_count();
        h->plan.grid = std::min(sms, std::max(h->plan.grid, static_cast<int>(std::min<int64_t>(sms, (plan_rows * 5 / 4 + h->plan.R - 1) / h->plan.R))));
        h->plan.grid = std::max(h->plan.grid, min_training_grid(sms));
        h->plan.n_active = h->plan.grid;
        h->plan.per_cta = round_up((po.Pstride - kNumStats + h->plan.grid - 1) / h->plan.grid, 4);
        if (mlp_update_smem_bytes(h->plan.R, il.floats, po.Pstride, Dp, Ap, h->plan.per_cta, h->plan.tc, il.tail, po.wv) > 227 * 1024)
            return fail(CPPRL_B200_EUNSUPPORTED, "update: observation / action size needs more than 227 KB of shared memory");
        if (static_cast<int64_t>(h->plan.grid) * (h->plan.per_cta + kNumStats) > view.x_stride)
            return fail(CPPRL_B200_EINVAL, "ppo_update: the communicator was created for fewer parameters than this policy has");
    }
    const size_t rms_ws = policy->d_norm_mean ? cpprl_b200_rms_workspace_bytes((ro->T + 1) * ro->N, policy->obs_dim) : 0;
    if (int rc = h->scratch.ensure(il.floats, static_cast<size_t>(h->plan.n_active) * po.Pstride, rms_ws))
        return rc;

    MlpUpdateArgs &a = h->args;
    std::memset(&a, 0, sizeof(a));
    fill_policy_args(a, *policy, po, il, Dp, Ap);
    a.image = h->scratch.image;
    a.obs = ro->d_observations;
    a.actions = ro->d_actions;
    a.value_preds = ro->d_value_preds;
    a.returns = ro->d_returns;
    a.old_logp = ro->d_action_log_probs;
    a.rows = local_rows;
    a.rows_norm = rows;
    a.B = static_cast<int>(mbs);
    a.B_norm = static_cast<int>(mbs);
    a.perm_stride = mbs;
    a.step_rows = nullptr;
    a.comm = view;
    a.n_steps = static_cast<int>(h->n_mb);
    a.n_active = h->plan.n_active;
    a.per_cta = h->plan.per_cta;
    a.tc = h->plan.tc ? 1 : 0;
    a.no_stage_ahead = mlp_no_stage_ahead();
    a.adv_sums = sharded ? comm_local_adv(h->comm) : h->scratch.adv_sums;
    if (sharded)
        a.adv_seq = ++*comm_update_seq(h->comm);
    a.algo = kAlgoPPO;
    a.clip_param = opt->clip_param;
    // clamp(ratio, 1.0 - clip, 1.0 + clip): bounds formed in double, applied to a float tensor (ppo.cpp:111-113)
    a.clip_lo = static_cast<float>(1.0 - static_cast<double>(opt->clip_param));
    a.clip_hi = static_cast<float>(1.0 + static_cast<double>(opt->clip_param));
    a.c_actor = opt->actor_loss_coef;
    a.c_value = opt->value_loss_coef;
    a.c_entropy = opt->entropy_coef;
    a.kl_limit = static_cast<double>(opt->kl_target) * 1.5; // ppo.cpp:100
    a.opt = kOptAdam;
    a.opt_m = h->d_m;
    a.opt_v = h->d_v;
    a.lr = static_cast<double>(opt->learning_rate);
    a.beta1 = 0.9f;
    a.beta2 = 0.999f;
    a.eps = opt->epsilon;
    a.gpart = h->scratch.gpart;
    a.barrier = h->scratch.barrier;
    a.state = h->scratch.state;

    cudaStream_t s = h->stream;
    KernelTimer timer("prologue", s);
    CPPRL_CUDA_TRY(cudaMemsetAsync(h->scratch.image, 0, il.floats * sizeof(float), s));
    reset_state_kernel<<<1, 1, 0, s>>>(h->scratch.state);
    CPPRL_LAUNCH_CHECK("reset_state_kernel");
    CPPRL_CUDA_TRY(cudaMemsetAsync(const_cast<double *>(a.adv_sums), 0, 2 * sizeof(double), s));
    if (int rc = launch_advantage_moments(ro->d_returns, ro->d_value_preds, local_rows, const_cast<double *>(a.adv_sums), h->scratch.adv_ws, s))
        return rc;
    h->active = true;
    return 0;
}

extern "C" int cpprl_b200_debug_tc_gemm(int variant, const float *d_a, const float *d_b, float *d_c, int M, int N, int K,
                                        void *stream)
{
    if (int rc = require_device())
        return rc;
    if (!d_a || !d_b || !d_c || M <= 0 || N <= 0 || K <= 0 || N % 4 || K % 4 || M % 4)
        return fail(CPPRL_B200_EINVAL, "debug_tc_gemm: bad arguments");
    return cnn_debug_tc_gemm(variant, d_a, d_b, d_c, M, N, K, static_cast<cudaStream_t>(stream));
}

static long long *g_dbg_clocks = nullptr;
static int g_dbg_step = 0;
extern "C" void cpprl_b200_debug_phase_clocks(long long *d_buffer, int step)
{
    g_dbg_clocks = d_buffer;
    g_dbg_step = step;
}

// Which variant of the fused MlpBase update a minibatch of `minibatch_rows` rows with `obs_dim` observation columns trains
// on: 1 = tensor-core (tcgen05 3xTF32, mlp_tc.cuh), 0 = fp32 CUDA-core.  What bench.py reports its roofline against.
extern "C" int cpprl_b200_mlp_update_variant(int64_t minibatch_rows, int64_t obs_dim, int64_t num_actions)
{
    if (require_device() != 0 || minibatch_rows <= 0 || obs_dim < 1 || obs_dim > 256 || num_actions < 1 || num_actions > kMaxActions)
        return 0;
    cpprl_b200_policy p{};
    p.base_kind = CPPRL_B200_BASE_MLP;
    p.head_kind = CPPRL_B200_HEAD_NORMAL;
    p.hidden = kHid;
    p.obs_dim = obs_dim;
    p.num_actions = num_actions;
    ParamOffsets po;
    if (!mlp_param_offsets(p, po))
        return 0;
    Plan plan;
    if (choose_plan(minibatch_rows, po, round_up(static_cast<int>(obs_dim), 8), round_up(static_cast<int>(num_actions), 4), true, plan))
        return 0;
    return plan.tc ? plan.R : 0;
}

static int ppo_shard_permutation(cpprl_b200_ppo *h, const int64_t *d_perm)
{
    const size_t need = static_cast<size_t>(h->mbs * h->n_mb);
    if (need > h->lperm_cap)
    {
        CPPRL_CUDA_TRY(cudaStreamSynchronize(h->stream));
        cudaFree(h->d_lperm);
        h->d_lperm = nullptr;
        CPPRL_CUDA_TRY(cudaMalloc(&h->d_lperm, need * sizeof(int64_t)));
        h->lperm_cap = need;
    }
    // n_mb totals + up to 64 piece counts per minibatch (comm.cu)
    if (static_cast<size_t>(h->n_mb) * 65 > h->counts_cap)
    {
        CPPRL_CUDA_TRY(cudaStreamSynchronize(h->stream));
        cudaFree(h->d_counts);
        h->d_counts = nullptr;
        CPPRL_CUDA_TRY(cudaMalloc(&h->d_counts, h->n_mb * 65 * sizeof(int)));
        h->counts_cap = h->n_mb * 65;
    }
    return launch_shard_permutation(d_perm, h->n_mb, h->mbs, h->global_envs, h->env_offset, h->rollout.N, h->d_lperm,
                                    h->d_counts, h->stream);
}

static int ppo_launch_epoch(cpprl_b200_ppo *h, const int64_t *d_perm)
{
    if (h->cnn)
    {
        // one pass over the epoch's minibatches, every kernel enqueued on the stream; the KL decision stays on the
        // device (kernels of later minibatches see state->stopped and return)
        std::vector<int> counts;
        if (h->comm)
        {
            // this rank's rows of every minibatch (stable compaction of the global permutation, comm.cu); the row
            // counts size the grids, so they come back to the host: one small synchronising copy per epoch
            if (int rc = ppo_shard_permutation(h, d_perm))
                return rc;
            counts.resize(h->n_mb);
            CPPRL_CUDA_TRY(cudaMemcpyAsync(counts.data(), h->d_counts, h->n_mb * sizeof(int), cudaMemcpyDeviceToHost, h->stream));
            CPPRL_CUDA_TRY(cudaStreamSynchronize(h->stream));
            d_perm = h->d_lperm;
        }
        KernelTimer timer("update", h->stream);
        h->cargs.perm = reinterpret_cast<const long long *>(d_perm);
        for (int64_t mb = 0; mb < h->n_mb; ++mb)
        {
            h->cargs.row_base = mb * h->mbs;
            if (h->comm)
            {
                h->cargs.B = counts[mb];
                h->cargs.xseq = ++*comm_step_seq(h->comm);
            }
            if (int rc = cnn_minibatch(h->cargs, h->cws, h->stream))
                return rc;
        }
        ++h->epoch;
        return 0;
    }
    if (h->gru)
    {
        h->gargs.perm = reinterpret_cast<const long long *>(d_perm);
        h->gargs.dbg_clocks = g_dbg_clocks;
        h->gargs.dbg_step = g_dbg_step;
        KernelTimer timer("update", h->stream);
        int rc = launch_gru_update(h->gargs, 1, h->stream);
        if (rc == 0)
            ++h->epoch;
        return rc;
    }
    if (h->comm)
    {
        // compact the global permutation to the rows this rank owns (stable, one CTA per minibatch)
        if (int rc = ppo_shard_permutation(h, d_perm))
            return rc;
        d_perm = h->d_lperm;
        h->args.step_rows = h->d_counts;
        unsigned int *seq = comm_step_seq(h->comm);
        h->args.xseq0 = *seq;
        *seq += static_cast<unsigned int>(h->n_mb);
    }
    h->args.perm = reinterpret_cast<const long long *>(d_perm);
    h->args.dbg_clocks = g_dbg_clocks;
    h->args.dbg_step = g_dbg_step;
    h->args.dbg_launch = static_cast<int>(h->scratch.barrier_launches);
    KernelTimer timer("update", h->stream);
    h->args.barrier_sel = static_cast<int>(h->scratch.barrier_launches & 1u);
    int rc = mlp_update_launch(h->plan.R, h->args, h->plan.grid, h->stream);
    if (rc == 0)
        ++h->scratch.barrier_launches; // (only a launch that ran has zeroed the other counter)
    if (rc == 0)
        ++h->epoch;
    return rc;
}

extern "C" int cpprl_b200_ppo_epoch_device(cpprl_b200_ppo *h, const int64_t *d_perm, int64_t perm_len)
{
    if (!h || !d_perm)
        return fail(CPPRL_B200_EINVAL, "ppo_epoch: null argument");
    if (!h->active)
        return fail(CPPRL_B200_ESTATE, "ppo_epoch called without ppo_begin");
    if (perm_len != h->mbs * h->n_mb)
        return fail(CPPRL_B200_EINVAL, h->gru ? "ppo_epoch: permutation length must be N for recurrent policies"
                                              : "ppo_epoch: permutation length must be T*N");
    return ppo_launch_epoch(h, d_perm);
}

// One epoch from a HOST permutation: H2D on the copy stream into device buffer (epoch & 1), kernel after it.  A
// pageable source goes through the handle's pinned staging buffer; a pinned source is copied from where it lies.
static int ppo_epoch_host(cpprl_b200_ppo *h, const int64_t *h_perm, int64_t perm_len, bool src_pinned)
{
    if (static_cast<size_t>(perm_len) > h->perm_cap)
    {
        CPPRL_CUDA_TRY(cudaStreamSynchronize(h->stream));
        CPPRL_CUDA_TRY(cudaStreamSynchronize(h->copy_stream));
        for (int b = 0; b < 2; ++b)
        {
            cudaFree(h->d_perm[b]);
            if (h->h_perm[b])
                cudaFreeHost(h->h_perm[b]);
            h->d_perm[b] = nullptr, h->h_perm[b] = nullptr;
            CPPRL_CUDA_TRY(cudaMalloc(&h->d_perm[b], perm_len * sizeof(int64_t)));
            CPPRL_CUDA_TRY(cudaMallocHost(&h->h_perm[b], perm_len * sizeof(int64_t)));
            h->consumed_valid[b] = false;
        }
        h->perm_cap = perm_len;
    }
    const int b = h->epoch & 1;
    // the pinned buffer is free once its previous H2D finished; the device buffer once its kernel finished
    if (h->consumed_valid[b])
    {
        if (!src_pinned)
            CPPRL_CUDA_TRY(cudaEventSynchronize(h->copied[b]));
        CPPRL_CUDA_TRY(cudaStreamWaitEvent(h->copy_stream, h->consumed[b], 0));
    }
    const int64_t *src = h_perm;
    if (!src_pinned)
    {
        std::memcpy(h->h_perm[b], h_perm, perm_len * sizeof(int64_t));
        src = h->h_perm[b];
    }
    CPPRL_CUDA_TRY(cudaMemcpyAsync(h->d_perm[b], src, perm_len * sizeof(int64_t), cudaMemcpyHostToDevice, h->copy_stream));
    CPPRL_CUDA_TRY(cudaEventRecord(h->copied[b], h->copy_stream));
    CPPRL_CUDA_TRY(cudaStreamWaitEvent(h->stream, h->copied[b], 0));
    int rc = ppo_launch_epoch(h, h->d_perm[b]);
    if (rc != 0)
        return rc;
    CPPRL_CUDA_TRY(cudaEventRecord(h->consumed[b], h->stream));
    h->consumed_valid[b] = true;
    return 0;
}

extern "C" int cpprl_b200_ppo_epoch(cpprl_b200_ppo *h, const int64_t *h_perm, int64_t perm_len)
{
    if (!h || !h_perm)
        return fail(CPPRL_B200_EINVAL, "ppo_epoch: null argument");
    if (!h->active)
        return fail(CPPRL_B200_ESTATE, "ppo_epoch called without ppo_begin");
    if (perm_len != h->mbs * h->n_mb)
        return fail(CPPRL_B200_EINVAL, h->gru ? "ppo_epoch: permutation length must be N for recurrent policies"
                                              : "ppo_epoch: permutation length must be T*N");
    return ppo_epoch_host(h, h_perm, perm_len, false);
}

extern "C" int cpprl_b200_ppo_stopped(cpprl_b200_ppo *h, int *stopped)
{
    if (!h || !stopped)
        return fail(CPPRL_B200_EINVAL, "ppo_stopped: null argument");
    if (!h->active)
        return fail(CPPRL_B200_ESTATE, "ppo_stopped called without ppo_begin");
    CPPRL_CUDA_TRY(cudaMemcpyAsync(h->h_state, h->scratch.state, sizeof(UpdateState), cudaMemcpyDeviceToHost, h->stream));
    CPPRL_CUDA_TRY(cudaStreamSynchronize(h->stream));
    *stopped = h->h_state->stopped;
    return 0;
}

extern "C" int cpprl_b200_ppo_end_async(cpprl_b200_ppo *h)
{
    if (!h)
        return fail(CPPRL_B200_EINVAL, "ppo_end: null handle");
    if (!h->active)
        return fail(CPPRL_B200_ESTATE, "ppo_end called without ppo_begin");
    const cpprl_b200_policy &p = h->policy;
    if (p.d_norm_mean)
    {
        // finish_update: the normaliser sees ALL (T+1)*N observation rows (ppo.cpp:149-154)
        const int64_t M = (h->rollout.T + 1) * h->rollout.N;
        if (int rc = launch_rms_update(h->rollout.d_observations, M, p.obs_dim, p.d_norm_count, p.d_norm_mean,
                                       p.d_norm_var, h->scratch.rms_ws, h->comm, h->stream, &h->scratch.state->comm_error))
            return rc;
    }
    CPPRL_CUDA_TRY(cudaMemcpyAsync(h->h_state, h->scratch.state, sizeof(UpdateState), cudaMemcpyDeviceToHost, h->stream));
    h->active = false;
    return 0;
}

extern "C" int cpprl_b200_ppo_fetch_stats(cpprl_b200_ppo *h, cpprl_b200_ppo_stats *stats)
{
    if (!h || !stats)
        return fail(CPPRL_B200_EINVAL, "ppo_fetch_stats: null argument");
    CPPRL_CUDA_TRY(cudaStreamSynchronize(h->stream));
    const UpdateState &st = *h->h_state;
    if (st.comm_error)
        return fail(CPPRL_B200_ENCCL, "ppo_update: a peer rank never published its exchange record (bounded wait expired)");
    const float n = static_cast<float>(st.num_updates);
    // means over the applied updates; 0/0 = NaN when the first minibatch tripped the KL stop (ppo.cpp:156-159)
    stats->value_loss = st.total_value_loss / n;
    stats->action_loss = st.total_action_loss / n;
    stats->entropy = st.total_entropy / n;
    stats->clip_fraction = st.total_clip_fraction / n;
    stats->kl_divergence = st.last_kl;
    stats->kl_early_stop_update = st.stopped ? st.kl_stop_update : -1.f;
    stats->num_updates = st.num_updates;
    h->host_step = st.opt_step;
    return 0;
}

extern "C" int cpprl_b200_ppo_end(cpprl_b200_ppo *h, cpprl_b200_ppo_stats *stats)
{
    if (int rc = cpprl_b200_ppo_end_async(h))
        return rc;
    return cpprl_b200_ppo_fetch_stats(h, stats);
}

extern "C" int cpprl_b200_ppo_update(cpprl_b200_ppo *h, const cpprl_b200_policy *policy, const cpprl_b200_rollout *rollout,
                                     const cpprl_b200_ppo_options *opt, const int64_t *h_perms, int64_t perm_len,
                                     cpprl_b200_ppo_stats *stats, void *stream)
{
    if (!h_perms || !stats)
        return fail(CPPRL_B200_EINVAL, "ppo_update: null argument");
    if (int rc = cpprl_b200_ppo_begin(h, policy, rollout, opt, stream))
        return rc;
    for (int e = 0; e < opt->num_epoch; ++e)
        if (int rc = cpprl_b200_ppo_epoch(h, h_perms + static_cast<int64_t>(e) * perm_len, perm_len))
        {
            h->active = false;
            return rc;
        }
    return cpprl_b200_ppo_end(h, stats);
}

// PPO::update with the permutations the reference itself would draw (ppo.cpp:63-84 builds one generator per epoch,
// each drawing torch::randperm from the default CPU generator): `rng_state` is that generator's state
// (at::CPUGeneratorImpl::get_state, cpprl_b200_rng_state_bytes() bytes) on entry and what it must be set to on return.
// The epochs' permutations are drawn by background threads while the device runs (host_rng.cu), and the next update's
// are drawn ahead: they are used when that update arrives with the state this one returned.  After a KL early stop the
// state is the one after the last epoch the reference would have started.
extern "C" int cpprl_b200_ppo_update_rng(cpprl_b200_ppo *h, const cpprl_b200_policy *policy, const cpprl_b200_rollout *rollout,
                                         const cpprl_b200_ppo_options *opt, void *rng_state, cpprl_b200_ppo_stats *stats,
                                         void *stream)
{
    if (!rng_state || !stats || !opt || !h)
        return fail(CPPRL_B200_EINVAL, "ppo_update_rng: null argument");
    if (int rc = cpprl_b200_ppo_begin(h, policy, rollout, opt, stream))
        return rc;
    const int64_t n = h->mbs * h->n_mb;
    const int epochs = opt->num_epoch;
    auto bail = [&](int rc) {
        h->active = false;
        return rc;
    };
    if (!PermStream::supported(n))
        return bail(fail(CPPRL_B200_EUNSUPPORTED, "ppo_update_rng: rollout too large for randperm's 32-bit branch"));
    // a ring of 2 * epochs permutations: this update's and, drawn ahead by the stream's threads, the next one's
    const size_t ring_elems = 2 * static_cast<size_t>(n) * epochs;
    if (ring_elems != h->rng_perms_cap)
    {
        delete h->perm_stream; // (stops its threads before the block they write goes away)
        h->perm_stream = nullptr;
        if (h->h_rng_perms)
            cudaFreeHost(h->h_rng_perms);
        h->h_rng_perms = nullptr, h->rng_perms_cap = 0;
        if (int rc = check_cuda(cudaMallocHost(&h->h_rng_perms, ring_elems * sizeof(int64_t)), "cudaMallocHost(perms)"))
            return bail(rc);
        h->rng_perms_cap = ring_elems;
    }
    if (!h->perm_stream)
        h->perm_stream = new PermStream();
    // Ring slots are rewritten only with permutations two updates ahead of the one that last read them, and that update's
    // H2D copies were complete when its ppo_end returned; the device buffers of the previous update may still be read by
    // its last kernels: the H2D below waits on the `consumed` events.
    if (int rc = check_cuda(cudaStreamSynchronize(h->copy_stream), "cudaStreamSynchronize(copy)"))
        return bail(rc);
    if (int rc = h->perm_stream->start(rng_state, n, epochs, h->h_rng_perms, 2 * epochs, 0))
        return bail(rc);
    const bool can_stop = static_cast<double>(opt->kl_target) * 1.5 < 1e3; // a mean log-prob difference cannot reach that
    int drawn = 0, rc = 0;
    for (int e = 0; e < epochs && rc == 0; ++e)
    {
        if (can_stop && e > 0)
        {
            int stopped = 0;
            rc = cpprl_b200_ppo_stopped(h, &stopped);
            if (rc || stopped)
                break; // the reference leaves its epoch loop here: no further randperm is drawn (ppo.cpp:100-104)
        }
        const int64_t *perm = h->perm_stream->wait(e);
        drawn = e + 1;
        rc = ppo_epoch_host(h, perm, n, true);
    }
    if (drawn > 0)
        h->perm_stream->state_after(drawn - 1, rng_state);
    if (rc == 0)
        rc = cpprl_b200_ppo_end(h, stats); // synchronises: every H2D of the pinned block has finished
    else
        cudaStreamSynchronize(h->copy_stream);
    if (rc)
        return bail(rc);
    return 0;
}

// =====================================================================================================
// A2C
// =====================================================================================================
struct cpprl_b200_a2c
{
    int64_t n_params = 0;
    float *d_sq = nullptr;
    Scratch scratch;
    UpdateState *h_state = nullptr;
    cudaStream_t stream = nullptr;
    cpprl_b200_comm *comm = nullptr;
    int64_t global_envs = 0;
};

extern "C" int cpprl_b200_a2c_create(cpprl_b200_a2c **out, int64_t n_params)
{
    if (!out || n_params <= 0)
        return fail(CPPRL_B200_EINVAL, "a2c_create: bad arguments");
    if (int rc = require_device())
        return rc;
    auto *h = new cpprl_b200_a2c();
    h->n_params = n_params;
    auto attempt = [&]() -> int {
        CPPRL_CUDA_TRY(cudaMalloc(&h->d_sq, n_params * sizeof(float)));
        CPPRL_CUDA_TRY(cudaMemset(h->d_sq, 0, n_params * sizeof(float)));
        CPPRL_CUDA_TRY(cudaMallocHost(&h->h_state, sizeof(UpdateState)));
        return h->scratch.ensure(0, 0, 0);
    };
    int rc = attempt();
    if (rc != 0)
    {
        cpprl_b200_a2c_destroy(h);
        return rc;
    }
    *out = h;
    return 0;
}

extern "C" void cpprl_b200_a2c_destroy(cpprl_b200_a2c *h)
{
    if (!h)
        return;
    cudaFree(h->d_sq);
    if (h->h_state)
        cudaFreeHost(h->h_state);
    h->scratch.release();
    delete h;
}

extern "C" int cpprl_b200_a2c_set_comm(cpprl_b200_a2c *h, cpprl_b200_comm *comm, int64_t global_envs)
{
    if (!h || (comm && global_envs <= 0))
        return fail(CPPRL_B200_EINVAL, "a2c_set_comm: bad arguments");
    h->comm = comm;
    h->global_envs = global_envs;
    return 0;
}

extern "C" int cpprl_b200_a2c_state(cpprl_b200_a2c *h, float **d_square_avg)
{
    if (!h || !d_square_avg)
        return fail(CPPRL_B200_EINVAL, "a2c_state: null argument");
    *d_square_avg = h->d_sq;
    return 0;
}

// Recurrent A2C: evaluate_actions sees all T*N rows in sequence mode with hidden_states[0] (a2c.cpp:58-63), so runs
// are cut where ANY environment's mask is zero.  One CTA per environment; the CTAs add their gradients into one
// buffer and the stand-alone RMSprop kernel applies it.
static int a2c_update_gru(cpprl_b200_a2c *h, const cpprl_b200_policy *policy, const cpprl_b200_rollout *ro,
                          const cpprl_b200_a2c_options *opt, cudaStream_t s)
{
    if (h->comm)
        return fail(CPPRL_B200_EUNSUPPORTED, "a2c_update: the recurrent path is single-GPU");
    if (!ro->d_hidden_states || !ro->d_masks)
        return fail(CPPRL_B200_EINVAL, "a2c_update: recurrent policy needs the rollout's hidden_states and masks");
    if (ro->hidden_dim != kHid)
        return fail(CPPRL_B200_EINVAL, "a2c_update: rollout hidden_dim must equal the GRU hidden size (64)");
    GruUpdateArgs a;
    if (int rc = fill_gru_policy_args(a, *policy, ro->T, "a2c_update"))
        return rc;
    const size_t rms_ws = policy->d_norm_mean ? cpprl_b200_rms_workspace_bytes((ro->T + 1) * ro->N, policy->obs_dim) : 0;
    if (int rc = h->scratch.ensure(0, static_cast<size_t>(a.po.P), rms_ws))
        return rc;
    if (int rc = h->scratch.ensure_tape(static_cast<size_t>(ro->N) * kGruTapes * ro->T * kHid))
        return rc;
    KernelTimer timer("update", s);
    if (policy->d_norm_mean)
    {
        if (int rc = launch_rms_update(ro->d_observations, (ro->T + 1) * ro->N, policy->obs_dim, policy->d_norm_count,
                                       policy->d_norm_mean, policy->d_norm_var, h->scratch.rms_ws, nullptr, s, nullptr))
            return rc;
    }
    a.obs = ro->d_observations;
    a.hxs = ro->d_hidden_states;
    a.masks = ro->d_masks;
    a.actions = ro->d_actions;
    a.value_preds = ro->d_value_preds;
    a.returns = ro->d_returns;
    a.T = static_cast<int>(ro->T);
    a.N = static_cast<int>(ro->N);
    a.cut_over_all_agents = 1;
    a.n_mb = 1;
    a.rows = ro->T * ro->N;
    a.algo = kAlgoA2C;
    a.grad_out = h->scratch.gpart;
    a.c_actor = 1.f;
    a.c_value = opt->value_loss_coef;
    a.c_entropy = opt->entropy_coef;
    a.tape = h->scratch.tape;
    a.state = h->scratch.state;
    CPPRL_CUDA_TRY(cudaMemsetAsync(h->scratch.gpart, 0, a.po.P * sizeof(float), s));
    reset_state_kernel<<<1, 1, 0, s>>>(h->scratch.state);
    CPPRL_LAUNCH_CHECK("reset_state_kernel");
    if (int rc = launch_gru_update(a, static_cast<int>(ro->N), s))
        return rc;
    if (int rc = cpprl_b200_rmsprop_step(policy->d_params, h->scratch.gpart, h->d_sq, a.po.P, opt->learning_rate,
                                         opt->alpha, opt->epsilon, s))
        return rc;
    CPPRL_CUDA_TRY(cudaMemcpyAsync(h->h_state, h->scratch.state, sizeof(UpdateState), cudaMemcpyDeviceToHost, s));
    return 0;
}

extern "C" int cpprl_b200_a2c_update_async(cpprl_b200_a2c *h, const cpprl_b200_policy *policy,
                                           const cpprl_b200_rollout *ro, const cpprl_b200_a2c_options *opt, void *stream)
{
    if (!h || !ro || !opt)
        return fail(CPPRL_B200_EINVAL, "a2c_update: null argument");
    if (int rc = require_device())
        return rc;
    if (int rc = validate_policy(policy, "a2c_update"))
        return rc;
    if (policy->n_params != h->n_params)
        return fail(CPPRL_B200_EINVAL, "a2c_update: handle was created for a different parameter count");
    if (!ro->d_observations || !ro->d_actions || !ro->d_returns)
        return fail(CPPRL_B200_EINVAL, "a2c_update: null rollout buffer");
    if (ro->T <= 0 || ro->N <= 0)
        return fail(CPPRL_B200_EINVAL, "a2c_update: non-positive rollout size");
    const int64_t rows = ro->T * ro->N;
    const bool sharded = h->comm != nullptr;
    const int64_t rows_global = sharded ? ro->T * h->global_envs : rows;
    if (rows_global > 0x7fffffffll)
        return fail(CPPRL_B200_EUNSUPPORTED, "a2c_update: more than 2^31 rows");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    h->stream = s;
    if (policy->base_kind == CPPRL_B200_BASE_GRU_MLP)
        return a2c_update_gru(h, policy, ro, opt, s);
    if (policy->base_kind == CPPRL_B200_BASE_CNN)
    {
        CnnUpdateArgs c;
        fill_cnn_policy_args(c, *policy);
        if (sharded)
        {
            // every rank holds the same number of environments; the loss means are over the box-wide T * global_envs rows
            c.comm = comm_view(h->comm);
            if (cnn_exchange_floats(c.C, c.A) > 2ll * c.comm.x_stride)
                return fail(CPPRL_B200_EINVAL, "a2c_update: the communicator was created for fewer parameters than this policy has");
            c.xseq = ++*comm_step_seq(h->comm);
        }
        const int chunk = cnn_chunk_rows(rows);
        if (int rc = h->scratch.ensure(0, 0, 0))
            return rc;
        if (int rc = h->scratch.ensure_tape(cnn_workspace_floats(c.C, c.A, chunk, true)))
            return rc;
        CnnWorkspace ws;
        cnn_workspace_carve(ws, h->scratch.tape, c.C, c.A, chunk, true);
        c.obs = ro->d_observations;
        if (int rc = cnn_scan_observations(c.obs, rows * policy->obs_dim, cnn_inexact_flag(ws), &c.obs_exact, s))
            return rc;
        c.actions = ro->d_actions;
        c.value_preds = ro->d_value_preds;
        c.returns = ro->d_returns;
        c.rows = rows_global;
        c.B = static_cast<int>(rows);
        c.B_norm = static_cast<int>(rows_global);
        c.algo = kAlgoA2C;
        c.c_actor = 1.f;
        c.c_value = opt->value_loss_coef;
        c.c_entropy = opt->entropy_coef;
        c.opt = kOptRmsprop;
        c.opt_v = h->d_sq;
        c.lr = static_cast<double>(opt->learning_rate);
        c.alpha = opt->alpha;
        c.eps = opt->epsilon;
        c.state = h->scratch.state;
        KernelTimer timer("update", s);
        reset_state_kernel<<<1, 1, 0, s>>>(h->scratch.state);
        CPPRL_LAUNCH_CHECK("reset_state_kernel");
        if (int rc = cnn_build_images(c, ws, s))
            return rc;
        if (int rc = cnn_minibatch(c, ws, s))
            return rc;
        CPPRL_CUDA_TRY(cudaMemcpyAsync(h->h_state, h->scratch.state, sizeof(UpdateState), cudaMemcpyDeviceToHost, s));
        return 0;
    }

    ParamOffsets po;
    mlp_param_offsets(*policy, po);
    const int Dp = round_up(static_cast<int>(policy->obs_dim), 8);
    const int Ap = round_up(static_cast<int>(policy->num_actions), 4);
    Plan plan;
    const CommView view = comm_view(h->comm);
    // sharded: CTA b pairs with CTA b of the peers and the exchange records sit at blockIdx * (per_cta + stats), so the
    // grid and per_cta must come from a RANK-INDEPENDENT quantity -- the mean shard, ceil(T * global_envs / world) --
    // not from this rank's own row count (shards may differ by one environment, api.shard_range); a rank that holds
    // more rows than the plan simply walks more than one tile per CTA
    const int64_t plan_rows = sharded ? (rows_global + view.world - 1) / view.world : rows;
    if (int rc = choose_plan(plan_rows, po, Dp, Ap, true, plan, sharded ? plan_rows + ro->T : plan_rows)) // (shards differ by one environment)
        return rc;
    const ImageLayout il = plan.il;
    if (sharded)
    {
        plan.n_active = plan.grid;
        if (static_cast<int64_t>(plan.grid) * (plan.per_cta + kNumStats) > view.x_stride)
            return fail(CPPRL_B200_EINVAL, "a2c_update: the communicator was created for fewer parameters than this policy has");
    }
    const size_t rms_ws = policy->d_norm_mean ? cpprl_b200_rms_workspace_bytes((ro->T + 1) * ro->N, policy->obs_dim) : 0;
    if (int rc = h->scratch.ensure(il.floats, static_cast<size_t>(plan.n_active) * po.Pstride, rms_ws))
        return rc;

    KernelTimer timer("update", s);
    // (the state record is reset first: the sharded normaliser update below may raise comm_error in it)
    reset_state_kernel<<<1, 1, 0, s>>>(h->scratch.state);
    CPPRL_LAUNCH_CHECK("reset_state_kernel");
    if (policy->d_norm_mean)
    {
        // the normaliser is updated BEFORE the forward pass (a2c.cpp:52-55)
        if (int rc = launch_rms_update(ro->d_observations, (ro->T + 1) * ro->N, policy->obs_dim, policy->d_norm_count,
                                       policy->d_norm_mean, policy->d_norm_var, h->scratch.rms_ws, h->comm, s, &h->scratch.state->comm_error))
            return rc;
    }
    MlpUpdateArgs a;
    std::memset(&a, 0, sizeof(a));
    fill_policy_args(a, *policy, po, il, Dp, Ap);
    a.image = h->scratch.image;
    a.obs = ro->d_observations;
    a.actions = ro->d_actions;
    a.value_preds = ro->d_value_preds;
    a.returns = ro->d_returns;
    a.old_logp = ro->d_action_log_probs ? ro->d_action_log_probs : ro->d_returns;
    a.rows = rows;
    a.perm = nullptr;
    a.B = static_cast<int>(rows);
    a.B_norm = static_cast<int>(rows_global);
    a.rows_norm = rows_global;
    a.perm_stride = rows;
    a.step_rows = nullptr;
    a.comm = view;
    if (sharded)
    {
        unsigned int *seq = comm_step_seq(h->comm);
        a.xseq0 = *seq;
        *seq += 1u;
    }
    a.n_steps = 1;
    a.n_active = plan.n_active;
    a.per_cta = plan.per_cta;
    a.tc = plan.tc ? 1 : 0;
    a.no_stage_ahead = mlp_no_stage_ahead();
    a.adv_sums = nullptr;
    a.algo = kAlgoA2C;
    a.c_actor = 1.f;
    a.c_value = opt->value_loss_coef;
    a.c_entropy = opt->entropy_coef;
    a.opt = kOptRmsprop;
    a.opt_m = nullptr;
    a.opt_v = h->d_sq;
    a.lr = static_cast<double>(opt->learning_rate);
    a.alpha = opt->alpha;
    a.eps = opt->epsilon;
    a.gpart = h->scratch.gpart;
    a.barrier = h->scratch.barrier;
    a.state = h->scratch.state;
    CPPRL_CUDA_TRY(cudaMemsetAsync(h->scratch.image, 0, il.floats * sizeof(float), s));
    a.barrier_sel = static_cast<int>(h->scratch.barrier_launches & 1u);
    a.dbg_clocks = g_dbg_clocks; // (cpprl_b200_debug_phase_clocks)
    a.dbg_step = g_dbg_step;
    if (int rc = mlp_update_launch(plan.R, a, plan.grid, s))
        return rc;
    ++h->scratch.barrier_launches;
    CPPRL_CUDA_TRY(cudaMemcpyAsync(h->h_state, h->scratch.state, sizeof(UpdateState), cudaMemcpyDeviceToHost, s));
    return 0;
}

extern "C" int cpprl_b200_a2c_fetch_stats(cpprl_b200_a2c *h, cpprl_b200_a2c_stats *stats)
{
    if (!h || !stats)
        return fail(CPPRL_B200_EINVAL, "a2c_fetch_stats: null argument");
    CPPRL_CUDA_TRY(cudaStreamSynchronize(h->stream));
    if (h->h_state->comm_error)
        return fail(CPPRL_B200_ENCCL, "a2c_update: a peer rank never published its exchange record (bounded wait expired)");
    stats->value_loss = h->h_state->a2c_value_loss;
    stats->action_loss = h->h_state->a2c_action_loss;
    stats->entropy = h->h_state->a2c_entropy;
    return 0;
}

extern "C" int cpprl_b200_a2c_update(cpprl_b200_a2c *h, const cpprl_b200_policy *policy, const cpprl_b200_rollout *rollout,
                                     const cpprl_b200_a2c_options *opt, cpprl_b200_a2c_stats *stats, void *stream)
{
    if (!stats)
        return fail(CPPRL_B200_EINVAL, "a2c_update: null stats");
    if (int rc = cpprl_b200_a2c_update_async(h, policy, rollout, opt, stream))
        return rc;
    return cpprl_b200_a2c_fetch_stats(h, stats);
}

// =====================================================================================================
// forward evaluation
// =====================================================================================================
// Scratch of the forward entry points (evaluate_actions, get_values, act, bootstrap_returns): weight image, barrier word,
// CNN / GRU tape, distribution buffer, placeholder actions.  These entry points take an arbitrary `void *stream`, so
// the scratch is keyed by (device, stream): work enqueued on different streams or devices never shares a buffer, and a
// buffer that has to grow is only ever in use on its own stream (synchronised before it is freed).  g_eval_mu guards
// the map and serialises the host-side enqueue.
struct EvalScratch
{
    Scratch sc;
    float *act_dist = nullptr;
    void *act_zero = nullptr;
    size_t act_dist_cap = 0, act_zero_cap = 0;
};
static std::mutex g_eval_mu;
static std::map<std::pair<int, cudaStream_t>, EvalScratch> g_eval_map;

// Caller holds g_eval_mu.
static EvalScratch &eval_scratch_for(cudaStream_t s)
{
    int dev = 0;
    cudaGetDevice(&dev);
    return g_eval_map[std::make_pair(dev, s)];
}

extern "C" int64_t cpprl_b200_policy_num_params(const cpprl_b200_policy *policy)
{
    if (!policy)
        return 0;
    ParamOffsets po;
    GruParamOffsets gpo;
    if (policy->base_kind == CPPRL_B200_BASE_CNN)
        return cnn_param_offsets(static_cast<int>(policy->obs_chw[0]), static_cast<int>(policy->num_actions),
                                 policy->head_kind == CPPRL_B200_HEAD_NORMAL ? 1 : 0).P;
    if (gru_param_offsets(*policy, gpo))
        return gpo.P;
    if (!mlp_param_offsets(*policy, po))
        return 0;
    return po.P;
}

// forward of any base on B explicit rows; d_dist (nullable, [B][A]) receives what act() samples from.  Caller holds g_eval_mu.
static int evaluate_impl(EvalScratch &es, const cpprl_b200_policy *policy, const float *d_obs, const float *d_hxs, int64_t n_hxs,
                         const float *d_masks, const void *d_actions, int64_t B, float *d_value, float *d_logp,
                         float *d_entropy, float *d_hxs_out, float *d_dist, cudaStream_t s)
{
    Scratch &g_eval_scratch = es.sc;

    if (policy->base_kind == CPPRL_B200_BASE_CNN)
    {
        CnnUpdateArgs c;
        fill_cnn_policy_args(c, *policy);
        const int chunk = cnn_chunk_rows(B);
        if (int rc = g_eval_scratch.ensure_tape(cnn_workspace_floats(c.C, c.A, chunk, false)))
            return rc;
        CnnWorkspace ws;
        cnn_workspace_carve(ws, g_eval_scratch.tape, c.C, c.A, chunk, false);
        c.obs = d_obs;
        c.actions = d_actions;
        c.rows = B;
        c.B = static_cast<int>(B);
        c.algo = kAlgoForward;
        c.out_value = d_value;
        c.out_logp = d_logp;
        c.out_entropy_sum = d_entropy;
        c.out_dist = d_dist;
        CPPRL_CUDA_TRY(cudaMemsetAsync(d_entropy, 0, sizeof(float), s));
        if (int rc = cnn_build_images(c, ws, s))
            return rc;
        if (int rc = cnn_forward(c, ws, s))
            return rc;
        finish_entropy_kernel<<<1, 1, 0, s>>>(d_entropy, 1.f / static_cast<float>(B));
        CPPRL_LAUNCH_CHECK("finish_entropy_kernel");
        return 0;
    }
    if (policy->base_kind == CPPRL_B200_BASE_GRU_MLP)
    {
        // NNBase::forward_gru (nn_base.cpp:47-114): n_hxs == B is the one-step mode, otherwise x is [T, n_hxs] flattened
        if (!d_hxs || !d_masks || n_hxs <= 0)
            return fail(CPPRL_B200_EINVAL, "evaluate_actions: recurrent policy needs hidden states and masks");
        if (B % n_hxs != 0)
            return fail(CPPRL_B200_EINVAL, "evaluate_actions: batch size is not a multiple of the number of hidden states");
        const int64_t T = B / n_hxs;
        GruUpdateArgs a;
        if (int rc = fill_gru_policy_args(a, *policy, T, "evaluate_actions"))
            return rc;
        if (int rc = g_eval_scratch.ensure(0, 0, 0))
            return rc;
        if (int rc = g_eval_scratch.ensure_tape(static_cast<size_t>(n_hxs) * T * kHid))
            return rc;
        a.obs = d_obs;
        a.hxs = d_hxs;
        a.masks = d_masks;
        a.actions = d_actions;
        a.T = static_cast<int>(T);
        a.N = static_cast<int>(n_hxs);
        a.cut_over_all_agents = 1;
        a.rows = B;
        a.algo = kAlgoForward;
        a.tape = g_eval_scratch.tape;
        a.out_value = d_value;
        a.out_logp = d_logp;
        a.out_entropy_sum = d_entropy;
        a.out_hxs = d_hxs_out;
        a.out_dist = d_dist;
        a.step_mode = T == 1;
        CPPRL_CUDA_TRY(cudaMemsetAsync(d_entropy, 0, sizeof(float), s));
        if (int rc = launch_gru_update(a, static_cast<int>(n_hxs), s))
            return rc;
        // sequence mode hands the incoming hidden states back unchanged (nn_base.cpp:110-112)
        if (T != 1 && d_hxs_out && d_hxs_out != d_hxs)
            CPPRL_CUDA_TRY(cudaMemcpyAsync(d_hxs_out, d_hxs, n_hxs * kHid * sizeof(float), cudaMemcpyDeviceToDevice, s));
        finish_entropy_kernel<<<1, 1, 0, s>>>(d_entropy, 1.f / static_cast<float>(B));
        CPPRL_LAUNCH_CHECK("finish_entropy_kernel");
        return 0;
    }

    ParamOffsets po;
    mlp_param_offsets(*policy, po);
    const int Dp = round_up(static_cast<int>(policy->obs_dim), 8);
    const int Ap = round_up(static_cast<int>(policy->num_actions), 4);
    Plan plan;
    if (int rc = choose_plan(B, po, Dp, Ap, false, plan))
        return rc;
    const ImageLayout il = plan.il;
    if (int rc = g_eval_scratch.ensure(il.floats, 0, 0))
        return rc;
    MlpUpdateArgs a;
    std::memset(&a, 0, sizeof(a));
    fill_policy_args(a, *policy, po, il, Dp, Ap);
    a.image = g_eval_scratch.image;
    a.obs = d_obs;
    a.actions = d_actions;
    a.rows = B;
    a.B = static_cast<int>(B);
    a.n_steps = 1;
    a.n_active = plan.n_active;
    a.per_cta = 0;
    a.B_norm = static_cast<int>(B);
    a.rows_norm = B;
    a.perm_stride = B;
    a.comm.world = 1;
    a.algo = kAlgoForward;
    a.barrier = g_eval_scratch.barrier;
    a.out_value = d_value;
    a.out_logp = d_logp;
    a.out_entropy_sum = d_entropy;
    a.out_dist = d_dist;
    CPPRL_CUDA_TRY(cudaMemsetAsync(g_eval_scratch.image, 0, il.floats * sizeof(float), s));
    CPPRL_CUDA_TRY(cudaMemsetAsync(d_entropy, 0, sizeof(float), s));
    a.barrier_sel = static_cast<int>(g_eval_scratch.barrier_launches & 1u);
    if (int rc = mlp_update_launch(plan.R, a, plan.grid, s))
        return rc;
    ++g_eval_scratch.barrier_launches;
    finish_entropy_kernel<<<1, 1, 0, s>>>(d_entropy, 1.f / static_cast<float>(B)); // entropy().mean() (policy.cpp:109)
    CPPRL_LAUNCH_CHECK("finish_entropy_kernel");
    return 0;
}

extern "C" int cpprl_b200_evaluate_actions(const cpprl_b200_policy *policy, const float *d_obs, const float *d_hxs,
                                           int64_t n_hxs, const float *d_masks, const void *d_actions, int64_t B,
                                           float *d_value, float *d_logp, float *d_entropy, float *d_hxs_out,
                                           void *stream)
{
    if (int rc = require_device())
        return rc;
    if (int rc = validate_policy(policy, "evaluate_actions"))
        return rc;
    if (!d_obs || !d_actions || !d_value || !d_logp || !d_entropy)
        return fail(CPPRL_B200_EINVAL, "evaluate_actions: null pointer");
    if (B <= 0 || B > 0x7fffffffll)
        return fail(CPPRL_B200_EINVAL, "evaluate_actions: bad batch size");
    std::lock_guard<std::mutex> lock(g_eval_mu);
    return evaluate_impl(eval_scratch_for(static_cast<cudaStream_t>(stream)), policy, d_obs, d_hxs, n_hxs, d_masks, d_actions, B,
                         d_value, d_logp, d_entropy, d_hxs_out, nullptr, static_cast<cudaStream_t>(stream));
}

// ---- PolicyImpl::act (policy.cpp:52-81): forward + sample + log-prob ----------------------------------------------------
// Categorical::sample is torch::multinomial(probs, 1) (categorical.cpp:63-80): inverse CDF of one uniform here;
// Normal::sample is loc + scale * N(0, 1) (normal.cpp:39-44).  Philox4x32-10 keyed by (seed, row), so a given
// (seed, offset) reproduces the same actions whatever the launch shape.  Parity with the reference is distributional
// (it draws from the CPU mt19937 stream); the log-probabilities are those evaluate_actions returns for the same actions.
__global__ void __launch_bounds__(256)
    act_sample_kernel(const float *__restrict__ dist, const float *__restrict__ params, int slog_offset, int head, int A,
                      long long B, unsigned long long seed, unsigned long long offset, void *__restrict__ actions,
                      float *__restrict__ logp)
{
    const long long row = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (row >= B)
        return;
    curandStatePhilox4_32_10_t rng;
    curand_init(seed, static_cast<unsigned long long>(row), offset, &rng);
    const float *d = dist + row * A;
    if (head == 0)
    {
        const float u = curand_uniform(&rng); // (0, 1]
        float cdf = 0.f;
        int pick = A - 1;
        for (int i = 0; i < A; ++i)
        {
            cdf += expf(d[i]);
            if (u <= cdf)
            {
                pick = i;
                break;
            }
        }
        static_cast<long long *>(actions)[row] = pick;
        logp[row] = d[pick];
    }
    else
    {
        float lp = 0.f;
        for (int i = 0; i < A; i += 4)
        {
            const float4 z = curand_normal4(&rng);
            const float zz[4] = {z.x, z.y, z.z, z.w};
            for (int j = 0; j < 4 && i + j < A; ++j)
            {
                const float log_sigma = params[slog_offset + i + j], sigma = expf(log_sigma);
                const float x = d[i + j] + sigma * zz[j];
                const float diff = x - d[i + j];
                static_cast<float *>(actions)[row * A + i + j] = x;
                lp += -(diff * diff) / (2.f * sigma * sigma) - log_sigma - 0.91893853320467274178f; // normal.cpp:24-37
            }
        }
        logp[row] = lp; // summed over the action dimensions (policy.cpp:72-75)
    }
}

// scratch of the act / get_values entry points: the distribution parameters [B][A] + one float of entropy, and
// placeholder actions (zeros) for the forward kernels' log-prob stage
static int ensure_act_scratch(EvalScratch &es, int64_t B, int A, bool normal, cudaStream_t s)
{
    float *&g_act_dist = es.act_dist;
    void *&g_act_zero = es.act_zero;
    size_t &g_act_dist_cap = es.act_dist_cap, &g_act_zero_cap = es.act_zero_cap;
    const size_t need_dist = static_cast<size_t>(B) * A + 4, need_zero = static_cast<size_t>(B) * (normal ? A * sizeof(float) : sizeof(int64_t));
    if (need_dist > g_act_dist_cap)
    {
        CPPRL_CUDA_TRY(cudaStreamSynchronize(s));
        cudaFree(g_act_dist);
        g_act_dist = nullptr;
        CPPRL_CUDA_TRY(cudaMalloc(&g_act_dist, need_dist * sizeof(float)));
        g_act_dist_cap = need_dist;
    }
    if (need_zero > g_act_zero_cap)
    {
        CPPRL_CUDA_TRY(cudaStreamSynchronize(s));
        cudaFree(g_act_zero);
        g_act_zero = nullptr;
        CPPRL_CUDA_TRY(cudaMalloc(&g_act_zero, need_zero));
        CPPRL_CUDA_TRY(cudaMemset(g_act_zero, 0, need_zero));
        g_act_zero_cap = need_zero;
    }
    return 0;
}

// PolicyImpl::get_values (policy.cpp:132-144): the fused forward with placeholder actions; log-probabilities and the
// entropy go to scratch.  Caller holds g_eval_mu.
static int get_values_locked(const cpprl_b200_policy *policy, const float *d_obs, const float *d_hxs, int64_t n_hxs,
                             const float *d_masks, int64_t B, float *d_value, cudaStream_t s)
{
    EvalScratch &es = eval_scratch_for(s);
    const int A = static_cast<int>(policy->num_actions);
    if (int rc = ensure_act_scratch(es, B, A, policy->head_kind == CPPRL_B200_HEAD_NORMAL, s))
        return rc;
    return evaluate_impl(es, policy, d_obs, d_hxs, n_hxs, d_masks, es.act_zero, B, d_value, es.act_dist, es.act_dist + B, nullptr,
                         nullptr, s);
}

extern "C" int cpprl_b200_get_values(const cpprl_b200_policy *policy, const float *d_obs, const float *d_hxs, int64_t n_hxs,
                                     const float *d_masks, int64_t B, float *d_value, void *stream)
{
    if (int rc = require_device())
        return rc;
    if (int rc = validate_policy(policy, "get_values"))
        return rc;
    if (!d_obs || !d_value)
        return fail(CPPRL_B200_EINVAL, "get_values: null pointer");
    if (B <= 0 || B > 0x7fffffffll)
        return fail(CPPRL_B200_EINVAL, "get_values: bad batch size");
    std::lock_guard<std::mutex> lock(g_eval_mu);
    return get_values_locked(policy, d_obs, d_hxs, n_hxs, d_masks, B, d_value, static_cast<cudaStream_t>(stream));
}

extern "C" int cpprl_b200_bootstrap_returns(const cpprl_b200_policy *policy, const cpprl_b200_storage *st, int use_gae,
                                            float gamma, float tau, void *stream)
{
    if (int rc = require_device())
        return rc;
    if (int rc = validate_policy(policy, "bootstrap_returns"))
        return rc;
    if (!st || st->T <= 0 || st->N <= 0 || !st->d_observations || !st->d_hidden_states || !st->d_rewards || !st->d_value_preds ||
        !st->d_returns || !st->d_masks)
        return fail(CPPRL_B200_EINVAL, "bootstrap_returns: bad storage description");
    if (st->obs_dim != policy->obs_dim)
        return fail(CPPRL_B200_EINVAL, "bootstrap_returns: observation width of the storage differs from the policy's");
    const bool recurrent = policy->base_kind == CPPRL_B200_BASE_GRU_MLP;
    if (recurrent && st->hidden_dim != policy->hidden)
        return fail(CPPRL_B200_EINVAL, "bootstrap_returns: hidden-state width of the storage differs from the policy's");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const int64_t T = st->T, N = st->N;
    // the value of the last observation is written where compute_returns would copy next_value to
    // (value_predictions[T] with GAE, storage.cpp:120; returns[T] otherwise, storage.cpp:135): no temporary
    float *next_value = (use_gae ? st->d_value_preds : st->d_returns) + T * N;
    {
        std::lock_guard<std::mutex> lock(g_eval_mu);
        if (int rc = get_values_locked(policy, st->d_observations + T * N * st->obs_dim,
                                       recurrent ? st->d_hidden_states + T * N * st->hidden_dim : nullptr, recurrent ? N : 0,
                                       st->d_masks + T * N, N, next_value, s))
            return rc;
    }
    return cpprl_b200_compute_returns(st->d_rewards, st->d_value_preds, st->d_masks, next_value, st->d_returns, T, N, use_gae,
                                      gamma, tau, stream);
}

extern "C" int cpprl_b200_act(const cpprl_b200_policy *policy, const float *d_obs, const float *d_hxs, int64_t n_hxs,
                              const float *d_masks, int64_t B, uint64_t seed, uint64_t offset, float *d_value,
                              void *d_actions, float *d_logp, float *d_hxs_out, void *stream)
{
    if (int rc = require_device())
        return rc;
    if (int rc = validate_policy(policy, "act"))
        return rc;
    if (!d_obs || !d_actions || !d_value || !d_logp)
        return fail(CPPRL_B200_EINVAL, "act: null pointer");
    if (B <= 0 || B > 0x7fffffffll)
        return fail(CPPRL_B200_EINVAL, "act: bad batch size");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const int A = static_cast<int>(policy->num_actions);
    const bool normal = policy->head_kind == CPPRL_B200_HEAD_NORMAL;
    std::lock_guard<std::mutex> lock(g_eval_mu);
    EvalScratch &es = eval_scratch_for(s);
    if (int rc = ensure_act_scratch(es, B, A, normal, s))
        return rc;
    float *const g_act_dist = es.act_dist;
    float *d_entropy = g_act_dist + static_cast<size_t>(B) * A;
    if (int rc = evaluate_impl(es, policy, d_obs, d_hxs, n_hxs, d_masks, es.act_zero, B, d_value, d_logp, d_entropy, d_hxs_out,
                               g_act_dist, s))
        return rc;
    int slog = 0;
    if (normal)
    {
        if (policy->base_kind == CPPRL_B200_BASE_CNN)
            slog = cnn_param_offsets(static_cast<int>(policy->obs_chw[0]), A, 1).slog;
        else if (policy->base_kind == CPPRL_B200_BASE_GRU_MLP)
        {
            GruUpdateArgs g;
            if (int rc = fill_gru_policy_args(g, *policy, 1, "act"))
                return rc;
            slog = g.po.slog;
        }
        else
        {
            ParamOffsets po;
            mlp_param_offsets(*policy, po);
            slog = po.slog;
        }
    }
    act_sample_kernel<<<static_cast<unsigned>((B + 255) / 256), 256, 0, s>>>(g_act_dist, policy->d_params, slog, normal ? 1 : 0, A,
                                                                            B, seed, offset, d_actions, d_logp);
    CPPRL_LAUNCH_CHECK("act_sample_kernel");
    return 0;
}
