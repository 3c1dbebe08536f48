// oracle/ref_driver.cpp — C-ABI driver over the UNMODIFIED reference classes.
//
// TEST INFRASTRUCTURE ONLY (oracle).  Compiled by oracle/build_ref.sh together with the
// reference's own translation units (read from /root/reference, namespace renamed to
// `cpprl_ref` by -Dcpprl=cpprl_ref) into oracle/_ref/libcpprl_ref.so.  Only tests/,
// __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it.
//
// Nothing here re-implements arithmetic: every function builds the reference's objects
// (RolloutStorage, Policy, PPO, A2C, generators, distributions, normalisers) and calls the
// reference's own methods, copying plain host buffers in and out so that Python (ctypes)
// can drive it side by side with the CUDA path.
#include <chrono>
#include <cstdint>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

#include <torch/torch.h>

#include "cpprl/cpprl.h"
#include "cpprl/distributions/normal.h"
#include "cpprl/generators/recurrent_generator.h"
#include "cpprl/running_mean_std.h"

#include "ref_driver.h"

using namespace cpprl;  // == cpprl_ref (macro-renamed)

namespace
{
thread_local std::string g_err;

struct RefSession
{
    RefConfig cfg;
    std::shared_ptr<NNBase> base;
    std::unique_ptr<Policy> policy;
    std::unique_ptr<RolloutStorage> storage;
    std::unique_ptr<Algorithm> algo;
    ActionSpace space;
    std::vector<int64_t> obs_shape;
};

torch::Tensor from_host(const float *p, std::vector<int64_t> shape)
{
    return torch::from_blob(const_cast<float *>(p), shape, torch::kFloat).clone();
}

void to_host(const torch::Tensor &t, float *out)
{
    auto c = t.detach().to(torch::kFloat).contiguous();
    std::memcpy(out, c.data_ptr<float>(), sizeof(float) * c.numel());
}

template <typename F>
int guarded(F &&f)
{
    try
    {
        f();
        return 0;
    }
    catch (const std::exception &e)
    {
        g_err = e.what();
        return 1;
    }
}
} // namespace

extern "C"
{
const char *ref_last_error() { return g_err.c_str(); }

const char *ref_torch_version() { return TORCH_VERSION; }

void ref_set_num_threads(int n) { at::set_num_threads(n); }
int ref_get_num_threads() { return at::get_num_threads(); }
void ref_manual_seed(uint64_t seed) { torch::manual_seed(seed); }

void *ref_create(const RefConfig *cfg, uint64_t init_seed)
{
    RefSession *s = nullptr;
    int rc = guarded([&] {
        torch::manual_seed(init_seed);
        auto sess = std::make_unique<RefSession>();
        sess->cfg = *cfg;
        for (int i = 0; i < cfg->obs_ndim; ++i)
            sess->obs_shape.push_back(cfg->obs_shape[i]);
        bool recurrent = cfg->base == REF_BASE_GRU_MLP;
        if (cfg->base == REF_BASE_CNN)
            sess->base = std::make_shared<CnnBase>(cfg->obs_shape[0], false, cfg->hidden);
        else
            sess->base = std::make_shared<MlpBase>(cfg->obs_shape[0], recurrent, cfg->hidden);
        sess->space = ActionSpace{cfg->head == REF_HEAD_CATEGORICAL ? "Discrete" : "Box",
                                  {cfg->num_actions}};
        sess->policy = std::make_unique<Policy>(sess->space, sess->base, cfg->normalize_obs != 0);
        sess->storage = std::make_unique<RolloutStorage>(cfg->T, cfg->N, sess->obs_shape, sess->space,
                                                         cfg->storage_hidden, torch::kCPU);
        if (cfg->algo == REF_ALGO_PPO)
            sess->algo = std::make_unique<PPO>(*sess->policy, cfg->clip_param, cfg->num_epoch,
                                               cfg->num_mini_batch, cfg->actor_loss_coef,
                                               cfg->value_loss_coef, cfg->entropy_coef,
                                               cfg->learning_rate, cfg->epsilon, cfg->max_grad_norm,
                                               cfg->kl_target);
        else
            sess->algo = std::make_unique<A2C>(*sess->policy, cfg->actor_loss_coef, cfg->value_loss_coef,
                                               cfg->entropy_coef, cfg->learning_rate, cfg->epsilon,
                                               cfg->alpha, cfg->max_grad_norm);
        s = sess.release();
    });
    return rc == 0 ? s : nullptr;
}

void ref_destroy(void *h) { delete static_cast<RefSession *>(h); }

int64_t ref_num_params(void *h)
{
    auto *s = static_cast<RefSession *>(h);
    int64_t n = 0;
    for (auto &p : (*s->policy)->parameters())
        n += p.numel();
    return n;
}

// Writes one line per parameter: "name numel\n" (module registration order = flat layout).
int ref_param_layout(void *h, char *buf, int buflen)
{
    auto *s = static_cast<RefSession *>(h);
    std::string out;
    for (auto &kv : (*s->policy)->named_parameters())
        out += kv.key() + " " + std::to_string(kv.value().numel()) + "\n";
    if (static_cast<int>(out.size()) + 1 > buflen)
        return -static_cast<int>(out.size()) - 1;
    std::memcpy(buf, out.c_str(), out.size() + 1);
    return static_cast<int>(out.size());
}

int ref_get_params(void *h, float *out)
{
    auto *s = static_cast<RefSession *>(h);
    return guarded([&] {
        for (auto &p : (*s->policy)->parameters())
        {
            to_host(p, out);
            out += p.numel();
        }
    });
}

int ref_set_params(void *h, const float *in)
{
    auto *s = static_cast<RefSession *>(h);
    return guarded([&] {
        torch::NoGradGuard g;
        for (auto &p : (*s->policy)->parameters())
        {
            p.copy_(from_host(in, p.sizes().vec()));
            in += p.numel();
        }
    });
}

// buffers: observation_normalizer.{clip, rms.count, rms.mean, rms.variance}
int ref_get_normalizer(void *h, float *clip, float *count, float *mean, float *var)
{
    auto *s = static_cast<RefSession *>(h);
    return guarded([&] {
        for (auto &kv : (*s->policy)->named_buffers())
        {
            const auto &k = kv.key();
            if (k == "observation_normalizer.clip")
                to_host(kv.value(), clip);
            else if (k == "observation_normalizer.rms.count")
                to_host(kv.value(), count);
            else if (k == "observation_normalizer.rms.mean")
                to_host(kv.value(), mean);
            else if (k == "observation_normalizer.rms.variance")
                to_host(kv.value(), var);
        }
    });
}

int ref_set_normalizer(void *h, const float *count, const float *mean, const float *var)
{
    auto *s = static_cast<RefSession *>(h);
    return guarded([&] {
        torch::NoGradGuard g;
        for (auto &kv : (*s->policy)->named_buffers())
        {
            const auto &k = kv.key();
            if (k == "observation_normalizer.rms.count")
                kv.value().copy_(from_host(count, {1}));
            else if (k == "observation_normalizer.rms.mean")
                kv.value().copy_(from_host(mean, kv.value().sizes().vec()));
            else if (k == "observation_normalizer.rms.variance")
                kv.value().copy_(from_host(var, kv.value().sizes().vec()));
        }
    });
}

// Fill the storage from host buffers laid out exactly as the storage keeps them
// ([T+1,N,*D], [T+1,N,Hs], [T,N,1], ...).  `actions` is int64 for Discrete, fp32 for Box.
int ref_set_rollout(void *h, const float *obs, const float *hidden, const float *rewards,
                    const float *value_preds, const float *returns, const float *logp,
                    const void *actions, const float *masks)
{
    auto *s = static_cast<RefSession *>(h);
    return guarded([&] {
        const auto &c = s->cfg;
        std::vector<int64_t> oshape{c.T + 1, c.N};
        oshape.insert(oshape.end(), s->obs_shape.begin(), s->obs_shape.end());
        s->storage->set_observations(from_host(obs, oshape));
        s->storage->set_hidden_states(from_host(hidden, {c.T + 1, c.N, c.storage_hidden}));
        s->storage->set_rewards(from_host(rewards, {c.T, c.N, 1}));
        s->storage->set_value_predictions(from_host(value_preds, {c.T + 1, c.N, 1}));
        s->storage->set_returns(from_host(returns, {c.T + 1, c.N, 1}));
        s->storage->set_action_log_probs(from_host(logp, {c.T, c.N, 1}));
        if (c.head == REF_HEAD_CATEGORICAL)
            s->storage->set_actions(
                torch::from_blob(const_cast<void *>(actions), {c.T, c.N, 1}, torch::kLong).clone());
        else
            s->storage->set_actions(from_host(static_cast<const float *>(actions), {c.T, c.N, c.num_actions}));
        s->storage->set_masks(from_host(masks, {c.T + 1, c.N, 1}));
    });
}

int ref_get_returns(void *h, float *out)
{
    auto *s = static_cast<RefSession *>(h);
    return guarded([&] { to_host(s->storage->get_returns(), out); });
}

int ref_get_value_predictions(void *h, float *out)
{
    auto *s = static_cast<RefSession *>(h);
    return guarded([&] { to_host(s->storage->get_value_predictions(), out); });
}

int ref_compute_returns(void *h, const float *next_value, int use_gae, float gamma, float tau, double *seconds)
{
    auto *s = static_cast<RefSession *>(h);
    return guarded([&] {
        auto nv = from_host(next_value, {s->cfg.N, 1});
        auto t0 = std::chrono::steady_clock::now();
        s->storage->compute_returns(nv, use_gae != 0, gamma, tau);
        auto t1 = std::chrono::steady_clock::now();
        if (seconds)
            *seconds = std::chrono::duration<double>(t1 - t0).count();
    });
}

// Runs Algorithm::update.  If perm_seed >= 0, torch::manual_seed(perm_seed) is called first so the
// generator's torch::randperm stream is reproducible from Python (same libtorch, same process).
// stats_out receives the UpdateDatum values in the reference's order; *n_stats their count.
int ref_update(void *h, float decay_level, int64_t perm_seed, float *stats_out, int *n_stats, double *seconds)
{
    auto *s = static_cast<RefSession *>(h);
    return guarded([&] {
        if (perm_seed >= 0)
            torch::manual_seed(static_cast<uint64_t>(perm_seed));
        auto t0 = std::chrono::steady_clock::now();
        auto data = s->algo->update(*s->storage, decay_level);
        auto t1 = std::chrono::steady_clock::now();
        if (seconds)
            *seconds = std::chrono::duration<double>(t1 - t0).count();
        *n_stats = static_cast<int>(data.size());
        for (size_t i = 0; i < data.size(); ++i)
            stats_out[i] = data[i].value;
    });
}

int ref_after_update(void *h)
{
    auto *s = static_cast<RefSession *>(h);
    return guarded([&] { s->storage->after_update(); });
}

// Policy::evaluate_actions on B rows (NoGrad).  hxs is [n_hxs, hidden] (n_hxs == B for step mode).
int ref_evaluate_actions(void *h, const float *obs, const float *hxs, int64_t n_hxs, const float *masks,
                         const void *actions, int64_t B, float *value, float *logp, float *entropy)
{
    auto *s = static_cast<RefSession *>(h);
    return guarded([&] {
        torch::NoGradGuard g;
        const auto &c = s->cfg;
        std::vector<int64_t> oshape{B};
        oshape.insert(oshape.end(), s->obs_shape.begin(), s->obs_shape.end());
        auto o = from_host(obs, oshape);
        int64_t hs = (*s->policy)->get_hidden_size();
        auto hx = from_host(hxs, {n_hxs, hs});
        auto m = from_host(masks, {B, 1});
        torch::Tensor a;
        if (c.head == REF_HEAD_CATEGORICAL)
            a = torch::from_blob(const_cast<void *>(actions), {B, 1}, torch::kLong).clone();
        else
            a = from_host(static_cast<const float *>(actions), {B, c.num_actions});
        auto r = (*s->policy)->evaluate_actions(o, hx, m, a);
        to_host(r[0], value);
        to_host(r[1], logp);
        to_host(r[2], entropy);
    });
}

int ref_get_values(void *h, const float *obs, const float *hxs, const float *masks, int64_t B, float *value)
{
    auto *s = static_cast<RefSession *>(h);
    return guarded([&] {
        torch::NoGradGuard g;
        std::vector<int64_t> oshape{B};
        oshape.insert(oshape.end(), s->obs_shape.begin(), s->obs_shape.end());
        int64_t hs = (*s->policy)->get_hidden_size();
        auto r = (*s->policy)->get_values(from_host(obs, oshape), from_host(hxs, {B, hs}), from_host(masks, {B, 1}));
        to_host(r, value);
    });
}

// Draws the minibatches the reference's generator would produce for the storage's current contents.
// Feed-forward: writes, per minibatch k, the 8 MiniBatch fields concatenated row-wise into the out_*
// arrays (capacity = T*N rows).  Returns number of minibatches in *n_mb and rows per minibatch in *mbs.
int ref_generate_minibatches(void *h, const float *advantages, int num_mini_batch, int64_t perm_seed,
                             float *out_obs, float *out_hidden, void *out_actions, float *out_values,
                             float *out_returns, float *out_masks, float *out_logp, float *out_adv,
                             int *n_mb, int *mbs)
{
    auto *s = static_cast<RefSession *>(h);
    return guarded([&] {
        const auto &c = s->cfg;
        if (perm_seed >= 0)
            torch::manual_seed(static_cast<uint64_t>(perm_seed));
        auto adv = from_host(advantages, {c.T, c.N, 1});
        std::unique_ptr<cpprl::Generator> gen;
        if ((*s->policy)->is_recurrent())
            gen = s->storage->recurrent_generator(adv, num_mini_batch);
        else
            gen = s->storage->feed_forward_generator(adv, num_mini_batch);
        int count = 0;
        int64_t rows = 0;
        while (!gen->done())
        {
            MiniBatch mb = gen->next();
            rows = mb.advantages.size(0);
            int64_t od = mb.observations.numel() / rows;
            to_host(mb.observations, out_obs + count * rows * od);
            int64_t hd = mb.hidden_states.numel() / mb.hidden_states.size(0);
            to_host(mb.hidden_states, out_hidden + count * mb.hidden_states.size(0) * hd);
            if (c.head == REF_HEAD_CATEGORICAL)
            {
                auto a = mb.actions.contiguous();
                std::memcpy(static_cast<int64_t *>(out_actions) + count * rows, a.data_ptr<int64_t>(),
                            sizeof(int64_t) * rows);
            }
            else
                to_host(mb.actions, static_cast<float *>(out_actions) + count * rows * c.num_actions);
            to_host(mb.value_predictions, out_values + count * rows);
            to_host(mb.returns, out_returns + count * rows);
            to_host(mb.masks, out_masks + count * rows);
            to_host(mb.action_log_probs, out_logp + count * rows);
            to_host(mb.advantages, out_adv + count * rows);
            ++count;
        }
        *n_mb = count;
        *mbs = static_cast<int>(rows);
    });
}

// ---- stand-alone pieces (known-answer helpers) ---------------------------------------------

// Categorical(logits): log_prob(actions) [B], entropy [B], probs [B,A]
int ref_categorical(const float *logits, const int64_t *actions, int64_t B, int64_t A, float *logp,
                    float *entropy, float *probs)
{
    return guarded([&] {
        auto l = from_host(logits, {B, A});
        Categorical dist(nullptr, &l);
        auto a = torch::from_blob(const_cast<int64_t *>(actions), {B}, torch::kLong).clone();
        to_host(dist.log_prob(a), logp);
        to_host(dist.entropy(), entropy);
        to_host(dist.get_probs(), probs);
    });
}

// Normal(loc [B,A], scale [A]): log_prob(value) [B,A], entropy [B]
int ref_normal(const float *loc, const float *scale, const float *value, int64_t B, int64_t A, float *logp,
               float *entropy)
{
    return guarded([&] {
        Normal dist(from_host(loc, {B, A}), from_host(scale, {A}));
        to_host(dist.log_prob(from_host(value, {B, A})), logp);
        to_host(dist.entropy(), entropy);
    });
}

// ObservationNormalizer(D, clip): state in/out (count, mean, var), optional update with x [M,D], then
// optional process_observation of y [B,D].
int ref_obs_normalizer(int64_t D, float clip, float *count, float *mean, float *var, const float *x,
                       int64_t M, const float *y, int64_t B, float *y_out)
{
    return guarded([&] {
        ObservationNormalizer norm(static_cast<int>(D), clip);
        {
            torch::NoGradGuard g;
            for (auto &kv : norm->named_buffers())
            {
                if (kv.key() == "rms.count")
                    kv.value().copy_(from_host(count, {1}));
                else if (kv.key() == "rms.mean")
                    kv.value().copy_(from_host(mean, {D}));
                else if (kv.key() == "rms.variance")
                    kv.value().copy_(from_host(var, {D}));
            }
        }
        if (x && M > 0)
            norm->update(from_host(x, {M, D}));
        if (y && B > 0)
            to_host(norm->process_observation(from_host(y, {B, D})), y_out);
        for (auto &kv : norm->named_buffers())
        {
            if (kv.key() == "rms.count")
                to_host(kv.value(), count);
            else if (kv.key() == "rms.mean")
                to_host(kv.value(), mean);
            else if (kv.key() == "rms.variance")
                to_host(kv.value(), var);
        }
    });
}
}
