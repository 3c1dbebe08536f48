// host_shim_test.cpp — exercises the C++ drop-in (namespace cpprl, include/cpprl/*.h) on a CUDA device the way the
// reference's in-source doctest cases drive the reference (shapes, error texts, the "learns a pattern" tests of
// src/algorithms/{ppo,a2c}.cpp) and checks it against the compiled reference (oracle/_ref/libcpprl_ref.so, driven
// through oracle/ref_driver.cpp's C interface) on identical rollouts and identical torch::manual_seed permutations.
// TEST CODE: built by tests/cpp/build.sh, run by tests/test_cpp_shim_gpu.py on the GPU box.
#include <cmath>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "cpprl/cpprl.h"
#include "cpprl_b200.h"
#include "oracle/ref_driver.h"

extern "C"
{
const char *ref_last_error();
void *ref_create(const RefConfig *cfg, uint64_t init_seed);
void ref_destroy(void *h);
int64_t ref_num_params(void *h);
int ref_get_params(void *h, float *out);
int ref_get_normalizer(void *h, float *clip, float *count, float *mean, float *var);
int ref_set_rollout(void *h, const float *obs, const float *hidden, const float *rewards, const float *value_preds,
                    const float *returns, const float *logp, const void *actions, const float *masks);
int ref_get_returns(void *h, float *out);
int ref_compute_returns(void *h, const float *next_value, int use_gae, float gamma, float tau, double *seconds);
int ref_update(void *h, float decay_level, int64_t perm_seed, float *stats_out, int *n_stats, double *seconds);
int ref_evaluate_actions(void *h, const float *obs, const float *hxs, int64_t n_hxs, const float *masks,
                         const void *actions, int64_t B, float *value, float *logp, float *entropy);
}

using namespace cpprl;

static int g_failed = 0, g_checks = 0;
#define CHECK(cond)                                                                  \
    do                                                                               \
    {                                                                                \
        ++g_checks;                                                                  \
        if (!(cond))                                                                 \
        {                                                                            \
            ++g_failed;                                                              \
            std::printf("  CHECK FAILED %s:%d: %s\n", __FILE__, __LINE__, #cond);     \
        }                                                                            \
    } while (0)
#define CHECK_THROWS(expr, text)                                                     \
    do                                                                               \
    {                                                                                \
        ++g_checks;                                                                  \
        bool threw = false;                                                          \
        try                                                                          \
        {                                                                            \
            expr;                                                                    \
        }                                                                            \
        catch (const std::exception &e)                                              \
        {                                                                            \
            threw = std::string(e.what()).find(text) != std::string::npos;           \
            if (!threw)                                                              \
                std::printf("  wrong message: %s\n", e.what());                      \
        }                                                                            \
        if (!threw)                                                                  \
        {                                                                            \
            ++g_failed;                                                              \
            std::printf("  CHECK_THROWS FAILED %s:%d: %s\n", __FILE__, __LINE__, #expr); \
        }                                                                            \
    } while (0)

static const torch::Device kDev(torch::kCUDA, 0);
static std::vector<int64_t> shape_of(const torch::Tensor &t) { return t.sizes().vec(); }

// ---- storage.cpp:341-437: the reference's golden vectors for compute_returns -------------------------------
static void test_compute_returns_known_answer()
{
    std::printf("compute_returns known answers\n");
    for (int use_gae = 0; use_gae < 2; ++use_gae)
    {
        RolloutStorage storage(3, 2, {4}, ActionSpace{"Discrete", {3}}, 5, kDev);
        float vp[3][2] = {{0, 1}, {1, 2}, {2, 3}}, rw[3][2] = {{0, 1}, {1, 2}, {2, 3}}, mk[3][2] = {{1, 1}, {1, 0}, {1, 1}};
        for (int t = 0; t < 3; ++t)
            storage.insert(torch::zeros({2, 4}), torch::zeros({2, 5}), torch::zeros({2, 1}, torch::kLong),
                           torch::zeros({2, 1}), torch::from_blob(vp[t], {2, 1}).clone(),
                           torch::from_blob(rw[t], {2, 1}).clone(), torch::from_blob(mk[t], {2, 1}).clone());
        float nv[2] = {0, 1};
        storage.compute_returns(torch::from_blob(nv, {2, 1}).clone().to(kDev), use_gae, 0.6, 0.6);
        auto r = storage.get_returns().cpu();
        const float want_gae[4][2] = {{1.032, 2.2}, {2.2, 2.0}, {2.0, 3.6}, {0, 0}};
        const float want_n[4][2] = {{1.32, 2.2}, {2.2, 2.0}, {2.0, 3.6}, {0, 1}};
        for (int t = 0; t < 4; ++t)
            for (int e = 0; e < 2; ++e)
                CHECK(std::fabs(r[t][e][0].item().toFloat() - (use_gae ? want_gae : want_n)[t][e]) < 1e-5);
    }
}

// ---- generator shapes and exception texts (feed_forward_generator.cpp:82-121, recurrent_generator.cpp:100-137) ----
static void test_generators()
{
    std::printf("generators\n");
    RolloutStorage storage(3, 5, {5, 2}, ActionSpace{"Discrete", {3}}, 10, kDev);
    auto adv = torch::rand({3, 5, 1}, torch::TensorOptions().device(kDev));
    auto gen = storage.feed_forward_generator(adv, 5);
    auto mb = gen->next();
    CHECK((shape_of(mb.observations) == std::vector<int64_t>{3, 5, 2}));
    CHECK((shape_of(mb.hidden_states) == std::vector<int64_t>{3, 10}));
    CHECK((shape_of(mb.actions) == std::vector<int64_t>{3, 1}));
    CHECK((shape_of(mb.advantages) == std::vector<int64_t>{3, 1}));
    CHECK(mb.actions.dtype() == torch::kLong);
    for (int i = 0; i < 4; ++i)
    {
        CHECK(!gen->done());
        gen->next();
    }
    CHECK(gen->done());
    CHECK_THROWS(gen->next(), "No minibatches left in generator.");
    CHECK_THROWS(storage.feed_forward_generator(adv, 16), "to be greater than or equal to the number of minibatches (16)");
    auto rgen = storage.recurrent_generator(adv, 5);
    auto rmb = rgen->next();
    CHECK((shape_of(rmb.observations) == std::vector<int64_t>{3, 5, 2}));
    CHECK((shape_of(rmb.hidden_states) == std::vector<int64_t>{1, 10}));
    CHECK((shape_of(rmb.masks) == std::vector<int64_t>{3, 1}));
    CHECK_THROWS(storage.recurrent_generator(adv, 6), "PPO needs the number of processes (5)");
    // gathered rows are bit-exact copies of rows t*N + env
    torch::manual_seed(3);
    RolloutStorage s2(4, 6, {7}, ActionSpace{"Box", {2}}, 1, kDev);
    s2.set_observations(torch::randn({5, 6, 7}, torch::TensorOptions().device(kDev)));
    auto adv2 = torch::randn({4, 6, 1}, torch::TensorOptions().device(kDev));
    FeedForwardGenerator g2(8, s2.get_observations(), s2.get_hidden_states(), s2.get_actions(), s2.get_value_predictions(),
                            s2.get_returns(), s2.get_masks(), s2.get_action_log_probs(), adv2);
    auto idx = g2.get_indices();
    auto m2 = g2.next();
    auto flat_obs = s2.get_observations().narrow(0, 0, 4).reshape({24, 7});
    CHECK(torch::equal(m2.observations, flat_obs.index({idx[0].to(kDev)})));
    CHECK(torch::equal(m2.advantages, adv2.reshape({24, 1}).index({idx[0].to(kDev)})));
}

// ---- the reference's storage scenarios on the device (storage.cpp:298-339, 438-488, 504-581) -----------------------
static void test_storage_write_side()
{
    std::printf("storage write side\n");
    const uint64_t before = cpprl_b200_launch_count();
    { // insert(): CPU inputs of convertible dtypes land on rows step+1 / step of the device buffers, ONE launch
        RolloutStorage storage(3, 4, {5, 2}, ActionSpace{"Discrete", {3}}, 10, kDev);
        auto obs = torch::rand({4, 5, 2}) + 1, hid = torch::rand({4, 10}) + 1;
        auto act = torch::randint(1, 3, {4, 1}), lp = torch::rand({4, 1}) + 1, v = torch::rand({4, 1}) + 1, r = torch::rand({4, 1}) + 1;
        const uint64_t l0 = cpprl_b200_launch_count();
        storage.insert(obs, hid, act, lp, v, r, torch::zeros({4, 1}));
        CHECK(cpprl_b200_launch_count() - l0 == 1);
        CHECK(torch::equal(storage.get_observations()[1].cpu(), obs));
        CHECK(torch::equal(storage.get_hidden_states()[1].cpu(), hid));
        CHECK(torch::equal(storage.get_actions()[0].cpu(), act.to(torch::kLong)));
        CHECK(torch::equal(storage.get_action_log_probs()[0].cpu(), lp));
        CHECK(torch::equal(storage.get_value_predictions()[0].cpu(), v));
        CHECK(torch::equal(storage.get_rewards()[0].cpu(), r));
        CHECK(storage.get_masks()[1].sum().item().toFloat() == 0.f);
        CHECK(storage.get_masks()[0].sum().item().toFloat() == 4.f);
        CHECK(storage.get_observations()[0].abs().sum().item().toFloat() == 0.f);
    }
    { // after_update(): expected 6 / 2 / 0 at the positions the reference probes
        RolloutStorage storage(3, 2, {3}, ActionSpace{"Discrete", {3}}, 2, kDev);
        float obs[3][6] = {{0, 1, 2, 1, 2, 3}, {0, 1, 2, 1, 2, 3}, {5, 6, 7, 7, 8, 9}};
        float hid[3][4] = {{0, 1, 0, 1}, {0, 1, 0, 1}, {1, 2, 3, 4}}, msk[3][2] = {{0, 1}, {0, 1}, {0, 0}};
        for (int t = 0; t < 3; ++t)
            storage.insert(torch::from_blob(obs[t], {2, 3}).clone(), torch::from_blob(hid[t], {2, 2}).clone(), torch::zeros({2, 1}),
                           torch::zeros({2, 1}), torch::zeros({2, 1}), torch::zeros({2, 1}), torch::from_blob(msk[t], {2, 1}).clone());
        storage.after_update();
        CHECK(storage.get_observations()[0][0][1].item().toFloat() == 6.f);
        CHECK(storage.get_hidden_states()[0][0][1].item().toFloat() == 2.f);
        CHECK(storage.get_masks()[0][0][0].item().toFloat() == 0.f);
    }
    { // the concatenating constructor: five one-environment storages, column i == storage i in every field
        std::vector<RolloutStorage> storages;
        for (int i = 0; i < 5; ++i)
        {
            storages.push_back({3, 1, {4}, ActionSpace{"Discrete", {3}}, 5, kDev});
            float vals[3] = {1, 2, 3}, msk[3] = {1, 0, 1};
            for (int t = 0; t < 3; ++t)
                storages[i].insert(torch::rand({1, 4}), torch::rand({1, 5}), torch::rand({1, 1}), torch::rand({1, 1}),
                                   torch::full({1, 1}, vals[t]), torch::full({1, 1}, vals[t]), torch::full({1, 1}, msk[t]));
        }
        std::vector<RolloutStorage *> pointers;
        for (auto &st : storages)
            pointers.push_back(&st);
        const uint64_t l0 = cpprl_b200_launch_count();
        RolloutStorage combined(pointers, kDev);
        CHECK(cpprl_b200_launch_count() - l0 == 5); // one launch per part, all eight fields
        CHECK((shape_of(combined.get_observations()) == std::vector<int64_t>{4, 5, 4}));
        CHECK((shape_of(combined.get_hidden_states()) == std::vector<int64_t>{4, 5, 5}));
        CHECK((shape_of(combined.get_rewards()) == std::vector<int64_t>{3, 5, 1}));
        CHECK((shape_of(combined.get_actions()) == std::vector<int64_t>{3, 5, 1}));
        CHECK(combined.get_actions().dtype() == torch::kLong);
        for (int i = 0; i < 5; ++i)
        {
            CHECK(torch::equal(combined.get_observations().narrow(1, i, 1), storages[i].get_observations()));
            CHECK(torch::equal(combined.get_hidden_states().narrow(1, i, 1), storages[i].get_hidden_states()));
            CHECK(torch::equal(combined.get_rewards().narrow(1, i, 1), storages[i].get_rewards()));
            CHECK(torch::equal(combined.get_value_predictions().narrow(1, i, 1), storages[i].get_value_predictions()));
            CHECK(torch::equal(combined.get_returns().narrow(1, i, 1), storages[i].get_returns()));
            CHECK(torch::equal(combined.get_action_log_probs().narrow(1, i, 1), storages[i].get_action_log_probs()));
            CHECK(torch::equal(combined.get_actions().narrow(1, i, 1), storages[i].get_actions()));
            CHECK(torch::equal(combined.get_masks().narrow(1, i, 1), storages[i].get_masks()));
        }
    }
    CHECK(cpprl_b200_launch_count() > before);
}

// ---- combine constructor, checkpoint round trip, reward normaliser ------------------------------
static void test_combine_checkpoint_reward_normaliser()
{
    std::printf("normaliser combine / checkpoint / reward normaliser\n");
    auto opts = torch::TensorOptions().device(kDev);
    { // observation_normalizer.cpp:145-190: 1, 2 and 3 device updates; unweighted means, counts summed to 6
        std::vector<ObservationNormalizer> normalizers;
        for (int i = 0; i < 3; ++i)
        {
            normalizers.push_back(ObservationNormalizer(3));
            normalizers[i]->to(kDev);
            for (int j = 0; j <= i; ++j)
                normalizers[i]->update(torch::rand({3}).to(kDev));
        }
        ObservationNormalizer combined(normalizers);
        auto mean = combined->get_mean(), var = combined->get_variance();
        for (int k = 0; k < 3; ++k)
        {
            float m = 0, v = 0;
            for (auto &n : normalizers)
                m += n->get_mean()[k], v += n->get_variance()[k];
            CHECK(std::fabs(mean[k] - m / 3) <= 1e-6f);
            CHECK(std::fabs(var[k] - v / 3) <= 1e-6f);
        }
        CHECK(combined->get_step_count() == 6);
        CHECK(std::fabs(combined->get_clip_value() - 10.f) <= 1e-6f);
        // the combined normaliser is usable on the device after ->to()
        combined->to(kDev);
        auto y = combined->process_observation(torch::rand({4, 3}, opts));
        CHECK((shape_of(y) == std::vector<int64_t>{4, 3}));
    }
    { // checkpoint: the kernels step the parameters and normaliser buffers IN PLACE; torch::save / torch::load must
      // see those values (gym_client-style resume), and the restored policy must evaluate identically
        torch::manual_seed(11);
        ActionSpace space{"Box", {3}};
        Policy policy(space, std::make_shared<MlpBase>(6, false, 64), true);
        policy->to(kDev);
        RolloutStorage storage(8, 16, {6}, space, 1, kDev);
        PPO ppo(policy, 0.2, 2, 2, 1, 0.5, 1e-3, 1e-3, 1e-5, 0.5, 0.01);
        auto obs = torch::randn({16, 6}, opts);
        storage.set_first_observation(obs);
        for (int t = 0; t < 8; ++t)
        {
            auto act = policy->act(obs, torch::Tensor(), torch::ones({16, 1}, opts));
            obs = torch::randn({16, 6}, opts);
            storage.insert(obs, torch::zeros({16, 1}, opts), act[1], act[2], act[0], torch::randn({16, 1}, opts), torch::ones({16, 1}, opts));
        }
        storage.compute_returns(policy->get_values(storage.get_observations()[-1], storage.get_hidden_states()[-1], storage.get_masks()[-1]),
                                true, 0.99, 0.95);
        auto before = policy->parameters()[0].clone();
        ppo.update(storage);
        CHECK(!torch::equal(before, policy->parameters()[0])); // the update moved the torch-visible tensors
        const std::string path = "/tmp/cpprl_b200_policy_checkpoint.pt";
        torch::save(policy, path);
        torch::manual_seed(12); // different initial weights
        Policy restored(space, std::make_shared<MlpBase>(6, false, 64), true);
        restored->to(kDev);
        torch::load(restored, path);
        restored->to(kDev);
        auto pa = policy->named_parameters(), pb = restored->named_parameters();
        CHECK(pa.size() == pb.size());
        for (auto &kv : pa)
            CHECK(torch::equal(kv.value().cpu(), pb[kv.key()].cpu()));
        auto ba = policy->named_buffers(), bb = restored->named_buffers();
        for (auto &kv : ba)
            CHECK(torch::equal(kv.value().cpu(), bb[kv.key()].cpu()));
        CHECK(ba["observation_normalizer.rms.count"].item().toFloat() > 100.f); // 9 x 16 rows went through the kernel
        auto x = torch::randn({32, 6}, opts), a = torch::randn({32, 3}, opts);
        auto ea = policy->evaluate_actions(x, torch::Tensor(), torch::Tensor(), a);
        auto eb = restored->evaluate_actions(x, torch::Tensor(), torch::Tensor(), a);
        CHECK(torch::equal(ea[0], eb[0]));
        CHECK(torch::equal(ea[1], eb[1]));
        std::remove(path.c_str());
    }
    { // reward normaliser as the example driver uses it (gym_client.cpp:167,217-220): RunningMeanStd(1) over the
      // discounted returns of all environments, on the device
        const int num_envs = 32;
        RunningMeanStd returns_rms(1);
        returns_rms->to(kDev);
        auto returns = torch::zeros({num_envs}, opts);
        double count = 1e-4, mean = 0, var = 1; // float64 restatement of running_mean_std.cpp:20-43
        torch::manual_seed(13);
        for (int step = 0; step < 20; ++step)
        {
            auto reward = torch::randn({num_envs}) * 3 + 1;
            returns = returns * 0.99 + reward.to(kDev);
            returns_rms->update(returns);
            auto r = returns.cpu().to(torch::kDouble);
            const double bm = r.mean().item().toDouble(), bv = r.var(false).item().toDouble(), bc = num_envs;
            const double delta = bm - mean, tot = count + bc;
            const double m2 = var * count + bv * bc + delta * delta * count * bc / tot;
            mean += delta * bc / tot;
            var = m2 / tot;
            count = tot;
        }
        CHECK(std::fabs(returns_rms->get_mean().item().toFloat() - mean) <= 1e-5 * std::fabs(mean) + 1e-6);
        CHECK(std::fabs(returns_rms->get_variance().item().toFloat() - var) <= 1e-5 * var);
        CHECK(returns_rms->get_count() == 20 * num_envs);
    }
}

// ---- observation_normalizer.cpp:95-129 ----------------------------------------------------------------------------
static void test_normalizer()
{
    std::printf("observation normaliser\n");
    ObservationNormalizer normalizer(5);
    normalizer->to(kDev);
    float rows[3][5] = {{-10., 0., 5., 3.2, 0.}, {-5., 2., 4., 3.7, -3.}, {1, 2, 3, 4, 5}};
    for (auto &row : rows)
        normalizer->update(torch::from_blob(row, {5}).clone().to(kDev));
    auto out = normalizer->process_observation(torch::from_blob(rows[2], {5}).clone().to(kDev)).cpu();
    const float want[5] = {1.26008659, 0.70712887, -1.2240818, 1.10914509, 1.31322402};
    for (int i = 0; i < 5; ++i)
        CHECK(std::fabs(out[i].item().toFloat() - want[i]) < 1e-4);
    ObservationNormalizer clipper(7, 1);
    clipper->to(kDev);
    float big[7] = {-1000, -100, -10, 0, 10, 100, 1000};
    auto clipped = clipper->process_observation(torch::from_blob(big, {7}).clone().to(kDev));
    CHECK(clipped.max().item().toFloat() <= 1.f && clipped.min().item().toFloat() >= -1.f);
    CHECK(normalizer->get_step_count() == 3);
}

// ---- policy.cpp:152-360 (shapes) and the parameter / buffer names of SURVEY.md §8b ------------------------------------
static void test_policy_surface()
{
    std::printf("policy surface\n");
    auto base = std::make_shared<MlpBase>(3, false, 64);
    Policy policy(ActionSpace{"Discrete", {5}}, base, true);
    policy->to(kDev);
    std::vector<std::string> names;
    for (const auto &p : policy->named_parameters())
        names.push_back(p.key());
    const std::vector<std::string> want{"base.actor.0.weight", "base.actor.0.bias", "base.actor.2.weight", "base.actor.2.bias",
                                        "base.critic.0.weight", "base.critic.0.bias", "base.critic.2.weight",
                                        "base.critic.2.bias", "base.critic_linear.weight", "base.critic_linear.bias",
                                        "output.linear.weight", "output.linear.bias"};
    CHECK(names == want);
    std::vector<std::string> buffers;
    for (const auto &b : policy->named_buffers())
        buffers.push_back(b.key());
    CHECK((buffers == std::vector<std::string>{"observation_normalizer.clip", "observation_normalizer.rms.count",
                                               "observation_normalizer.rms.mean", "observation_normalizer.rms.variance"}));
    auto opts = torch::TensorOptions().device(kDev);
    auto ev = policy->evaluate_actions(torch::rand({4, 3}, opts), torch::rand({4, 1}, opts), torch::ones({4, 1}, opts),
                                       torch::randint(5, {4, 1}, opts.dtype(torch::kLong)));
    CHECK((shape_of(ev[0]) == std::vector<int64_t>{4, 1}));
    CHECK((shape_of(ev[1]) == std::vector<int64_t>{4, 1}));
    CHECK(ev[2].dim() == 0);
    auto act = policy->act(torch::rand({4, 3}, opts), torch::rand({4, 1}, opts), torch::ones({4, 1}, opts));
    CHECK((shape_of(act[0]) == std::vector<int64_t>{4, 1}));
    CHECK((shape_of(act[1]) == std::vector<int64_t>{4, 1}) && act[1].dtype() == torch::kLong);
    CHECK((shape_of(act[2]) == std::vector<int64_t>{4, 1}));
    CHECK((shape_of(policy->get_probs(torch::rand({4, 3}, opts), torch::rand({4, 1}, opts), torch::ones({4, 1}, opts))) ==
           std::vector<int64_t>{4, 5}));
    CHECK((shape_of(policy->get_values(torch::rand({4, 3}, opts), torch::rand({4, 1}, opts), torch::ones({4, 1}, opts))) ==
           std::vector<int64_t>{4, 1}));
    CHECK(policy->get_hidden_size() == 1 && !policy->is_recurrent() && policy->using_observation_normalizer());
    // the fused forward and the module forward read the same memory: log-probs agree
    {
        Policy plain(ActionSpace{"Box", {2}}, std::make_shared<MlpBase>(6, false, 64), false);
        plain->to(kDev);
        auto x = torch::randn({33, 6}, opts), a = torch::randn({33, 2}, opts);
        auto fused = plain->evaluate_actions(x, torch::Tensor(), torch::Tensor(), a);
        auto out = std::static_pointer_cast<NNBase>(plain->named_children()["base"])->forward(x, torch::Tensor(), torch::Tensor());
        CHECK(torch::allclose(fused[0], out[0], 1e-4, 1e-5));
        // parameters stay aliased to the flat buffer after the call, and survive a device round trip
        auto flat = plain->flat_parameters();
        CHECK(plain->parameters()[0].data_ptr<float>() == flat.data_ptr<float>());
        plain->to(torch::kCPU);
        plain->to(kDev);
        auto again = plain->evaluate_actions(x, torch::Tensor(), torch::Tensor(), a);
        CHECK(torch::allclose(again[1], fused[1], 1e-6, 1e-6));
    }
    CHECK_THROWS(Policy(ActionSpace{"Continuous", {2}}, std::make_shared<MlpBase>(3, false, 64)), "Action space Continuous not supported");
    // no CPU fallback
    Policy cpu_policy(ActionSpace{"Discrete", {2}}, std::make_shared<MlpBase>(3, false, 64), false);
    CHECK_THROWS(cpu_policy->evaluate_actions(torch::rand({4, 3}), torch::Tensor(), torch::Tensor(), torch::zeros({4, 1}, torch::kLong)),
                 "no CPU fallback");
}

// ---- the reference's directional learning tests (ppo.cpp:180-367, a2c.cpp:92-279) on the device --------------------------
template <typename Algo>
static void play(Policy &policy, RolloutStorage &storage, Algo &algo, bool game)
{
    auto opts = torch::TensorOptions().device(kDev);
    auto observation = torch::randint(0, 2, {2, 1}, torch::kFloat).to(kDev);
    if (game)
        storage.set_first_observation(observation);
    for (int i = 0; i < 10; ++i)
    {
        for (int j = 0; j < 5; ++j)
        {
            if (!game)
                observation = torch::randint(0, 2, {2, 1}, torch::kFloat).to(kDev);
            auto act = policy->act(observation, torch::Tensor(), torch::ones({2, 1}, opts));
            auto actions = act[1];
            torch::Tensor rewards;
            if (game)
            { // reward +1 when the action matches the observation, else -1
                rewards = (actions == observation.to(torch::kLong)).to(torch::kFloat) * 2 - 1;
                observation = torch::randint(0, 2, {2, 1}, torch::kFloat).to(kDev);
            }
            else // the reward is the action
                rewards = actions.to(torch::kFloat);
            storage.insert(observation, torch::zeros({2, 5}, opts), actions, act[2], act[0], rewards, torch::ones({2, 1}, opts));
        }
        auto next_value = policy->get_values(storage.get_observations()[-1], storage.get_hidden_states()[-1],
                                             storage.get_masks()[-1]);
        storage.compute_returns(next_value, false, game ? 0.1 : 0., 0.9);
        algo.update(storage);
        storage.after_update();
    }
}

static void test_learning()
{
    std::printf("learning tests\n");
    auto opts = torch::TensorOptions().device(kDev);
    for (int algo_kind = 0; algo_kind < 2; ++algo_kind)
        for (int game = 0; game < 2; ++game)
            for (int normalize = 0; normalize <= game; ++normalize)
            {
                torch::manual_seed(0);
                ActionSpace space{"Discrete", {2}};
                Policy policy(space, std::make_shared<MlpBase>(1, false, 64), normalize != 0);
                policy->to(kDev);
                RolloutStorage storage(algo_kind == 0 ? 20 : 5, 2, {1}, space, 5, kDev);
                auto probe = [&] { return policy->get_probs(torch::ones({2, 1}, opts), torch::zeros({2, 5}, opts), torch::ones({2, 1}, opts)).cpu(); };
                auto before = probe();
                if (algo_kind == 0)
                {
                    PPO ppo(policy, 0.2, 3, 5, 1, 0.5, 1e-3, 0.001);
                    play(policy, storage, ppo, game != 0);
                }
                else
                {
                    A2C a2c(policy, 1.0, 0.5, 1e-3, 0.001);
                    play(policy, storage, a2c, game != 0);
                }
                auto after = probe();
                std::printf("  %s %s norm=%d: P(action 1 | obs 1) %.4f -> %.4f\n", algo_kind == 0 ? "PPO" : "A2C",
                            game ? "game" : "pattern", normalize, before[0][1].item().toFloat(), after[0][1].item().toFloat());
                CHECK(after[0][0].item().toFloat() < before[0][0].item().toFloat());
                CHECK(after[0][1].item().toFloat() > before[0][1].item().toFloat());
            }
}

// ---- parity with the compiled reference on identical rollouts and identical torch::manual_seed permutations ----------------
// base: REF_BASE_MLP, REF_BASE_GRU_MLP (nn_base.cpp:47-114 through recurrent_generator.cpp) or REF_BASE_CNN (cnn_base.cpp)
static void test_parity(bool ppo_algo, bool box, int base = REF_BASE_MLP)
{
    const bool gru = base == REF_BASE_GRU_MLP, cnn = base == REF_BASE_CNN;
    std::printf("parity vs compiled reference: %s %s %s\n", ppo_algo ? "PPO" : "A2C", box ? "Box/Normal" : "Discrete/Categorical",
                gru ? "GRU + MlpBase" : (cnn ? "CnnBase" : "MlpBase"));
    const int T = cnn ? 8 : (gru ? 16 : (ppo_algo ? 32 : 5)), N = cnn || gru ? 8 : (ppo_algo ? 8 : 64);
    const int C = 4, D = cnn ? C * 84 * 84 : (box ? 17 : 8), A = box ? 6 : 4, Hs = gru ? 64 : 1;
    RefConfig cfg{};
    cfg.algo = ppo_algo ? REF_ALGO_PPO : REF_ALGO_A2C;
    cfg.base = base;
    cfg.head = box ? REF_HEAD_NORMAL : REF_HEAD_CATEGORICAL;
    cfg.T = T, cfg.N = N, cfg.num_actions = A;
    if (cnn)
        cfg.obs_ndim = 3, cfg.obs_shape[0] = C, cfg.obs_shape[1] = 84, cfg.obs_shape[2] = 84;
    else
        cfg.obs_ndim = 1, cfg.obs_shape[0] = D;
    cfg.hidden = cnn ? 512 : 64, cfg.storage_hidden = Hs, cfg.normalize_obs = cnn ? 0 : 1;
    cfg.num_epoch = cnn ? 2 : 3, cfg.num_mini_batch = gru ? N : (cnn ? 2 : 4); // recurrent: one environment per minibatch
    cfg.clip_param = 0.2f, cfg.actor_loss_coef = 1.f, cfg.value_loss_coef = 0.5f, cfg.entropy_coef = 0.01f;
    cfg.learning_rate = ppo_algo ? 3e-4f : 7e-4f, cfg.epsilon = 1e-8f, cfg.alpha = 0.99f, cfg.max_grad_norm = 0.5f;
    cfg.kl_target = 1e9f;
    void *ref = ref_create(&cfg, 7);
    if (!ref)
    {
        std::printf("  ref_create failed: %s\n", ref_last_error());
        ++g_failed;
        return;
    }
    const int64_t P = ref_num_params(ref);
    std::vector<float> params(P);
    ref_get_params(ref, params.data());

    ActionSpace space{box ? "Box" : "Discrete", {A}};
    std::shared_ptr<NNBase> nn_base;
    if (cnn)
        nn_base = std::make_shared<CnnBase>(C, false, 512);
    else
        nn_base = std::make_shared<MlpBase>(D, gru, 64);
    Policy policy(space, nn_base, !cnn);
    policy->to(kDev);
    CHECK(policy->flat_parameters().numel() == P);
    policy->flat_parameters().copy_(torch::from_blob(params.data(), {P}).clone());

    torch::manual_seed(11);
    auto obs = cnn ? torch::floor(torch::rand({T + 1, N, C, 84, 84}) * 256) : torch::randn({T + 1, N, D});
    auto rewards = torch::randn({T, N, 1});
    auto masks = (torch::rand({T + 1, N, 1}) < 0.95).to(torch::kFloat);
    masks[0].fill_(1);
    auto hidden = torch::zeros({T + 1, N, Hs}), returns = torch::zeros({T + 1, N, 1}), next_value = torch::randn({N, 1});
    torch::Tensor actions = box ? torch::randn({T, N, A}) : torch::randint(A, {T, N, 1}, torch::kLong);
    auto values = torch::zeros({T + 1, N, 1}), logp = torch::zeros({T, N, 1});
    {
        auto flat_obs = obs.narrow(0, 0, T).reshape({T * N, D}).contiguous();
        (void)D;
        // recurrent: sequence mode, N initial states and the rollout's masks (nn_base.cpp:57-112)
        auto hxs = gru ? torch::zeros({N, Hs}) : torch::zeros({T * N, 1});
        auto ones = gru ? masks.narrow(0, 0, T).reshape({T * N, 1}).contiguous() : torch::ones({T * N, 1});
        auto v = torch::zeros({T * N, 1}), lp = torch::zeros({T * N, 1});
        float ent = 0;
        auto act_flat = actions.reshape({T * N, -1}).contiguous();
        CHECK(ref_evaluate_actions(ref, flat_obs.data_ptr<float>(), hxs.data_ptr<float>(), hxs.size(0), ones.data_ptr<float>(),
                                   act_flat.data_ptr(), T * N, v.data_ptr<float>(), lp.data_ptr<float>(), &ent) == 0);
        values.narrow(0, 0, T).copy_(v.view({T, N, 1}));
        logp.copy_(lp.view({T, N, 1}));
        // the drop-in's own forward agrees with the reference's on the same weights
        auto mine = gru ? policy->evaluate_actions(flat_obs.to(kDev), hxs.to(kDev), ones.to(kDev), act_flat.to(kDev))
                        : policy->evaluate_actions(flat_obs.to(kDev), torch::Tensor(), torch::Tensor(), act_flat.to(kDev));
        CHECK(torch::allclose(mine[0].cpu(), v, 1e-4, 2e-5));
        CHECK(torch::allclose(mine[1].cpu(), lp, 1e-4, 2e-5));
        CHECK(std::fabs(mine[2].item().toFloat() - ent) <= 1e-4 * std::fabs(ent));
    }
    ref_set_rollout(ref, obs.data_ptr<float>(), hidden.data_ptr<float>(), rewards.data_ptr<float>(), values.data_ptr<float>(),
                    returns.data_ptr<float>(), logp.data_ptr<float>(), actions.data_ptr(), masks.data_ptr<float>());
    RolloutStorage storage(T, N, cnn ? std::vector<int64_t>{C, 84, 84} : std::vector<int64_t>{D}, space, Hs, kDev);
    storage.set_observations(obs.to(kDev));
    storage.set_rewards(rewards.to(kDev));
    storage.set_masks(masks.to(kDev));
    storage.set_actions(actions.to(kDev));
    storage.set_value_predictions(values.to(kDev));
    storage.set_action_log_probs(logp.to(kDev));

    PPO ppo(policy, cfg.clip_param, cfg.num_epoch, cfg.num_mini_batch, cfg.actor_loss_coef, cfg.value_loss_coef,
            cfg.entropy_coef, cfg.learning_rate, cfg.epsilon, cfg.max_grad_norm, cfg.kl_target);
    A2C a2c(policy, 1.f, cfg.value_loss_coef, cfg.entropy_coef, cfg.learning_rate, cfg.epsilon, cfg.alpha);
    for (int upd = 0; upd < 2; ++upd)
    {
        const bool gae = ppo_algo;
        ref_compute_returns(ref, next_value.data_ptr<float>(), gae, 0.99f, 0.95f, nullptr);
        storage.compute_returns(next_value.to(kDev), gae, 0.99f, 0.95f);
        auto ref_returns = torch::zeros({T + 1, N, 1});
        ref_get_returns(ref, ref_returns.data_ptr<float>());
        CHECK(torch::equal(storage.get_returns().cpu(), ref_returns)); // bit-exact
        float want[8];
        int n_want = 0;
        CHECK(ref_update(ref, 1.f, 500 + upd, want, &n_want, nullptr) == 0);
        torch::manual_seed(500 + upd); // same CPU RNG stream for the generators' randperm
        auto got = ppo_algo ? ppo.update(storage) : a2c.update(storage);
        CHECK(static_cast<int>(got.size()) == n_want);
        for (int i = 0; i < n_want && i < static_cast<int>(got.size()); ++i)
        {
            // (CnnBase / GRU: later updates inherit the near-zero-gradient parameter tail of the first, tests/test_cnn_gpu.py)
            const float tol = ((cnn || gru) && upd > 0 ? 1e-3f : 1e-4f) * std::fabs(want[i]) * (i == 4 && (cnn || gru) ? 10.f : 1.f) + 5e-6f;
            if (std::fabs(got[i].value - want[i]) > tol)
                std::printf("  stat %s: got %.8g want %.8g\n", got[i].name.c_str(), got[i].value, want[i]);
            CHECK(std::fabs(got[i].value - want[i]) <= tol);
        }
        ref_get_params(ref, params.data());
        auto mine = policy->flat_parameters().cpu();
        auto want_p = torch::from_blob(params.data(), {P});
        auto rel = (mine - want_p).abs() / torch::max(want_p.abs(), torch::full({P}, cfg.learning_rate));
        const float err = rel.max().item().toFloat();
        const float q99 = std::get<0>(rel.kthvalue(static_cast<int64_t>(0.99 * (P - 1)) + 1)).item().toFloat();
        std::printf("  update %d: parameter error relative to max(|ref|, lr): max %.3e, 99th percentile %.3e\n", upd, err, q99);
        // MlpBase: every parameter; CnnBase / GRU: the bulk to 1e-4, the sign-like Adam steps of near-zero gradients bounded
        CHECK(cnn || gru ? (q99 <= 1e-4f && err <= 0.5f) : err <= 1e-3f);
        if (cnn)
            continue; // no observation normaliser on image observations
        float clip, count;
        std::vector<float> mean(D), var(D);
        ref_get_normalizer(ref, &clip, &count, mean.data(), var.data());
        auto buffers = policy->named_buffers();
        CHECK(std::fabs(buffers["observation_normalizer.rms.count"].item().toFloat() - count) <= 1e-6f * count);
        CHECK(torch::allclose(buffers["observation_normalizer.rms.mean"].cpu(), torch::from_blob(mean.data(), {D}), 1e-5, 1e-6));
        CHECK(torch::allclose(buffers["observation_normalizer.rms.variance"].cpu(), torch::from_blob(var.data(), {D}), 1e-5, 1e-6));
    }
    ref_destroy(ref);
}

int main()
{
    if (!torch::cuda::is_available())
    {
        std::printf("no CUDA device: the drop-in has no CPU fallback\n");
        return 2;
    }
    // The reference's CPU convolutions partition their sums by thread count, and a pre-activation within rounding of zero
    // lands on different sides of the ReLU with a different partition (the reference differs from
    // itself between 1 and 16 host threads the same way).  The CnnBase parity bounds below were set with 16 threads
    // (the 1-GPU box's core count); fix that count so that boxes with more cores check the same thing.
    at::set_num_threads(16);
    test_compute_returns_known_answer();
    test_generators();
    test_storage_write_side();
    test_normalizer();
    test_combine_checkpoint_reward_normaliser();
    test_policy_surface();
    test_learning();
    test_parity(true, false);
    test_parity(true, true);
    test_parity(false, false);
    test_parity(true, false, REF_BASE_GRU_MLP);
    test_parity(true, false, REF_BASE_CNN);
    std::printf("%d checks, %d failed\n", g_checks, g_failed);
    return g_failed == 0 ? 0 : 1;
}
