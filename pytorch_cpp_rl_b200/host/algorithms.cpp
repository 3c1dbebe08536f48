// algorithms.cpp — PPO::update / A2C::update over the fused CUDA update (cpprl_b200_ppo_* / cpprl_b200_a2c_*).
// Reference behaviour: src/algorithms/ppo.cpp:18-178, src/algorithms/a2c.cpp:16-90.
#include "host_common.h"

namespace cpprl
{
using detail::check;

PPO::PPO(Policy &policy, float clip_param, int num_epoch, int num_mini_batch, float actor_loss_coef,
         float value_loss_coef, float entropy_coef, float learning_rate, float epsilon, float max_grad_norm,
         float kl_target)
    : policy(policy), actor_loss_coef(actor_loss_coef), value_loss_coef(value_loss_coef), entropy_coef(entropy_coef),
      max_grad_norm(max_grad_norm), original_learning_rate(learning_rate), original_clip_param(clip_param),
      kl_target(kl_target), epsilon(epsilon), num_epoch(num_epoch), num_mini_batch(num_mini_batch)
{
}

PPO::~PPO() { cpprl_b200_ppo_destroy(handle); }

std::vector<UpdateDatum> PPO::update(RolloutStorage &rollouts, float decay_level)
{
    cpprl_b200_policy pd;
    policy->describe(pd);
    cpprl_b200_rollout rd;
    rollouts.describe(rd);
    if (handle == nullptr || handle_params != pd.n_params)
    {
        cpprl_b200_ppo_destroy(handle);
        handle = nullptr;
        check(cpprl_b200_ppo_create(&handle, pd.n_params));
        handle_params = pd.n_params;
    }
    cpprl_b200_ppo_options opt{};
    opt.clip_param = original_clip_param * decay_level;       // ppo.cpp:47
    opt.learning_rate = original_learning_rate * decay_level; // ppo.cpp:48
    opt.num_epoch = num_epoch;
    opt.num_mini_batch = num_mini_batch;
    opt.actor_loss_coef = actor_loss_coef;
    opt.value_loss_coef = value_loss_coef;
    opt.entropy_coef = entropy_coef;
    opt.epsilon = epsilon;
    opt.kl_target = kl_target;

    // The permutations the reference's generator constructors draw (one torch::randperm per epoch on the default CPU
    // generator, none after a KL early stop: ppo.cpp:63-104) come from the library: the generator's state goes in,
    // the epochs' Fisher-Yates walks run on background threads while the device works on the previous epoch, and the
    // state torch::randperm would have left behind comes back (bit-identical permutations: tests/test_host_rng_cpu.py).
    const int64_t rows = rd.T * rd.N;
    const bool recurrent = policy->is_recurrent();
    const int64_t perm_len = recurrent ? rd.N : rows;
    cpprl_b200_ppo_stats st{};
    auto gen = at::detail::getDefaultCPUGenerator();
    if (perm_len < static_cast<int64_t>(0xffffffffu / 20))
    {
        std::lock_guard<std::mutex> lock(gen.mutex());
        auto state = gen.get_state().contiguous();
        TORCH_CHECK(state.numel() == cpprl_b200_rng_state_bytes(), "unexpected CPU generator state size");
        check(cpprl_b200_ppo_update_rng(handle, &pd, &rd, &opt, state.data_ptr(), &st,
                                        detail::current_stream(rollouts.get_rewards())));
        gen.set_state(state);
    }
    else
    {
        // randperm's 64-bit branch (more than 2^32 / 20 rows): drawn by torch itself, epoch by epoch
        check(cpprl_b200_ppo_begin(handle, &pd, &rd, &opt, detail::current_stream(rollouts.get_rewards())));
        // a mean log-prob difference cannot reach 1e3: with such a target the stop can never trigger and the host
        // need not synchronise between epochs to learn whether the next epoch's generator would have been constructed
        const bool can_stop = static_cast<double>(kl_target) * 1.5 < 1e3;
        for (int epoch = 0; epoch < num_epoch; ++epoch)
        {
            if (can_stop && epoch > 0)
            {
                int stopped = 0;
                check(cpprl_b200_ppo_stopped(handle, &stopped));
                if (stopped)
                    break;
            }
            auto perm = torch::randperm(perm_len, torch::TensorOptions(torch::kLong));
            check(cpprl_b200_ppo_epoch(handle, perm.data_ptr<int64_t>(), perm_len));
        }
        check(cpprl_b200_ppo_end(handle, &st));
    }
    std::vector<UpdateDatum> out{{"Value loss", st.value_loss},
                                 {"Action loss", st.action_loss},
                                 {"Clip fraction", st.clip_fraction},
                                 {"Entropy", st.entropy},
                                 {"KL divergence", st.kl_divergence}};
    if (st.kl_early_stop_update > -1)
        out.push_back({"KL divergence early stop update", st.kl_early_stop_update});
    return out;
}

A2C::A2C(Policy &policy, float actor_loss_coef, float value_loss_coef, float entropy_coef, float learning_rate,
         float epsilon, float alpha, float max_grad_norm)
    : policy(policy), actor_loss_coef(actor_loss_coef), value_loss_coef(value_loss_coef), entropy_coef(entropy_coef),
      max_grad_norm(max_grad_norm), original_learning_rate(learning_rate), epsilon(epsilon), alpha(alpha)
{
}

A2C::~A2C() { cpprl_b200_a2c_destroy(handle); }

std::vector<UpdateDatum> A2C::update(RolloutStorage &rollouts, float decay_level)
{
    cpprl_b200_policy pd;
    policy->describe(pd);
    cpprl_b200_rollout rd;
    rollouts.describe(rd);
    if (handle == nullptr || handle_params != pd.n_params)
    {
        cpprl_b200_a2c_destroy(handle);
        handle = nullptr;
        check(cpprl_b200_a2c_create(&handle, pd.n_params));
        handle_params = pd.n_params;
    }
    cpprl_b200_a2c_options opt{};
    opt.value_loss_coef = value_loss_coef;
    opt.entropy_coef = entropy_coef;
    opt.learning_rate = original_learning_rate * decay_level; // a2c.cpp:39
    opt.epsilon = epsilon;
    opt.alpha = alpha;
    cpprl_b200_a2c_stats st{};
    check(cpprl_b200_a2c_update(handle, &pd, &rd, &opt, &st, detail::current_stream(rollouts.get_rewards())));
    return {{"Value loss", st.value_loss}, {"Action loss", st.action_loss}, {"Entropy", st.entropy}};
}
} // namespace cpprl
