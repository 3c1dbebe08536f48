// storage.cpp — RolloutStorage and the two generators.  Reference behaviour: src/storage.cpp,
// src/generators/{feed_forward_generator,recurrent_generator}.cpp.  Buffer bookkeeping (allocation, insert, views)
// stays tensor-handle code; compute_returns and the minibatch gathers go through the C ABI.
#include "host_common.h"

namespace cpprl
{
using detail::check;
using detail::current_stream;
using detail::fptr;

// ---- generators -----------------------------------------------------------------------------------------
FeedForwardGenerator::FeedForwardGenerator(int mini_batch_size, torch::Tensor observations, torch::Tensor hidden_states,
                                           torch::Tensor actions, torch::Tensor value_predictions, torch::Tensor returns,
                                           torch::Tensor masks, torch::Tensor action_log_probs, torch::Tensor advantages)
{
    const int64_t T = actions.size(0), N = actions.size(1);
    // flat rows t*N + env of the first T time rows of every field (feed_forward_generator.cpp:51-76)
    auto rows = [&](const torch::Tensor &t) { return t.narrow(0, 0, T).reshape({T * N, -1}).contiguous(); };
    fields_[0] = observations.narrow(0, 0, T).reshape({T * N, -1}).contiguous();
    fields_[1] = rows(hidden_states);
    fields_[2] = rows(actions);
    fields_[3] = rows(value_predictions);
    fields_[4] = rows(returns);
    fields_[5] = rows(masks);
    fields_[6] = rows(action_log_probs);
    fields_[7] = rows(advantages);
    obs_shape_ = observations.sizes().vec();
    indices_host_ = torch::randperm(advantages.numel(), torch::TensorOptions(torch::kLong)).view({-1, mini_batch_size});
    indices_dev_ = indices_host_.to(actions.device());
}

bool FeedForwardGenerator::done() const { return index_ >= indices_host_.size(0); }

MiniBatch FeedForwardGenerator::next()
{
    if (index_ >= indices_host_.size(0))
        throw std::runtime_error("No minibatches left in generator.");
    auto idx = indices_dev_[index_].contiguous();
    const int64_t B = idx.size(0);
    torch::Tensor out[8];
    for (int f = 0; f < 8; ++f)
    {
        const auto &src = detail::need_cuda(fields_[f], "FeedForwardGenerator: rollout field");
        out[f] = torch::empty({B, src.size(1)}, src.options());
        check(cpprl_b200_gather_rows(src.data_ptr(), src.size(0), src.size(1) * src.element_size(), idx.data_ptr<int64_t>(),
                                     B, out[f].data_ptr(), current_stream(src)));
    }
    std::vector<int64_t> obs_shape(obs_shape_.begin() + 2, obs_shape_.end());
    obs_shape.insert(obs_shape.begin(), B);
    ++index_;
    return MiniBatch(out[0].view(obs_shape), out[1], out[2], out[3], out[4], out[5], out[6], out[7]);
}

RecurrentGenerator::RecurrentGenerator(int num_processes, int num_mini_batch, torch::Tensor observations,
                                       torch::Tensor hidden_states, torch::Tensor actions, torch::Tensor value_predictions,
                                       torch::Tensor returns, torch::Tensor masks, torch::Tensor action_log_probs,
                                       torch::Tensor advantages)
    : observations_(observations), hidden_states_(hidden_states), actions_(actions), value_predictions_(value_predictions),
      returns_(returns), masks_(masks), action_log_probs_(action_log_probs), advantages_(advantages),
      indices_(torch::randperm(num_processes, torch::TensorOptions(torch::kLong))),
      num_envs_per_batch_(num_processes / num_mini_batch)
{
}

bool RecurrentGenerator::done() const { return index_ >= indices_.size(0); }

MiniBatch RecurrentGenerator::next()
{
    if (index_ >= indices_.size(0))
        throw std::runtime_error("No minibatches left in generator.");
    const int64_t env = indices_[index_].item().toLong(), npb = num_envs_per_batch_;
    const int64_t T = actions_.size(0);
    // window of npb environments, all T steps, flattened to [T*npb, ...] (throws like the reference unless npb == 1)
    auto window = [&](const torch::Tensor &t) {
        auto w = t.narrow(0, 0, T).narrow(1, env, npb);
        std::vector<int64_t> shape = w.sizes().vec();
        shape.erase(shape.begin());
        shape[0] = T * npb;
        return w.view(shape);
    };
    MiniBatch mb;
    mb.observations = window(observations_);
    mb.hidden_states = hidden_states_[0].narrow(0, env, npb).view({npb, -1});
    mb.actions = window(actions_);
    mb.value_predictions = window(value_predictions_);
    mb.returns = window(returns_);
    mb.masks = window(masks_);
    mb.action_log_probs = window(action_log_probs_);
    mb.advantages = window(advantages_);
    ++index_;
    return mb;
}

// ---- RolloutStorage ---------------------------------------------------------------------------------------
RolloutStorage::RolloutStorage(int64_t num_steps, int64_t num_processes, c10::ArrayRef<int64_t> obs_shape,
                               ActionSpace action_space, int64_t hidden_state_size, torch::Device device)
    : device(device), num_steps(num_steps), step(0)
{
    auto f32 = torch::TensorOptions(torch::kFloat).device(device);
    std::vector<int64_t> obs_dims{num_steps + 1, num_processes};
    obs_dims.insert(obs_dims.end(), obs_shape.begin(), obs_shape.end());
    observations = torch::zeros(obs_dims, f32);
    hidden_states = torch::zeros({num_steps + 1, num_processes, hidden_state_size}, f32);
    rewards = torch::zeros({num_steps, num_processes, 1}, f32);
    value_predictions = torch::zeros({num_steps + 1, num_processes, 1}, f32);
    returns = torch::zeros({num_steps + 1, num_processes, 1}, f32);
    action_log_probs = torch::zeros({num_steps, num_processes, 1}, f32);
    if (action_space.type == "Discrete")
        actions = torch::zeros({num_steps, num_processes, 1}, torch::TensorOptions(torch::kLong).device(device));
    else
        actions = torch::zeros({num_steps, num_processes, action_space.shape[0]}, f32);
    masks = torch::ones({num_steps + 1, num_processes, 1}, f32);
}

RolloutStorage::RolloutStorage(std::vector<RolloutStorage *> parts, torch::Device device)
    : device(device), num_steps(parts.at(0)->get_rewards().size(0)), step(0)
{
    // concatenate the workers' storages along the environment axis (storage.cpp:51-104)
    if (!device.is_cuda())
    {
        // host-resident bookkeeping copy (nothing on the update path runs on it until ->to(cuda))
        auto join = [&](const torch::Tensor &(RolloutStorage::*get)() const) {
            std::vector<torch::Tensor> pieces;
            for (auto *p : parts)
                pieces.push_back((p->*get)());
            return torch::cat(pieces, 1).to(device);
        };
        observations = join(&RolloutStorage::get_observations);
        hidden_states = join(&RolloutStorage::get_hidden_states);
        rewards = join(&RolloutStorage::get_rewards);
        value_predictions = join(&RolloutStorage::get_value_predictions);
        returns = join(&RolloutStorage::get_returns);
        action_log_probs = join(&RolloutStorage::get_action_log_probs);
        actions = join(&RolloutStorage::get_actions);
        masks = join(&RolloutStorage::get_masks);
        return;
    }
    // device path: allocate the joined buffers, then ONE launch per part scatters its eight fields into its
    // environment columns (cpprl_b200_storage_concat)
    int64_t total = 0;
    for (auto *p : parts)
        total += p->get_rewards().size(1);
    auto like = [&](const torch::Tensor &t) {
        auto shape = t.sizes().vec();
        shape[1] = total;
        return torch::empty(shape, t.options().device(device));
    };
    auto *first = parts.at(0);
    observations = like(first->observations);
    hidden_states = like(first->hidden_states);
    rewards = like(first->rewards);
    value_predictions = like(first->value_predictions);
    returns = like(first->returns);
    action_log_probs = like(first->action_log_probs);
    actions = like(first->actions);
    masks = like(first->masks);
    cpprl_b200_storage dst;
    describe_storage(dst);
    int64_t offset = 0;
    for (auto *p : parts)
    {
        // a part living elsewhere (host, another GPU) is brought over first; the staging copy is the H2D transfer itself
        RolloutStorage staged = *p;
        staged.to(device);
        cpprl_b200_storage src;
        staged.describe_storage(src);
        check(cpprl_b200_storage_concat(&dst, &src, offset, current_stream(observations)));
        offset += src.N;
    }
}

void RolloutStorage::describe_storage(cpprl_b200_storage &out)
{
    for (auto *t : {&observations, &hidden_states, &rewards, &value_predictions, &returns, &action_log_probs, &actions,
                    &masks})
    {
        detail::need_cuda(*t, "RolloutStorage: rollout buffer");
        if (!t->is_contiguous()) // a rebound, non-contiguous buffer is made dense and stays the storage's
            *t = t->contiguous();
    }
    out.T = rewards.size(0);
    out.N = rewards.size(1);
    out.obs_dim = observations.numel() / (observations.size(0) * observations.size(1));
    out.hidden_dim = hidden_states.size(2);
    out.action_cols = actions.size(2);
    out.action_elem_bytes = static_cast<int32_t>(actions.element_size());
    out.d_observations = fptr(observations);
    out.d_hidden_states = fptr(hidden_states);
    out.d_rewards = fptr(rewards);
    out.d_value_preds = fptr(value_predictions);
    out.d_returns = fptr(returns);
    out.d_action_log_probs = fptr(action_log_probs);
    out.d_actions = actions.data_ptr();
    out.d_masks = fptr(masks);
}

// what Tensor::copy_ accepts (any device, convertible dtype, broadcastable shape) as a dense device row of the
// destination's dtype
static torch::Tensor as_row(const torch::Tensor &value, const torch::Tensor &row)
{
    return value.to(row.options()).expand_as(row).contiguous();
}

void RolloutStorage::after_update()
{
    if (!observations.is_cuda())
    { // host-resident storage: handle bookkeeping only
        observations[0].copy_(observations[-1]);
        hidden_states[0].copy_(hidden_states[-1]);
        masks[0].copy_(masks[-1]);
        return;
    }
    cpprl_b200_storage st;
    describe_storage(st);
    check(cpprl_b200_storage_after_update(&st, current_stream(observations)));
}

void RolloutStorage::compute_returns(torch::Tensor next_value, bool use_gae, float gamma, float tau)
{
    detail::need_cuda(rewards, "RolloutStorage::compute_returns: storage");
    // the kernel writes in place; a rebound, non-contiguous buffer is made dense first (and stays the storage's)
    value_predictions = value_predictions.contiguous();
    returns = returns.contiguous();
    auto r = detail::dense(rewards, "rewards"), m = detail::dense(masks, "masks");
    auto nv = detail::dense(next_value.to(torch::kFloat), "compute_returns: next_value");
    check(cpprl_b200_compute_returns(fptr(r), fptr(value_predictions), fptr(m), fptr(nv), fptr(returns), rewards.size(0),
                                     rewards.size(1), use_gae ? 1 : 0, gamma, tau, current_stream(r)));
}

void RolloutStorage::bootstrap_returns(Policy &policy, bool use_gae, float gamma, float tau)
{
    // get_values(observations[-1], hidden_states[-1], masks[-1]) -> compute_returns (gym_client.cpp:266-275) in one call
    cpprl_b200_policy pd;
    policy->describe(pd);
    cpprl_b200_storage st;
    describe_storage(st);
    check(cpprl_b200_bootstrap_returns(&pd, &st, use_gae ? 1 : 0, gamma, tau, current_stream(rewards)));
}

std::unique_ptr<Generator> RolloutStorage::feed_forward_generator(torch::Tensor advantages, int num_mini_batch)
{
    const auto num_processes = actions.size(1);
    const auto batch_size = num_processes * num_steps;
    if (batch_size < num_mini_batch)
        throw std::runtime_error("PPO needs the number of processes (" + std::to_string(num_processes) +
                                 ") * the number of steps (" + std::to_string(num_steps) + ") = " +
                                 std::to_string(num_processes * num_steps) +
                                 " to be greater than or equal to the number of minibatches (" +
                                 std::to_string(num_mini_batch) + ")");
    const auto mini_batch_size = batch_size / num_mini_batch;
    return std::make_unique<FeedForwardGenerator>(mini_batch_size, observations, hidden_states, actions, value_predictions,
                                                  returns, masks, action_log_probs, advantages);
}

void RolloutStorage::insert(torch::Tensor observation, torch::Tensor hidden_state, torch::Tensor action,
                            torch::Tensor action_log_prob, torch::Tensor value_prediction, torch::Tensor reward,
                            torch::Tensor mask)
{
    if (!observations.is_cuda())
    { // host-resident storage: handle bookkeeping only
        observations[step + 1].copy_(observation);
        hidden_states[step + 1].copy_(hidden_state);
        actions[step].copy_(action);
        action_log_probs[step].copy_(action_log_prob);
        value_predictions[step].copy_(value_prediction);
        rewards[step].copy_(reward);
        masks[step + 1].copy_(mask);
    }
    else
    {
        // all seven fields in ONE launch (cpprl_b200_storage_insert) instead of seven copy kernels
        cpprl_b200_storage st;
        describe_storage(st);
        auto o = as_row(observation, observations[step + 1]), h = as_row(hidden_state, hidden_states[step + 1]);
        auto a = as_row(action, actions[step]), lp = as_row(action_log_prob, action_log_probs[step]);
        auto v = as_row(value_prediction, value_predictions[step]), r = as_row(reward, rewards[step]);
        auto m = as_row(mask, masks[step + 1]);
        check(cpprl_b200_storage_insert(&st, step, fptr(o), fptr(h), a.data_ptr(), fptr(lp), fptr(v), fptr(r), fptr(m),
                                        current_stream(observations)));
    }
    step = (step + 1) % num_steps;
}

std::unique_ptr<Generator> RolloutStorage::recurrent_generator(torch::Tensor advantages, int num_mini_batch)
{
    const auto num_processes = actions.size(1);
    if (num_processes < num_mini_batch)
        throw std::runtime_error("PPO needs the number of processes (" + std::to_string(num_processes) +
                                 ") to be greater than or equal to the number of minibatches (" +
                                 std::to_string(num_mini_batch) + ")");
    return std::make_unique<RecurrentGenerator>(num_processes, num_mini_batch, observations, hidden_states, actions,
                                                value_predictions, returns, masks, action_log_probs, advantages);
}

void RolloutStorage::set_first_observation(torch::Tensor observation)
{
    if (!observations.is_cuda())
    {
        observations[0].copy_(observation);
        return;
    }
    cpprl_b200_storage st;
    describe_storage(st);
    auto o = as_row(observation, observations[0]);
    check(cpprl_b200_storage_set_first_observation(&st, fptr(o), current_stream(observations)));
}

void RolloutStorage::to(torch::Device dev)
{
    device = dev;
    for (auto *t : {&observations, &hidden_states, &rewards, &value_predictions, &returns, &action_log_probs, &actions,
                    &masks})
        *t = t->to(dev);
}

void RolloutStorage::describe(cpprl_b200_rollout &out) const
{
    auto need = [](const torch::Tensor &t, const char *what) -> const torch::Tensor & {
        detail::need_cuda(t, what);
        if (!t.is_contiguous())
            throw std::runtime_error(std::string("cpprl_b200: rollout buffer '") + what + "' must be contiguous");
        return t;
    };
    out.T = rewards.size(0);
    out.N = rewards.size(1);
    out.hidden_dim = hidden_states.size(2);
    out.action_cols = actions.size(2);
    out.d_observations = fptr(need(observations, "observations"));
    out.d_hidden_states = fptr(need(hidden_states, "hidden_states"));
    out.d_actions = need(actions, "actions").data_ptr();
    out.d_value_preds = fptr(need(value_predictions, "value_predictions"));
    out.d_returns = fptr(need(returns, "returns"));
    out.d_masks = fptr(need(masks, "masks"));
    out.d_action_log_probs = fptr(need(action_log_probs, "action_log_probs"));
}
} // namespace cpprl
