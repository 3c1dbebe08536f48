// model.cpp — NNBase / MlpBase / CnnBase / output layers / Policy.  Reference behaviour: src/model/*.cpp.
// The torch modules exist so that names, shapes, initialisation, parameters(), to() and torch::save match the
// reference; `forward` (plain tensor expressions) serves the rollout side (act / get_probs).  The update path —
// evaluate_actions, get_values and everything in PPO/A2C::update — reads the SAME memory through the flat parameter
// buffer and runs in the fused CUDA kernels.
#include <cmath>

#include "host_common.h"

namespace cpprl
{
using detail::check;
using detail::current_stream;
using detail::fptr;

// ---- initialisation -------------------------------------------------------------------------------------------
static void orthogonal_fill(torch::Tensor tensor, double gain)
{
    torch::NoGradGuard guard;
    TORCH_CHECK(tensor.dim() >= 2, "Only tensors with 2 or more dimensions are supported");
    const int64_t rows = tensor.size(0), cols = tensor.numel() / rows;
    auto a = torch::randn({rows, cols});
    if (rows < cols)
        a = a.t();
    auto qr = torch::linalg_qr(a, "reduced");
    auto q = std::get<0>(qr) * torch::diag(std::get<1>(qr), 0).sign(); // uniform Haar measure (Mezzadri 2006)
    if (rows < cols)
        q = q.t();
    tensor.view({rows, cols}).copy_(q.to(tensor.device()));
    tensor.mul_(gain);
}

void init_weights(torch::OrderedDict<std::string, torch::Tensor> parameters, double weight_gain, double bias_gain)
{
    for (const auto &item : parameters)
    {
        if (item.value().size(0) == 0)
            continue;
        if (item.key().find("bias") != std::string::npos)
            torch::nn::init::constant_(item.value(), bias_gain);
        else if (item.key().find("weight") != std::string::npos)
            orthogonal_fill(item.value(), weight_gain);
    }
}

// ---- NNBase -------------------------------------------------------------------------------------------------------
NNBase::NNBase(bool recurrent, unsigned int recurrent_input_size, unsigned int hidden_size)
    : gru(nullptr), hidden_size(hidden_size), recurrent(recurrent)
{
    if (recurrent)
    {
        gru = register_module("gru", torch::nn::GRU(torch::nn::GRUOptions(recurrent_input_size, hidden_size)));
        init_weights(gru->named_parameters(), 1, 0);
    }
}

std::vector<torch::Tensor> NNBase::forward(torch::Tensor, torch::Tensor, torch::Tensor) { return {}; }

unsigned int NNBase::get_hidden_size() const { return recurrent ? hidden_size : 1; }

std::vector<torch::Tensor> NNBase::forward_gru(torch::Tensor x, torch::Tensor hxs, torch::Tensor masks)
{
    if (x.size(0) == hxs.size(0))
    {
        auto out = gru->forward(x.unsqueeze(0), (hxs * masks).unsqueeze(0));
        return {std::get<0>(out).squeeze(0), std::get<1>(out).squeeze(0)};
    }
    // sequence mode: [T*n, D] with n = hxs rows; runs are cut at every step t >= 1 where any agent's mask is 0 and
    // each run restarts from the ORIGINAL hxs times the run's first mask row — the state is never carried across
    // runs (the reference's behaviour, nn_base.cpp:98-104)
    const int64_t n = hxs.size(0), T = x.size(0) / n;
    auto xs = x.view({T, n, x.size(1)});
    auto ms = masks.view({T, n});
    auto ends = (ms.narrow(0, 1, T - 1) == 0).any(-1).to(torch::kCPU);
    std::vector<int64_t> cuts{0};
    for (int64_t t = 0; t < T - 1; ++t)
        if (ends[t].item().toBool())
            cuts.push_back(t + 1);
    cuts.push_back(T);
    std::vector<torch::Tensor> outs;
    for (size_t i = 0; i + 1 < cuts.size(); ++i)
    {
        auto h0 = hxs.unsqueeze(0) * ms[cuts[i]].view({1, -1, 1});
        outs.push_back(std::get<0>(gru->forward(xs.narrow(0, cuts[i], cuts[i + 1] - cuts[i]), h0)));
    }
    return {torch::cat(outs, 0).view({T * n, -1}), hxs};
}

// ---- MlpBase / CnnBase ------------------------------------------------------------------------------------------------
static torch::nn::Sequential tanh_tower(int64_t in, int64_t hidden)
{
    return torch::nn::Sequential(torch::nn::Linear(in, hidden), torch::nn::Functional(torch::tanh),
                                 torch::nn::Linear(hidden, hidden), torch::nn::Functional(torch::tanh));
}

MlpBase::MlpBase(unsigned int num_inputs, bool recurrent, unsigned int hidden_size)
    : NNBase(recurrent, num_inputs, hidden_size), actor(nullptr), critic(nullptr), critic_linear(nullptr),
      num_inputs(num_inputs)
{
    const int64_t tower_in = recurrent ? hidden_size : num_inputs; // the GRU feeds the towers when present
    actor = register_module("actor", tanh_tower(tower_in, hidden_size));
    critic = register_module("critic", tanh_tower(tower_in, hidden_size));
    critic_linear = register_module("critic_linear", torch::nn::Linear(hidden_size, 1));
    for (auto params : {actor->named_parameters(), critic->named_parameters(), critic_linear->named_parameters()})
        init_weights(params, std::sqrt(2.), 0);
    train();
}

std::vector<torch::Tensor> MlpBase::forward(torch::Tensor inputs, torch::Tensor hxs, torch::Tensor masks)
{
    auto x = inputs;
    if (is_recurrent())
    {
        auto g = forward_gru(x, hxs, masks);
        x = g[0];
        hxs = g[1];
    }
    return {critic_linear->forward(critic->forward(x)), actor->forward(x), hxs};
}

CnnBase::CnnBase(unsigned int num_inputs, bool recurrent, unsigned int hidden_size)
    : NNBase(recurrent, hidden_size, hidden_size), main(nullptr), critic_linear(nullptr), num_inputs(num_inputs)
{
    using namespace torch::nn;
    main = register_module(
        "main", Sequential(Conv2d(Conv2dOptions(num_inputs, 32, 8).stride(4)), Functional(torch::relu),
                           Conv2d(Conv2dOptions(32, 64, 4).stride(2)), Functional(torch::relu),
                           Conv2d(Conv2dOptions(64, 32, 3).stride(1)), Functional(torch::relu), Flatten(),
                           Linear(32 * 7 * 7, hidden_size), Functional(torch::relu)));
    critic_linear = register_module("critic_linear", Sequential(Linear(hidden_size, 1)));
    init_weights(main->named_parameters(), std::sqrt(2.), 0);
    init_weights(critic_linear->named_parameters(), 1, 0);
    train();
}

std::vector<torch::Tensor> CnnBase::forward(torch::Tensor inputs, torch::Tensor hxs, torch::Tensor masks)
{
    auto x = main->forward(inputs / 255.);
    if (is_recurrent())
    {
        auto g = forward_gru(x, hxs, masks);
        x = g[0];
        hxs = g[1];
    }
    return {critic_linear->forward(x), x, hxs};
}

// ---- output layers ----------------------------------------------------------------------------------------------------
CategoricalOutput::CategoricalOutput(unsigned int num_inputs, unsigned int num_outputs)
    : linear(register_module("linear", torch::nn::Linear(num_inputs, num_outputs)))
{
    init_weights(linear->named_parameters(), 0.01, 0);
}

std::unique_ptr<Distribution> CategoricalOutput::forward(torch::Tensor x)
{
    auto logits = linear(x);
    return std::make_unique<Categorical>(nullptr, &logits);
}

NormalOutput::NormalOutput(unsigned int num_inputs, unsigned int num_outputs) : linear_loc(nullptr)
{
    // scale_log is registered BEFORE linear_loc (output_layers.cpp:47-50): it precedes it in parameters()
    scale_log = register_parameter("scale_log", torch::zeros({static_cast<int64_t>(num_outputs)}));
    linear_loc = register_module("linear_loc", torch::nn::Linear(num_inputs, num_outputs));
    init_weights(linear_loc->named_parameters(), 1, 0);
}

std::unique_ptr<Distribution> NormalOutput::forward(torch::Tensor x)
{
    return std::make_unique<Normal>(linear_loc(x), scale_log.exp());
}

// ---- Policy -----------------------------------------------------------------------------------------------------------
PolicyImpl::PolicyImpl(ActionSpace action_space, std::shared_ptr<NNBase> base, bool normalize_observations)
    : action_space(action_space), base(register_module("base", base)), observation_normalizer(nullptr)
{
    const int num_outputs = action_space.shape.at(0);
    if (action_space.type == "Discrete")
        output_layer = std::make_shared<CategoricalOutput>(base->get_output_size(), num_outputs);
    else if (action_space.type == "Box")
        output_layer = std::make_shared<NormalOutput>(base->get_output_size(), num_outputs);
    else if (action_space.type == "MultiBinary")
        throw std::runtime_error("Action space MultiBinary (Bernoulli head) is not implemented in cpprl_b200");
    else
        throw std::runtime_error("Action space " + action_space.type + " not supported");
    register_module("output", output_layer);
    if (normalize_observations)
    {
        // only MlpBase observations can be normalised (policy.cpp:44-49)
        auto *mlp = dynamic_cast<MlpBase *>(base.get());
        if (mlp == nullptr)
            throw std::runtime_error("Observation normalisation is only supported for MlpBase");
        observation_normalizer =
            register_module("observation_normalizer", ObservationNormalizer(static_cast<int>(mlp->get_num_inputs())));
    }
}

// Every parameter must be the slice [offset, offset + numel) of one flat fp32 CUDA buffer, in parameters() order.
// When that does not hold (first call, after to(device), after torch::load) the values are packed into a fresh
// buffer and the parameter tensors are re-pointed at their slices; values are preserved.
torch::Tensor PolicyImpl::flat_parameters() const
{
    auto params = parameters();
    int64_t total = 0;
    for (const auto &p : params)
        total += p.numel();
    bool aliased = flat_.defined() && flat_.numel() == total && flat_.is_cuda();
    if (aliased)
    {
        int64_t off = 0;
        for (const auto &p : params)
        {
            aliased = aliased && p.is_cuda() && p.is_contiguous() && p.data_ptr<float>() == flat_.data_ptr<float>() + off;
            off += p.numel();
        }
    }
    if (aliased)
        return flat_;
    if (params.empty() || !params[0].is_cuda())
        throw std::runtime_error("cpprl_b200: the policy must live on a CUDA device (call policy->to(device)); there is no CPU fallback");
    torch::NoGradGuard guard;
    auto flat = torch::empty({total}, torch::TensorOptions(torch::kFloat).device(params[0].device()));
    int64_t off = 0;
    for (auto &p : params)
    {
        auto slice = flat.narrow(0, off, p.numel()).view(p.sizes());
        slice.copy_(p);
        p.set_data(slice);
        off += p.numel();
    }
    flat_ = flat;
    return flat_;
}

void PolicyImpl::describe(cpprl_b200_policy &out) const
{
    auto flat = flat_parameters();
    out = cpprl_b200_policy{};
    const bool cnn = base->kind() == NNBase::Kind::Cnn;
    out.base_kind = cnn ? CPPRL_B200_BASE_CNN : (base->is_recurrent() ? CPPRL_B200_BASE_GRU_MLP : CPPRL_B200_BASE_MLP);
    out.head_kind = action_space.type == "Discrete" ? CPPRL_B200_HEAD_CATEGORICAL : CPPRL_B200_HEAD_NORMAL;
    if (cnn)
    {
        auto *c = static_cast<CnnBase *>(base.get());
        out.obs_chw[0] = c->get_num_inputs(), out.obs_chw[1] = 84, out.obs_chw[2] = 84;
        out.obs_dim = out.obs_chw[0] * 84 * 84;
    }
    else if (auto *m = dynamic_cast<MlpBase *>(base.get()))
        out.obs_dim = m->get_num_inputs();
    out.hidden = base->get_output_size();
    out.num_actions = action_space.shape.at(0);
    out.d_params = flat.data_ptr<float>();
    out.n_params = flat.numel();
    if (!observation_normalizer.is_empty())
    {
        const auto &rms = observation_normalizer->running();
        detail::need_cuda(rms->mean_buffer(), "observation normaliser buffers");
        out.d_norm_clip = fptr(observation_normalizer->clip_buffer());
        out.d_norm_count = fptr(rms->count_buffer());
        out.d_norm_mean = fptr(rms->mean_buffer());
        out.d_norm_var = fptr(rms->variance_buffer());
    }
}

std::vector<torch::Tensor> PolicyImpl::act(torch::Tensor inputs, torch::Tensor rnn_hxs, torch::Tensor masks) const
{
    // policy.cpp:52-81 as ONE C-ABI call (fused forward + Philox sample + log-prob).  The Philox seed is drawn from
    // torch's CPU generator, so torch::manual_seed makes a run reproducible as it does for the reference.
    cpprl_b200_policy desc;
    describe(desc);
    const int64_t B = inputs.size(0);
    auto obs = detail::dense(inputs.reshape({B, -1}).to(torch::kFloat), "act: inputs");
    const bool discrete = action_space.type == "Discrete";
    auto opts = obs.options();
    auto value = torch::empty({B, 1}, opts), logp = torch::empty({B, 1}, opts);
    auto action = discrete ? torch::empty({B, 1}, opts.dtype(torch::kLong)) : torch::empty({B, action_space.shape.at(0)}, opts);
    torch::Tensor hxs, msk, hxs_out;
    if (base->is_recurrent())
    {
        hxs = detail::dense(rnn_hxs.to(torch::kFloat), "act: rnn_hxs");
        msk = detail::dense(masks.to(torch::kFloat).reshape({-1}), "act: masks");
        hxs_out = torch::empty_like(hxs);
    }
    const uint64_t seed = static_cast<uint64_t>(torch::randint(0, static_cast<int64_t>(1) << 62, {1}, torch::kLong).item<int64_t>());
    check(cpprl_b200_act(&desc, fptr(obs), hxs.defined() ? fptr(hxs) : nullptr, hxs.defined() ? hxs.size(0) : 0,
                         msk.defined() ? fptr(msk) : nullptr, B, seed, 0, fptr(value), action.data_ptr(), fptr(logp),
                         hxs_out.defined() ? fptr(hxs_out) : nullptr, current_stream(obs)));
    return {value, action, logp, hxs_out.defined() ? hxs_out : rnn_hxs};
}

std::vector<torch::Tensor> PolicyImpl::evaluate_actions(torch::Tensor inputs, torch::Tensor rnn_hxs, torch::Tensor masks,
                                                        torch::Tensor actions) const
{
    cpprl_b200_policy desc;
    describe(desc);
    const int64_t B = inputs.size(0);
    auto obs = detail::dense(inputs.reshape({B, -1}).to(torch::kFloat), "evaluate_actions: inputs");
    const bool discrete = action_space.type == "Discrete";
    auto act = detail::dense(discrete ? actions.to(torch::kLong).reshape({B}) : actions.to(torch::kFloat).reshape({B, -1}),
                             "evaluate_actions: actions");
    auto opts = obs.options();
    auto value = torch::empty({B, 1}, opts), logp = torch::empty({B, 1}, opts), entropy = torch::empty({1}, opts);
    torch::Tensor hxs, msk, hxs_out;
    if (base->is_recurrent())
    {
        hxs = detail::dense(rnn_hxs.to(torch::kFloat), "evaluate_actions: rnn_hxs");
        msk = detail::dense(masks.to(torch::kFloat).reshape({-1}), "evaluate_actions: masks");
        hxs_out = torch::empty_like(hxs); // new state in step mode, the incoming one in sequence mode (nn_base.cpp:55,112)
    }
    check(cpprl_b200_evaluate_actions(&desc, fptr(obs), hxs.defined() ? fptr(hxs) : nullptr, hxs.defined() ? hxs.size(0) : 0,
                                      msk.defined() ? fptr(msk) : nullptr, act.data_ptr(), B, fptr(value), fptr(logp),
                                      fptr(entropy), hxs_out.defined() ? fptr(hxs_out) : nullptr, current_stream(obs)));
    return {value, logp, entropy.squeeze(0), hxs_out.defined() ? hxs_out : rnn_hxs};
}

torch::Tensor PolicyImpl::get_probs(torch::Tensor inputs, torch::Tensor rnn_hxs, torch::Tensor masks) const
{
    if (!observation_normalizer.is_empty())
        inputs = observation_normalizer->process_observation(inputs);
    auto out = base->forward(inputs, rnn_hxs, masks);
    auto dist = output_layer->forward(out[1]);
    auto *categorical = dynamic_cast<Categorical *>(dist.get());
    if (categorical == nullptr)
        throw std::runtime_error("get_probs is only defined for Discrete action spaces");
    return categorical->get_probs();
}

torch::Tensor PolicyImpl::get_values(torch::Tensor inputs, torch::Tensor rnn_hxs, torch::Tensor masks) const
{
    // bootstrap value for compute_returns (policy.cpp:132-144): the value head of the fused forward
    cpprl_b200_policy desc;
    describe(desc);
    const int64_t B = inputs.size(0);
    auto obs = detail::dense(inputs.reshape({B, -1}).to(torch::kFloat), "get_values: inputs");
    auto value = torch::empty({B, 1}, obs.options());
    torch::Tensor hxs, msk;
    if (base->is_recurrent())
    {
        hxs = detail::dense(rnn_hxs.to(torch::kFloat), "get_values: rnn_hxs");
        msk = detail::dense(masks.to(torch::kFloat).reshape({-1}), "get_values: masks");
    }
    check(cpprl_b200_get_values(&desc, fptr(obs), hxs.defined() ? fptr(hxs) : nullptr, hxs.defined() ? hxs.size(0) : 0,
                                msk.defined() ? fptr(msk) : nullptr, B, fptr(value), current_stream(obs)));
    return value;
}

void PolicyImpl::update_observation_normalizer(torch::Tensor observations)
{
    TORCH_CHECK(!observation_normalizer.is_empty(), "the policy was built without an observation normaliser");
    observation_normalizer->update(observations);
}
} // namespace cpprl
