// cpprl/b200/host_api.h — the C++ drop-in: the reference's `cpprl::` classes over the cpprl_b200 C ABI.
//
// One header declares the whole surface; the reference-named headers (cpprl/storage.h, cpprl/algorithms/ppo.h, ...)
// are one-line forwarders to it, so `#include "cpprl/cpprl.h"` (or any individual header) and code written against
// Omegastick/pytorch-cpp-rl keep compiling.  Signatures, default arguments, exception texts and the
// `named_parameters()` / `named_buffers()` names follow the reference (file:line cited per class, paths relative to
// the reference repository); the bodies (pytorch_cpp_rl_b200/host/*.cpp) do no tensor math on the update path — they
// pull raw device pointers out of the torch::Tensor arguments and call include/cpprl_b200.h.
//
// Ownership rules that differ from the reference and why:
//  * PolicyImpl keeps every parameter of its sub-modules as a VIEW into one flat fp32 buffer (registration order), so
//    the fused update kernel can step them in place; `parameters()`, `torch::save`, `to()` keep working — after a
//    `to(device)` the next call re-packs the buffer (PolicyImpl::flat_parameters).
//  * PPO / A2C own an opaque C-ABI handle holding the optimiser state (Adam exp_avg / exp_avg_sq / step, RMSprop
//    square_avg) instead of a torch::optim object.
//  * Tensors handed to the update path must live on a CUDA device (sm_100); there is no CPU fallback: CPU tensors
//    make the call throw std::runtime_error.
#pragma once

#include <torch/torch.h>

#include <memory>
#include <string>
#include <vector>

struct cpprl_b200_ppo;
struct cpprl_b200_a2c;
struct cpprl_b200_policy;
struct cpprl_b200_rollout;
struct cpprl_b200_storage;

namespace cpprl
{
// ---------------------------------------------------------------------------------------------------------
// spaces.h:9-13, algorithms/algorithm.h:10-24
// ---------------------------------------------------------------------------------------------------------
struct ActionSpace
{
    std::string type; // "Discrete" | "Box" | "MultiBinary"
    std::vector<int64_t> shape;
};

struct UpdateDatum
{
    std::string name;
    float value;
};

class RolloutStorage;
class Policy;

class Algorithm
{
  public:
    virtual ~Algorithm() = default;
    virtual std::vector<UpdateDatum> update(RolloutStorage &rollouts, float decay_level = 1) = 0;
};

// ---------------------------------------------------------------------------------------------------------
// distributions (distribution.h, categorical.h, normal.h) — API-visible objects returned by the output layers;
// rollout side (act / get_probs), plain tensor expressions.
// ---------------------------------------------------------------------------------------------------------
class Distribution
{
  public:
    virtual ~Distribution() = default;
    virtual torch::Tensor entropy() = 0;
    virtual torch::Tensor log_prob(torch::Tensor value) = 0;
    virtual torch::Tensor sample(c10::ArrayRef<int64_t> sample_shape = {}) = 0;

  protected:
    std::vector<int64_t> batch_shape, event_shape;
    std::vector<int64_t> extended_shape(c10::ArrayRef<int64_t> sample_shape);
};

class Categorical : public Distribution
{
  public:
    // exactly one of probs / logits (categorical.cpp:9-45); probs are normalised and clamped to [1.21e-7, 1-1.21e-7]
    Categorical(const torch::Tensor *probs, const torch::Tensor *logits);
    torch::Tensor entropy() override;
    torch::Tensor log_prob(torch::Tensor value) override;
    torch::Tensor sample(c10::ArrayRef<int64_t> sample_shape = {}) override;
    torch::Tensor get_logits() { return logits_; }
    torch::Tensor get_probs() { return probs_; }

  private:
    torch::Tensor probs_, logits_;
    int64_t num_events_ = 0;
};

class Normal : public Distribution
{
  public:
    Normal(const torch::Tensor loc, const torch::Tensor scale); // normal.cpp:14-22
    torch::Tensor entropy() override;
    torch::Tensor log_prob(torch::Tensor value) override;
    torch::Tensor sample(c10::ArrayRef<int64_t> sample_shape = {}) override;
    torch::Tensor get_loc() { return loc_; }
    torch::Tensor get_scale() { return scale_; }

  private:
    torch::Tensor loc_, scale_;
};

// ---------------------------------------------------------------------------------------------------------
// running_mean_std.h:20-28, observation_normalizer.h:20-32
// ---------------------------------------------------------------------------------------------------------
class RunningMeanStdImpl : public torch::nn::Module
{
  public:
    explicit RunningMeanStdImpl(int size);
    RunningMeanStdImpl(std::vector<float> means, std::vector<float> variances);
    // batch moments + Chan merge in ONE kernel (cpprl_b200_rms_update); observation must be a CUDA tensor
    void update(torch::Tensor observation);
    int get_count() const { return static_cast<int>(count.item().toFloat()); }
    torch::Tensor get_mean() const { return mean.clone(); }
    torch::Tensor get_variance() const { return variance.clone(); }
    void set_count(int c) { count[0] = c + 1e-8; }
    // raw buffers (no clone) for the C-ABI descriptors
    const torch::Tensor &count_buffer() const { return count; }
    const torch::Tensor &mean_buffer() const { return mean; }
    const torch::Tensor &variance_buffer() const { return variance; }

  private:
    torch::Tensor count, mean, variance; // registered buffers "count", "mean", "variance"
    torch::Tensor workspace_;
};
TORCH_MODULE(RunningMeanStd);

class ObservationNormalizer;
class ObservationNormalizerImpl : public torch::nn::Module
{
  public:
    explicit ObservationNormalizerImpl(int size, float clip = 10.);
    ObservationNormalizerImpl(const std::vector<float> &means, const std::vector<float> &variances, float clip = 10.);
    // unweighted mean of the others' means / variances, counts summed (observation_normalizer.cpp:19-67), host code
    explicit ObservationNormalizerImpl(const std::vector<ObservationNormalizer> &others);
    torch::Tensor process_observation(torch::Tensor observation) const; // cpprl_b200_normalize_observations
    std::vector<float> get_mean() const;
    std::vector<float> get_variance() const;
    void update(torch::Tensor observations);
    float get_clip_value() const { return clip.item().toFloat(); }
    int get_step_count() const { return rms->get_count(); }
    const torch::Tensor &clip_buffer() const { return clip; }
    const RunningMeanStd &running() const { return rms; }

  private:
    torch::Tensor clip; // registered buffer "clip"
    RunningMeanStd rms; // registered module "rms"
};
TORCH_MODULE(ObservationNormalizer);

// ---------------------------------------------------------------------------------------------------------
// generators (generator.h:9-40, feed_forward_generator.h, recurrent_generator.h)
// ---------------------------------------------------------------------------------------------------------
struct MiniBatch
{
    torch::Tensor observations, hidden_states, actions, value_predictions, returns, masks, action_log_probs, advantages;
    MiniBatch() = default;
    MiniBatch(torch::Tensor observations, torch::Tensor hidden_states, torch::Tensor actions,
              torch::Tensor value_predictions, torch::Tensor returns, torch::Tensor masks,
              torch::Tensor action_log_probs, torch::Tensor advantages)
        : observations(observations), hidden_states(hidden_states), actions(actions),
          value_predictions(value_predictions), returns(returns), masks(masks), action_log_probs(action_log_probs),
          advantages(advantages)
    {
    }
};

class Generator
{
  public:
    virtual ~Generator() = default;
    virtual bool done() const = 0;
    virtual MiniBatch next() = 0;
};

// One torch::randperm(T*N) on the CPU generator per construction (= per epoch), reshaped [n_minibatches, mbs];
// next() gathers the eight fields with cpprl_b200_gather_rows (bit-exact copies of rows t*N + env).
class FeedForwardGenerator : public Generator
{
  public:
    FeedForwardGenerator(int mini_batch_size, torch::Tensor observations, torch::Tensor hidden_states,
                         torch::Tensor actions, torch::Tensor value_predictions, torch::Tensor returns,
                         torch::Tensor masks, torch::Tensor action_log_probs, torch::Tensor advantages);
    bool done() const override;
    MiniBatch next() override;
    const torch::Tensor &get_indices() const { return indices_host_; } // [n_minibatches, mbs] int64, CPU

  private:
    torch::Tensor fields_[8];     // time-major sources, last time row dropped where the field has T+1 rows
    std::vector<int64_t> obs_shape_;
    torch::Tensor indices_host_;  // as drawn
    torch::Tensor indices_dev_;   // same, on the rollout's device
    int64_t index_ = 0;
};

// One torch::randperm(N) per construction; minibatch k is environment perm[k], all T steps, as strided views
// (works only for num_mini_batch == N, like the reference: SURVEY.md §0 quirk 3).
class RecurrentGenerator : public Generator
{
  public:
    RecurrentGenerator(int num_processes, int num_mini_batch, torch::Tensor observations, torch::Tensor hidden_states,
                       torch::Tensor actions, torch::Tensor value_predictions, torch::Tensor returns,
                       torch::Tensor masks, torch::Tensor action_log_probs, torch::Tensor advantages);
    bool done() const override;
    MiniBatch next() override;

  private:
    torch::Tensor observations_, hidden_states_, actions_, value_predictions_, returns_, masks_, action_log_probs_,
        advantages_, indices_;
    int64_t index_ = 0, num_envs_per_batch_ = 1;
};

// ---------------------------------------------------------------------------------------------------------
// storage.h:13-73
// ---------------------------------------------------------------------------------------------------------
class RolloutStorage
{
  public:
    RolloutStorage(int64_t num_steps, int64_t num_processes, c10::ArrayRef<int64_t> obs_shape, ActionSpace action_space,
                   int64_t hidden_state_size, torch::Device device);
    RolloutStorage(std::vector<RolloutStorage *> individual_storages, torch::Device device);

    void after_update();
    // GAE / n-step scan on the device (cpprl_b200_compute_returns); side effects as storage.cpp:113-144
    void compute_returns(torch::Tensor next_value, bool use_gae, float gamma, float tau);
    // extension: compute_returns(policy->get_values(observations[-1], hidden_states[-1], masks[-1]), ...) as ONE
    // C-ABI call without the next_value temporary (cpprl_b200_bootstrap_returns)
    void bootstrap_returns(Policy &policy, bool use_gae, float gamma, float tau);
    std::unique_ptr<Generator> feed_forward_generator(torch::Tensor advantages, int num_mini_batch);
    void insert(torch::Tensor observation, torch::Tensor hidden_state, torch::Tensor action,
                torch::Tensor action_log_prob, torch::Tensor value_prediction, torch::Tensor reward, torch::Tensor mask);
    std::unique_ptr<Generator> recurrent_generator(torch::Tensor advantages, int num_mini_batch);
    void set_first_observation(torch::Tensor observation);
    void to(torch::Device device);

    const torch::Tensor &get_actions() const { return actions; }
    const torch::Tensor &get_action_log_probs() const { return action_log_probs; }
    const torch::Tensor &get_hidden_states() const { return hidden_states; }
    const torch::Tensor &get_masks() const { return masks; }
    const torch::Tensor &get_observations() const { return observations; }
    const torch::Tensor &get_returns() const { return returns; }
    const torch::Tensor &get_rewards() const { return rewards; }
    const torch::Tensor &get_value_predictions() const { return value_predictions; }
    void set_actions(torch::Tensor t) { actions = t; }
    void set_action_log_probs(torch::Tensor t) { action_log_probs = t; }
    void set_hidden_states(torch::Tensor t) { hidden_states = t; }
    void set_masks(torch::Tensor t) { masks = t; }
    void set_observations(torch::Tensor t) { observations = t; }
    void set_returns(torch::Tensor t) { returns = t; }
    void set_rewards(torch::Tensor t) { rewards = t; }
    void set_value_predictions(torch::Tensor t) { value_predictions = t; }

    // C-ABI view of the buffers (re-read on every call: the handles above may have been rebound)
    void describe(cpprl_b200_rollout &out) const;
    // writable view for the write-side entry points (insert / after_update / concatenation); makes rebound,
    // non-contiguous buffers dense first
    void describe_storage(cpprl_b200_storage &out);

  private:
    torch::Tensor observations, hidden_states, rewards, value_predictions, returns, action_log_probs, actions, masks;
    torch::Device device;
    int64_t num_steps;
    int64_t step;
};

// ---------------------------------------------------------------------------------------------------------
// model (model_utils.h, nn_base.h, mlp_base.h, cnn_base.h, output_layers.h, policy.h)
// ---------------------------------------------------------------------------------------------------------
struct FlattenImpl : torch::nn::Module
{
    torch::Tensor forward(torch::Tensor x) { return x.view({x.size(0), -1}); }
};
TORCH_MODULE(Flatten);

// orthogonal weights (QR of a normal matrix, sign-fixed) scaled by weight_gain, constant biases (model_utils.cpp:40-69)
void init_weights(torch::OrderedDict<std::string, torch::Tensor> parameters, double weight_gain, double bias_gain);

class NNBase : public torch::nn::Module
{
  public:
    NNBase(bool recurrent, unsigned int recurrent_input_size, unsigned int hidden_size);
    virtual std::vector<torch::Tensor> forward(torch::Tensor inputs, torch::Tensor hxs, torch::Tensor masks);
    // step mode (x rows == hxs rows) or sequence mode with the reference's run splitting (nn_base.cpp:47-114)
    std::vector<torch::Tensor> forward_gru(torch::Tensor x, torch::Tensor hxs, torch::Tensor masks);
    unsigned int get_hidden_size() const; // 1 when not recurrent (nn_base.cpp:38-45)
    int get_output_size() const { return hidden_size; }
    bool is_recurrent() const { return recurrent; }
    enum class Kind { Plain, Mlp, Cnn };
    virtual Kind kind() const { return Kind::Plain; }

  private:
    torch::nn::GRU gru;
    unsigned int hidden_size;
    bool recurrent;
};

class MlpBase : public NNBase
{
  public:
    MlpBase(unsigned int num_inputs, bool recurrent = false, unsigned int hidden_size = 64);
    std::vector<torch::Tensor> forward(torch::Tensor inputs, torch::Tensor hxs, torch::Tensor masks) override;
    unsigned int get_num_inputs() const { return num_inputs; }
    Kind kind() const override { return Kind::Mlp; }

  private:
    torch::nn::Sequential actor, critic;
    torch::nn::Linear critic_linear;
    unsigned int num_inputs;
};

class CnnBase : public NNBase
{
  public:
    CnnBase(unsigned int num_inputs, bool recurrent = false, unsigned int hidden_size = 512);
    std::vector<torch::Tensor> forward(torch::Tensor inputs, torch::Tensor hxs, torch::Tensor masks) override;
    Kind kind() const override { return Kind::Cnn; }
    unsigned int get_num_inputs() const { return num_inputs; }

  private:
    torch::nn::Sequential main, critic_linear;
    unsigned int num_inputs;
};

class OutputLayer : public torch::nn::Module
{
  public:
    virtual ~OutputLayer() = default;
    virtual std::unique_ptr<Distribution> forward(torch::Tensor x) = 0;
};

class CategoricalOutput : public OutputLayer
{
  public:
    CategoricalOutput(unsigned int num_inputs, unsigned int num_outputs);
    std::unique_ptr<Distribution> forward(torch::Tensor x) override;

  private:
    torch::nn::Linear linear;
};

class NormalOutput : public OutputLayer
{
  public:
    NormalOutput(unsigned int num_inputs, unsigned int num_outputs);
    std::unique_ptr<Distribution> forward(torch::Tensor x) override;

  private:
    torch::nn::Linear linear_loc;
    torch::Tensor scale_log;
};

class PolicyImpl : public torch::nn::Module
{
  public:
    PolicyImpl(ActionSpace action_space, std::shared_ptr<NNBase> base, bool normalize_observations = false);

    // rollout side: {value, action, action_log_probs, rnn_hxs} (policy.cpp:52-81)
    std::vector<torch::Tensor> act(torch::Tensor inputs, torch::Tensor rnn_hxs, torch::Tensor masks) const;
    // {value [B,1], action_log_probs [B,1], entropy [], rnn_hxs}; ONE fused forward kernel on CUDA tensors
    // (cpprl_b200_evaluate_actions).  The outputs carry no autograd graph: the backward pass lives inside
    // PPO::update / A2C::update.
    std::vector<torch::Tensor> evaluate_actions(torch::Tensor inputs, torch::Tensor rnn_hxs, torch::Tensor masks,
                                                torch::Tensor actions) const;
    torch::Tensor get_probs(torch::Tensor inputs, torch::Tensor rnn_hxs, torch::Tensor masks) const;
    torch::Tensor get_values(torch::Tensor inputs, torch::Tensor rnn_hxs, torch::Tensor masks) const;
    void update_observation_normalizer(torch::Tensor observations);
    bool is_recurrent() const { return base->is_recurrent(); }
    unsigned int get_hidden_size() const { return base->get_hidden_size(); }
    bool using_observation_normalizer() const { return !observation_normalizer.is_empty(); }

    // The flat fp32 parameter buffer (registration order) every parameter tensor is a view of; re-packed when a
    // parameter no longer aliases it (after to(device) / load).  `describe` fills the C-ABI policy descriptor.
    torch::Tensor flat_parameters() const;
    void describe(cpprl_b200_policy &out) const;
    const ActionSpace &get_action_space() const { return action_space; }

  private:
    ActionSpace action_space;
    std::shared_ptr<NNBase> base;
    ObservationNormalizer observation_normalizer;
    std::shared_ptr<OutputLayer> output_layer;
    mutable torch::Tensor flat_;
};
TORCH_MODULE(Policy);

// ---------------------------------------------------------------------------------------------------------
// algorithms (ppo.h:15-37, a2c.h:15-33)
// ---------------------------------------------------------------------------------------------------------
class PPO : public Algorithm
{
  public:
    PPO(Policy &policy, float clip_param, int num_epoch, int num_mini_batch, float actor_loss_coef,
        float value_loss_coef, float entropy_coef, float learning_rate, float epsilon = 1e-8,
        float max_grad_norm = 0.5, float kl_target = 0.01);
    ~PPO() override;
    PPO(const PPO &) = delete;
    PPO &operator=(const PPO &) = delete;
    // Whole update on the device: one persistent kernel per epoch, statistics read back once.  Permutations are drawn
    // with torch::randperm on the CPU generator, one per epoch, none after a KL early stop — the same RNG stream the
    // reference's generators consume, so torch::manual_seed(s) gives the same minibatches on both sides.
    std::vector<UpdateDatum> update(RolloutStorage &rollouts, float decay_level = 1) override;

  private:
    Policy &policy;
    float actor_loss_coef, value_loss_coef, entropy_coef, max_grad_norm, original_learning_rate, original_clip_param,
        kl_target, epsilon;
    int num_epoch, num_mini_batch;
    cpprl_b200_ppo *handle = nullptr;
    int64_t handle_params = 0;
};

class A2C : public Algorithm
{
  public:
    A2C(Policy &policy, float actor_loss_coef, float value_loss_coef, float entropy_coef, float learning_rate,
        float epsilon = 1e-8, float alpha = 0.99, float max_grad_norm = 0.5);
    ~A2C() override;
    A2C(const A2C &) = delete;
    A2C &operator=(const A2C &) = delete;
    std::vector<UpdateDatum> update(RolloutStorage &rollouts, float decay_level = 1) override;

  private:
    Policy &policy;
    float actor_loss_coef, value_loss_coef, entropy_coef, max_grad_norm, original_learning_rate, epsilon, alpha;
    cpprl_b200_a2c *handle = nullptr;
    int64_t handle_params = 0;
};
} // namespace cpprl
