// normalizer.cpp — RunningMeanStd / ObservationNormalizer over cpprl_b200_rms_update and
// cpprl_b200_normalize_observations.  Reference behaviour: src/running_mean_std.cpp, src/observation_normalizer.cpp.
#include <numeric>

#include "host_common.h"

namespace cpprl
{
using detail::check;
using detail::current_stream;
using detail::fptr;

RunningMeanStdImpl::RunningMeanStdImpl(int size)
    : count(register_buffer("count", torch::full({1}, 1e-4, torch::kFloat))),
      mean(register_buffer("mean", torch::zeros({size}))),
      variance(register_buffer("variance", torch::ones({size})))
{
}

RunningMeanStdImpl::RunningMeanStdImpl(std::vector<float> means, std::vector<float> variances)
    : count(register_buffer("count", torch::full({1}, 1e-4, torch::kFloat))),
      mean(register_buffer("mean", torch::from_blob(means.data(), {static_cast<int64_t>(means.size())}).clone())),
      variance(register_buffer("variance",
                               torch::from_blob(variances.data(), {static_cast<int64_t>(variances.size())}).clone()))
{
}

void RunningMeanStdImpl::update(torch::Tensor observation)
{
    const int64_t D = mean.size(0);
    auto x = detail::dense(observation.reshape({-1, D}), "RunningMeanStd::update: observation");
    if (!mean.is_cuda())
        throw std::runtime_error("cpprl_b200: RunningMeanStd buffers must live on a CUDA device (call ->to(device))");
    const int64_t M = x.size(0);
    const int64_t need = cpprl_b200_rms_workspace_bytes(M, D);
    if (!workspace_.defined() || workspace_.numel() < need || workspace_.device() != x.device())
        workspace_ = torch::empty({need}, torch::TensorOptions(torch::kByte).device(x.device()));
    check(cpprl_b200_rms_update(fptr(x), M, D, fptr(count), fptr(mean), fptr(variance), workspace_.data_ptr(),
                                current_stream(x)));
}

ObservationNormalizerImpl::ObservationNormalizerImpl(int size, float clip_value)
    : clip(register_buffer("clip", torch::full({1}, clip_value, torch::kFloat))),
      rms(register_module("rms", RunningMeanStd(size)))
{
}

ObservationNormalizerImpl::ObservationNormalizerImpl(const std::vector<float> &means, const std::vector<float> &variances,
                                                     float clip_value)
    : clip(register_buffer("clip", torch::full({1}, clip_value, torch::kFloat))),
      rms(register_module("rms", RunningMeanStd(means, variances)))
{
}

ObservationNormalizerImpl::ObservationNormalizerImpl(const std::vector<ObservationNormalizer> &others)
    : clip(register_buffer("clip", torch::zeros({1}, torch::kFloat))), rms(nullptr)
{
    const float n = static_cast<float>(others.size());
    float clip_sum = 0;
    std::vector<float> means(others.at(0)->get_mean().size(), 0.f), variances(means.size(), 0.f);
    int total_count = 0;
    for (const auto &other : others)
    {
        clip_sum += other->get_clip_value();
        const auto m = other->get_mean(), v = other->get_variance();
        for (size_t i = 0; i < means.size(); ++i)
            means[i] += m[i], variances[i] += v[i];
        total_count += other->get_step_count();
    }
    for (size_t i = 0; i < means.size(); ++i)
        means[i] /= others.size(), variances[i] /= others.size();
    clip[0] = clip_sum / n;
    rms = register_module("rms", RunningMeanStd(means, variances));
    rms->set_count(total_count);
}

torch::Tensor ObservationNormalizerImpl::process_observation(torch::Tensor observation) const
{
    const int64_t D = rms->mean_buffer().size(0);
    auto x = detail::dense(observation.reshape({-1, D}), "ObservationNormalizer::process_observation: observation");
    if (!clip.is_cuda())
        throw std::runtime_error("cpprl_b200: ObservationNormalizer buffers must live on a CUDA device (call ->to(device))");
    auto out = torch::empty_like(x);
    check(cpprl_b200_normalize_observations(fptr(x), x.size(0), D, fptr(rms->mean_buffer()), fptr(rms->variance_buffer()),
                                            fptr(clip), fptr(out), current_stream(x)));
    return out.view(observation.sizes());
}

static std::vector<float> to_vector(const torch::Tensor &t)
{
    auto c = t.to(torch::kCPU).contiguous();
    return std::vector<float>(c.data_ptr<float>(), c.data_ptr<float>() + c.numel());
}

std::vector<float> ObservationNormalizerImpl::get_mean() const { return to_vector(rms->mean_buffer()); }
std::vector<float> ObservationNormalizerImpl::get_variance() const { return to_vector(rms->variance_buffer()); }
void ObservationNormalizerImpl::update(torch::Tensor observations) { rms->update(observations); }
} // namespace cpprl
