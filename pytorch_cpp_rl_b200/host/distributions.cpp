// distributions.cpp — Categorical / Normal as the output layers return them (rollout side: act, get_probs).
// Reference behaviour: src/distributions/{distribution,categorical,normal}.cpp.
#include <cmath>
#include <limits>

#include "host_common.h"

namespace cpprl
{
std::vector<int64_t> Distribution::extended_shape(c10::ArrayRef<int64_t> sample_shape)
{
    std::vector<int64_t> out(sample_shape.begin(), sample_shape.end());
    out.insert(out.end(), batch_shape.begin(), batch_shape.end());
    out.insert(out.end(), event_shape.begin(), event_shape.end());
    return out;
}

Categorical::Categorical(const torch::Tensor *probs, const torch::Tensor *logits)
{
    if ((probs == nullptr) == (logits == nullptr))
        throw std::runtime_error("Either probs or logits is required, but not both");
    if (probs != nullptr)
    {
        if (probs->dim() < 1)
            throw std::runtime_error("Probabilities tensor must have at least one dimension");
        probs_ = (*probs / probs->sum(-1, true)).clamp(1.21e-7, 1. - 1.21e-7);
        logits_ = probs_.log();
    }
    else
    {
        if (logits->dim() < 1)
            throw std::runtime_error("Logits tensor must have at least one dimension");
        logits_ = *logits - logits->logsumexp(-1, true);
        probs_ = torch::softmax(logits_, -1);
    }
    num_events_ = probs_.size(-1);
    auto sizes = probs_.sizes().vec();
    if (sizes.size() > 1)
        batch_shape.assign(sizes.begin(), sizes.end() - 1);
}

torch::Tensor Categorical::entropy() { return -(logits_ * probs_).sum(-1); }

torch::Tensor Categorical::log_prob(torch::Tensor value)
{
    auto idx = value.to(torch::kLong).unsqueeze(-1);
    auto both = torch::broadcast_tensors({idx, logits_});
    return both[1].gather(-1, both[0].narrow(-1, 0, 1)).squeeze(-1);
}

torch::Tensor Categorical::sample(c10::ArrayRef<int64_t> sample_shape)
{
    auto shape = extended_shape(sample_shape);
    int64_t draws = 1;
    for (auto s : sample_shape)
        draws *= s;
    auto flat = probs_.reshape({-1, num_events_});
    auto picked = torch::multinomial(flat, draws, true); // [batch, draws]
    if (shape.empty())
        return picked.reshape({}).contiguous();
    return picked.transpose(0, 1).contiguous().view(shape);
}

Normal::Normal(const torch::Tensor loc, const torch::Tensor scale)
{
    auto both = torch::broadcast_tensors({loc, scale});
    loc_ = both[0];
    scale_ = both[1];
    batch_shape = loc_.sizes().vec();
}

torch::Tensor Normal::entropy() { return (0.5 + 0.5 * std::log(2 * M_PI) + scale_.log()).sum(-1); }

torch::Tensor Normal::log_prob(torch::Tensor value)
{
    auto variance = scale_.pow(2);
    return -(value - loc_).pow(2) / (2 * variance) - scale_.log() - std::log(std::sqrt(2 * M_PI));
}

torch::Tensor Normal::sample(c10::ArrayRef<int64_t> sample_shape)
{
    auto shape = extended_shape(sample_shape);
    torch::NoGradGuard no_grad;
    return at::normal(loc_.expand(shape), scale_.expand(shape));
}
} // namespace cpprl
