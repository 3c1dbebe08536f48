// host_common.h — helpers shared by the host shim translation units.
#pragma once
#include <c10/cuda/CUDAStream.h>
#include <torch/torch.h>

#include <stdexcept>
#include <string>

#include "cpprl/b200/host_api.h"
#include "cpprl_b200.h"

namespace cpprl
{
namespace detail
{
// non-zero C-ABI status -> std::runtime_error carrying the library's message (the reference's error convention is
// C++ exceptions: SURVEY.md §8b)
inline void check(int rc)
{
    if (rc != 0)
        throw std::runtime_error(std::string(cpprl_b200_last_error()));
}

inline void *current_stream(const torch::Tensor &t)
{
    return static_cast<void *>(c10::cuda::getCurrentCUDAStream(t.device().index()).stream());
}

// The update path has no CPU fallback: tensors must be CUDA fp32 (or int64 for Discrete actions) and contiguous.
inline const torch::Tensor &need_cuda(const torch::Tensor &t, const char *what)
{
    if (!t.defined() || !t.is_cuda())
        throw std::runtime_error(std::string("cpprl_b200: ") + what + " must live on a CUDA device (there is no CPU fallback)");
    return t;
}

inline torch::Tensor dense(const torch::Tensor &t, const char *what) { return need_cuda(t, what).contiguous(); }

inline float *fptr(const torch::Tensor &t) { return t.data_ptr<float>(); }
} // namespace detail
} // namespace cpprl
