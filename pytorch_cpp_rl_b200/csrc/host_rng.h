// host_rng.h — permutations of the reference's CPU random stream, produced by background threads (host_rng.cu).
#pragma once
#include <stdint.h>

namespace cpprl_b200
{
// `count` permutations torch::randperm(n, kLong) would draw back to back from the generator whose state blob
// (at::CPUGeneratorImpl::get_state, 5056 bytes) is given; permutation e lands in out[e] (n int64, caller-owned, may
// be pinned memory).  wait(e) blocks until permutation e is complete; state_after(e) is the generator state after
// permutation e was drawn (what torch's generator must be set to when e was the last one used).
class PermStream
{
  public:
    PermStream();
    ~PermStream();
    PermStream(const PermStream &) = delete;
    PermStream &operator=(const PermStream &) = delete;
    static bool supported(int64_t n);
    int start(const void *rng_state, int64_t n, int count, int64_t *const *out, int max_threads);
    void wait(int e);
    void state_after(int e, void *rng_state_out);
    void join();

  private:
    struct Impl;
    Impl *impl_;
};
} // namespace cpprl_b200
