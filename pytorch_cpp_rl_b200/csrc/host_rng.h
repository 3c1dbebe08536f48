// host_rng.h — permutations of the reference's CPU random stream, produced by background threads (host_rng.cu).
#pragma once
#include <stdint.h>

namespace cpprl_b200
{
// A rolling stream of the permutations torch::randperm(n, kLong) would draw back to back from the generator whose state
// blob (at::CPUGeneratorImpl::get_state, 5056 bytes) is given.  An update takes `count` consecutive permutations:
//   start(state, n, count, ring, slots, ...)   begin an update; `ring` holds `slots` permutations (slots * n int64,
//                                              caller-owned, may be pinned memory); slots == count or 2 * count
//   wait(e)                                    permutation e of this update (blocks until it is complete)
//   state_after(e, out)                        the generator state after permutation e was drawn (what torch's generator
//                                              must be set to when e was the last one used)
// With slots == 2 * count the threads keep drawing AHEAD of the update: the permutations the next update will ask for
// if the generator is left alone in between.  start() recognises that case by comparing the incoming state with the
// states recorded after each permutation drawn so far (any of them: an update that stopped early used fewer than
// `count`) and continues the stream where that state left it; any other state restarts the stream from it.  What comes
// out is the reference's stream either way; only when the first permutation of an update is ready changes.
class PermStream
{
  public:
    PermStream();
    ~PermStream();
    PermStream(const PermStream &) = delete;
    PermStream &operator=(const PermStream &) = delete;
    static bool supported(int64_t n);
    int start(const void *rng_state, int64_t n, int count, int64_t *ring, int slots, int max_threads);
    const int64_t *wait(int e);
    void state_after(int e, void *rng_state_out);
    void stop();                  // stops and joins the threads (also done by the destructor)
    long long restarts() const;   // how many start() calls had to restart the stream from the given state

  private:
    struct Impl;
    Impl *impl_;
};
} // namespace cpprl_b200
