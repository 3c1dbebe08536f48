// host_rng.cu — the reference's minibatch permutations, drawn on the host faster than torch::randperm draws them.
//
// FeedForwardGenerator draws ONE torch::randperm(T*N, kLong) per epoch from the default CPU generator
// (src/generators/feed_forward_generator.cpp:31-34).  A drop-in must consume that same random stream: identical
// seeds have to give identical minibatches ("indices bit-exact").  torch's CPU randperm is a serial
// Fisher-Yates walk, `z = mt19937() % (n - i); swap(r[i], r[i + z])`, ~6 ns per element — 1.5 ms per epoch at
// cfg3 (262 144 rows), ten epochs per update: longer than the whole update takes on the GPU.
//
// This file restates that algorithm (at::mt19937 + the small-n branch of randperm_cpu) so that an update's epochs
// overlap: one thread advances the Mersenne Twister and hands each epoch's raw draws to a worker, the workers run the
// epochs' swap walks in parallel (the walk of epoch e needs only its own draws), and the caller picks the
// permutations up in order while the GPU runs the previous epoch.  The generator state goes in and out as the
// 5056-byte blob of at::CPUGeneratorImpl::get_state / set_state, so torch's generator ends where torch::randperm
// would have left it.  Bit-exactness against torch.randperm is pinned by tests/test_host_rng_cpu.py.
//
// Host code only (no kernel here); it lives in the C-ABI library because the reference's host language is C++.
#include <atomic>
#include <condition_variable>
#include <cstdint>
#include <cstring>
#include <memory>
#include <mutex>
#include <thread>
#include <vector>

#include "common.cuh"
#include "host_rng.h"

namespace cpprl_b200
{
namespace
{
constexpr int kMtN = 624, kMtM = 397;
// at::CPUGeneratorImplState (legacy POD + the float normal cache): field offsets in the 5056-byte blob
constexpr size_t kOffLeft = 8, kOffSeeded = 12, kOffNext = 16, kOffState = 24;
constexpr size_t kStateBytes = 5056;

struct Mt19937
{
    uint32_t state[kMtN];
    int left;
    uint32_t next;

    void refill()
    {
        auto twist = [](uint32_t u, uint32_t v) {
            return (((u & 0x80000000u) | (v & 0x7fffffffu)) >> 1) ^ ((v & 1u) ? 0x9908b0dfu : 0u);
        };
        uint32_t *p = state;
        left = kMtN;
        next = 0;
        for (int j = kMtN - kMtM + 1; --j; p++)
            *p = p[kMtM] ^ twist(p[0], p[1]);
        for (int j = kMtM; --j; p++)
            *p = p[kMtM - kMtN] ^ twist(p[0], p[1]);
        *p = p[kMtM - kMtN] ^ twist(p[0], state[0]);
    }
    static uint32_t temper(uint32_t y)
    {
        y ^= (y >> 11);
        y ^= (y << 7) & 0x9d2c5680u;
        y ^= (y << 15) & 0xefc60000u;
        y ^= (y >> 18);
        return y;
    }
    // `count` outputs of at::mt19937::operator() into out
    void fill(uint32_t *out, int64_t count)
    {
        int64_t done = 0;
        while (done < count)
        {
            if (left <= 1) // operator(): if (--left == 0) next_state();
                refill();
            else
                --left;
            // after the decrement `left` outputs are available starting at state[next]; take as many as fit, keeping
            // the invariant (left counts the current output) by consuming left-1 further decrements lazily
            int64_t take = left < count - done ? left : count - done;
            for (int64_t i = 0; i < take; ++i)
                out[done + i] = temper(state[next + i]);
            next += static_cast<uint32_t>(take);
            left -= static_cast<int>(take - 1);
            done += take;
        }
    }
};

bool load_state(const void *blob, Mt19937 &mt)
{
    const unsigned char *b = static_cast<const unsigned char *>(blob);
    int32_t left, seeded;
    uint64_t next;
    std::memcpy(&left, b + kOffLeft, 4);
    std::memcpy(&seeded, b + kOffSeeded, 4);
    std::memcpy(&next, b + kOffNext, 8);
    if (!seeded || left < 0 || left > kMtN || next > static_cast<uint64_t>(kMtN))
        return false;
    for (int i = 0; i < kMtN; ++i)
    {
        uint64_t v;
        std::memcpy(&v, b + kOffState + 8 * i, 8);
        mt.state[i] = static_cast<uint32_t>(v);
    }
    mt.left = left;
    mt.next = static_cast<uint32_t>(next);
    return true;
}

void store_state(void *blob, const Mt19937 &mt)
{
    unsigned char *b = static_cast<unsigned char *>(blob);
    const int32_t left = mt.left;
    const uint64_t next = mt.next;
    std::memcpy(b + kOffLeft, &left, 4);
    std::memcpy(b + kOffNext, &next, 8);
    for (int i = 0; i < kMtN; ++i)
    {
        const uint64_t v = mt.state[i];
        std::memcpy(b + kOffState + 8 * i, &v, 8);
    }
}

// The walk of randperm_cpu's small-n branch on the raw draws of one permutation.  It moves 4-byte ids, so the whole
// array stays in the core's L2, and the result is widened to the int64 ids the generator hands out.  The 32-bit
// modulo equals torch's `int64_t(random()) % (n - i)` (both operands are below 2^32) and, having no dependence on the
// swaps, overlaps them in the out-of-order window (0.8 ms per 262 144 ids against 1.4 ms as two passes).
void walk(const uint32_t *raw, int64_t n, int64_t *out, std::vector<uint32_t> &ids)
{
    ids.resize(static_cast<size_t>(n));
    for (int64_t i = 0; i < n; ++i)
        ids[i] = static_cast<uint32_t>(i);
    const uint32_t un = static_cast<uint32_t>(n);
    for (int64_t i = 0; i + 1 < n; ++i)
    {
        const uint32_t j = static_cast<uint32_t>(i) + raw[i] % (un - static_cast<uint32_t>(i));
        const uint32_t a = ids[i], b = ids[j];
        ids[i] = b;
        ids[j] = a;
    }
    for (int64_t i = 0; i < n; ++i)
        out[i] = static_cast<int64_t>(ids[i]);
}
} // namespace

// ---- PermStream: `count` permutations of n ids, produced in order by background threads ------------------------------
struct PermStream::Impl
{
    int64_t n = 0;
    int count = 0;
    std::vector<int64_t *> out;
    std::vector<std::vector<uint32_t>> raw;              // raw draws per permutation
    std::vector<std::vector<unsigned char>> state_after; // generator state after permutation e
    std::vector<int> raw_ready, perm_ready;
    std::mutex mu;
    std::condition_variable cv;
    std::vector<std::thread> threads;
    Mt19937 mt;
    std::vector<unsigned char> blob;
    std::atomic<int> next_job{0};

    void producer()
    {
        for (int e = 0; e < count; ++e)
        {
            raw[e].resize(static_cast<size_t>(n > 0 ? n - 1 : 0));
            mt.fill(raw[e].data(), n > 0 ? n - 1 : 0);
            state_after[e] = blob;
            store_state(state_after[e].data(), mt);
            {
                std::lock_guard<std::mutex> lk(mu);
                raw_ready[e] = 1;
            }
            cv.notify_all();
        }
    }
    void worker()
    {
        std::vector<uint32_t> ids;
        for (;;)
        {
            const int e = next_job.fetch_add(1);
            if (e >= count)
                return;
            {
                std::unique_lock<std::mutex> lk(mu);
                cv.wait(lk, [&] { return raw_ready[e] != 0; });
            }
            walk(raw[e].data(), n, out[e], ids);
            std::vector<uint32_t>().swap(raw[e]);
            {
                std::lock_guard<std::mutex> lk(mu);
                perm_ready[e] = 1;
            }
            cv.notify_all();
        }
    }
};

PermStream::PermStream() : impl_(new Impl) {}
PermStream::~PermStream()
{
    join();
    delete impl_;
}

bool PermStream::supported(int64_t n)
{
    // randperm_cpu keeps this algorithm ("preserve old behavior") below uint32 max / 20; larger n draws 64-bit numbers
    return n >= 0 && n < static_cast<int64_t>(0xffffffffu / 20);
}

int PermStream::start(const void *rng_state, int64_t n, int count, int64_t *const *out, int max_threads)
{
    join();
    Impl &s = *impl_;
    if (!supported(n) || count < 0)
        return fail(CPPRL_B200_EUNSUPPORTED, "host_randperm: n outside the range of randperm's 32-bit branch");
    s.blob.assign(static_cast<const unsigned char *>(rng_state), static_cast<const unsigned char *>(rng_state) + kStateBytes);
    if (!load_state(rng_state, s.mt))
        return fail(CPPRL_B200_EINVAL, "host_randperm: not a seeded at::CPUGeneratorImpl state (5056 bytes expected)");
    s.n = n;
    s.count = count;
    s.out.assign(out, out + count);
    s.raw.assign(count, {});
    s.state_after.assign(count, {});
    s.raw_ready.assign(count, 0);
    s.perm_ready.assign(count, 0);
    s.next_job = 0;
    if (count == 0)
        return 0;
    int hw = static_cast<int>(std::thread::hardware_concurrency());
    if (hw <= 0)
        hw = 4;
    int workers = count;
    if (max_threads > 0)
        workers = std::min(workers, std::max(1, max_threads - 1));
    workers = std::max(1, std::min(workers, std::max(1, hw - 2)));
    s.threads.emplace_back([&s] { s.producer(); });
    for (int w = 0; w < workers; ++w)
        s.threads.emplace_back([&s] { s.worker(); });
    return 0;
}

void PermStream::wait(int e)
{
    Impl &s = *impl_;
    std::unique_lock<std::mutex> lk(s.mu);
    s.cv.wait(lk, [&] { return s.perm_ready[e] != 0; });
}

void PermStream::state_after(int e, void *rng_state_out)
{
    Impl &s = *impl_;
    {
        std::unique_lock<std::mutex> lk(s.mu);
        s.cv.wait(lk, [&] { return s.raw_ready[e] != 0; });
    }
    std::memcpy(rng_state_out, s.state_after[e].data(), kStateBytes);
}

void PermStream::join()
{
    for (auto &t : impl_->threads)
        if (t.joinable())
            t.join();
    impl_->threads.clear();
}
} // namespace cpprl_b200

using namespace cpprl_b200;

extern "C" int64_t cpprl_b200_rng_state_bytes(void) { return static_cast<int64_t>(kStateBytes); }

extern "C" int cpprl_b200_host_randperm(void *rng_state, int64_t n, int64_t count, int64_t *out, int max_threads)
{
    if (!rng_state || (!out && n * count > 0) || n < 0 || count < 0 || count > (1 << 20))
        return fail(CPPRL_B200_EINVAL, "host_randperm: bad arguments");
    if (count == 0)
        return 0;
    std::vector<int64_t *> outs(static_cast<size_t>(count));
    for (int64_t e = 0; e < count; ++e)
        outs[e] = out + e * n;
    PermStream ps;
    if (int rc = ps.start(rng_state, n, static_cast<int>(count), outs.data(), max_threads))
        return rc;
    for (int e = 0; e < count; ++e)
        ps.wait(e);
    ps.state_after(static_cast<int>(count) - 1, rng_state);
    ps.join();
    return 0;
}
