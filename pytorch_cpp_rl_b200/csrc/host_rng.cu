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
// would have left it.  The threads also draw AHEAD of the update, into the second half of a ring: when the next update
// arrives with the state this one returned (nobody drew from the generator in between), its first permutation is
// already there instead of 1.3 ms away (host_rng.h).  Bit-exactness against torch.randperm is pinned by
// tests/test_host_rng_cpu.py.
//
// Host code only (no kernel here); it lives in the C-ABI library because the reference's host language is C++.
#include <algorithm>
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
// `avail` counts the raw draws the producer has written so far: the walk follows the Mersenne Twister a block behind
// instead of waiting for all n - 1 draws (the producer is about twice as fast as the walk, so the wait is one block).
constexpr int64_t kHandOff = 8192;
void walk(const uint32_t *raw, int64_t n, int64_t *out, std::vector<uint32_t> &ids, const std::atomic<int64_t> *avail)
{
    ids.resize(static_cast<size_t>(n));
    for (int64_t i = 0; i < n; ++i)
        ids[i] = static_cast<uint32_t>(i);
    const uint32_t un = static_cast<uint32_t>(n);
    int64_t ready = 0;
    for (int64_t i = 0; i + 1 < n; ++i)
    {
        if (i >= ready)
        {
            const int64_t want = std::min<int64_t>(i + kHandOff, n - 1);
            while ((ready = avail->load(std::memory_order_acquire)) < want && ready <= i)
                std::this_thread::yield();
        }
        const uint32_t j = static_cast<uint32_t>(i) + raw[i] % (un - static_cast<uint32_t>(i));
        const uint32_t a = ids[i], b = ids[j];
        ids[i] = b;
        ids[j] = a;
    }
    for (int64_t i = 0; i < n; ++i)
        out[i] = static_cast<int64_t>(ids[i]);
}
} // namespace

// ---- PermStream: a rolling stream of permutations of n ids, produced in order by background threads --------------------
// Permutations carry a global index g = 0, 1, ... from the state the stream was (re)started with; g lives in ring slot
// g % slots.  The producer may run up to `slots` permutations past `base`, the first permutation of the current update.
struct PermStream::Impl
{
    int64_t n = 0;
    int count = 0, slots = 0, max_threads = 0;
    int64_t *ring = nullptr;
    long long base = 0;     // global index of the current update's first permutation
    long long produced = 0; // all raw draws (and the state after them) exist for [.., produced)
    long long started = 0;  // the producer has begun (or finished) the draws of [.., started): their walks may start
    long long next_job = 0; // next global index a worker walks
    std::unique_ptr<std::atomic<int64_t>[]> filled; // per slot: raw draws written so far (producer -> walker hand-off)
    std::vector<std::vector<uint32_t>> raw;              // per slot: raw draws of its permutation
    std::vector<std::vector<unsigned char>> state_after; // per slot: generator state after its permutation
    std::vector<long long> raw_g, perm_g;                // per slot: which g its draws / its finished permutation belong to
                                                         // (-1: none yet, -2: being written)
    bool quit = false;
    long long restarts = 0;
    std::mutex mu;
    std::condition_variable cv;
    std::vector<std::thread> threads;
    Mt19937 mt;
    std::vector<unsigned char> blob;

    void producer()
    {
        for (;;)
        {
            long long g;
            int s;
            {
                std::unique_lock<std::mutex> lk(mu);
                cv.wait(lk, [&] {
                    if (quit)
                        return true;
                    if (produced >= base + slots)
                        return false;
                    // the slot's previous permutation (never asked for after an early stop, perhaps) must have left the walk
                    const int sl = static_cast<int>(produced % slots);
                    return produced < slots || perm_g[sl] == produced - slots;
                });
                if (quit)
                    return;
                g = produced;
                s = static_cast<int>(g % slots);
                raw_g[s] = -2;
            }
            const int64_t draws = n > 0 ? n - 1 : 0;
            raw[s].resize(static_cast<size_t>(draws));
            filled[s].store(0, std::memory_order_relaxed);
            {
                std::lock_guard<std::mutex> lk(mu);
                started = g + 1;
            }
            cv.notify_all();
            for (int64_t off = 0; off < draws; off += kHandOff)
            {
                const int64_t len = std::min<int64_t>(kHandOff, draws - off);
                mt.fill(raw[s].data() + off, len);
                filled[s].store(off + len, std::memory_order_release);
            }
            state_after[s] = blob;
            store_state(state_after[s].data(), mt);
            {
                std::lock_guard<std::mutex> lk(mu);
                raw_g[s] = g;
                produced = g + 1;
            }
            cv.notify_all();
        }
    }
    void worker()
    {
        std::vector<uint32_t> ids;
        for (;;)
        {
            long long g;
            {
                std::unique_lock<std::mutex> lk(mu);
                cv.wait(lk, [&] { return quit || next_job < started; });
                if (quit)
                    return;
                g = next_job++;
            }
            const int s = static_cast<int>(g % slots);
            walk(raw[s].data(), n, ring + static_cast<int64_t>(s) * n, ids, &filled[s]);
            {
                std::lock_guard<std::mutex> lk(mu);
                perm_g[s] = g;
            }
            cv.notify_all();
        }
    }
    void stop()
    {
        {
            std::lock_guard<std::mutex> lk(mu);
            quit = true;
        }
        cv.notify_all();
        for (auto &t : threads)
            if (t.joinable())
                t.join();
        threads.clear();
        quit = false;
    }
};

PermStream::PermStream() : impl_(new Impl) {}
PermStream::~PermStream()
{
    impl_->stop();
    delete impl_;
}

bool PermStream::supported(int64_t n)
{
    // randperm_cpu keeps this algorithm ("preserve old behavior") below uint32 max / 20; larger n draws 64-bit numbers
    return n >= 0 && n < static_cast<int64_t>(0xffffffffu / 20);
}

int PermStream::start(const void *rng_state, int64_t n, int count, int64_t *ring, int slots, int max_threads)
{
    Impl &s = *impl_;
    if (count < 0 || slots < count || (count > 0 && n > 0 && !ring))
        return fail(CPPRL_B200_EINVAL, "host_randperm: bad ring");
    if (!supported(n))
        return fail(CPPRL_B200_EUNSUPPORTED, "host_randperm: n outside the range of randperm's 32-bit branch");
    if (!s.threads.empty() && n == s.n && count == s.count && slots == s.slots && ring == s.ring && max_threads == s.max_threads)
    {
        // does the incoming state continue the stream?  It does when it is the state recorded after one of the permutations
        // of the window the ring still holds (the full update: base + count - 1; an early stop: an earlier one)
        std::lock_guard<std::mutex> lk(s.mu);
        for (int sl = 0; sl < slots; ++sl)
        {
            const long long g = s.raw_g[sl];
            if (g >= s.base && std::memcmp(s.state_after[sl].data(), rng_state, kStateBytes) == 0)
            {
                s.base = g + 1;
                s.cv.notify_all();
                return 0;
            }
        }
    }
    s.stop();
    Mt19937 mt;
    if (!load_state(rng_state, mt))
        return fail(CPPRL_B200_EINVAL, "host_randperm: not a seeded at::CPUGeneratorImpl state (5056 bytes expected)");
    s.mt = mt;
    s.blob.assign(static_cast<const unsigned char *>(rng_state), static_cast<const unsigned char *>(rng_state) + kStateBytes);
    s.n = n, s.count = count, s.slots = slots, s.ring = ring, s.max_threads = max_threads;
    s.base = s.produced = s.started = s.next_job = 0;
    s.filled.reset(new std::atomic<int64_t>[static_cast<size_t>(slots > 0 ? slots : 1)]);
    s.raw.assign(static_cast<size_t>(slots), {});
    s.state_after.assign(static_cast<size_t>(slots), {});
    s.raw_g.assign(static_cast<size_t>(slots), -1);
    s.perm_g.assign(static_cast<size_t>(slots), -1);
    ++s.restarts;
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

const int64_t *PermStream::wait(int e)
{
    Impl &s = *impl_;
    const long long g = s.base + e;
    const int sl = static_cast<int>(g % s.slots);
    std::unique_lock<std::mutex> lk(s.mu);
    s.cv.wait(lk, [&] { return s.perm_g[sl] == g; });
    return s.ring + static_cast<int64_t>(sl) * s.n;
}

void PermStream::state_after(int e, void *rng_state_out)
{
    Impl &s = *impl_;
    const long long g = s.base + e;
    const int sl = static_cast<int>(g % s.slots);
    std::unique_lock<std::mutex> lk(s.mu);
    s.cv.wait(lk, [&] { return s.raw_g[sl] == g; });
    std::memcpy(rng_state_out, s.state_after[sl].data(), kStateBytes);
}

void PermStream::stop() { impl_->stop(); }
long long PermStream::restarts() const { return impl_->restarts; }
} // namespace cpprl_b200

using namespace cpprl_b200;

extern "C" int64_t cpprl_b200_rng_state_bytes(void) { return static_cast<int64_t>(kStateBytes); }

extern "C" int cpprl_b200_host_randperm(void *rng_state, int64_t n, int64_t count, int64_t *out, int max_threads)
{
    if (!rng_state || (!out && n * count > 0) || n < 0 || count < 0 || count > (1 << 20))
        return fail(CPPRL_B200_EINVAL, "host_randperm: bad arguments");
    if (count == 0)
        return 0;
    PermStream ps; // the caller's buffer is the ring: `count` slots, nothing drawn ahead
    if (int rc = ps.start(rng_state, n, static_cast<int>(count), out, static_cast<int>(count), max_threads))
        return rc;
    for (int e = 0; e < count; ++e)
        ps.wait(e);
    ps.state_after(static_cast<int>(count) - 1, rng_state);
    return 0;
}

// Test hook (tests/test_host_rng_cpu.py): `updates` consecutive updates on ONE stream with a ring of 2 * count slots, the
// way cpprl_b200_ppo_update_rng drives it; update u takes used[u] <= count permutations (an early stop) and hands the state
// after its last one to the next.  reseed_before[u] != 0 replaces the state with `reseed_state` before update u (someone
// else drew from the generator).  out receives the sum(used) permutations in order; *restarts the number of stream restarts.
extern "C" int cpprl_b200_debug_perm_stream(void *rng_state, const void *reseed_state, int64_t n, int count, int updates,
                                            const int *used, const int *reseed_before, int64_t *out, long long *restarts)
{
    if (!rng_state || !used || !out || n < 0 || count <= 0 || updates < 0)
        return fail(CPPRL_B200_EINVAL, "debug_perm_stream: bad arguments");
    std::vector<int64_t> ring(static_cast<size_t>(2 * count) * static_cast<size_t>(n));
    PermStream ps;
    int64_t written = 0;
    for (int u = 0; u < updates; ++u)
    {
        if (reseed_before && reseed_before[u] && reseed_state)
            std::memcpy(rng_state, reseed_state, kStateBytes);
        if (used[u] < 0 || used[u] > count)
            return fail(CPPRL_B200_EINVAL, "debug_perm_stream: used[u] outside [0, count]");
        if (int rc = ps.start(rng_state, n, count, ring.data(), 2 * count, 0))
            return rc;
        for (int e = 0; e < used[u]; ++e)
        {
            std::memcpy(out + written * n, ps.wait(e), static_cast<size_t>(n) * sizeof(int64_t));
            ++written;
        }
        if (used[u] > 0)
            ps.state_after(used[u] - 1, rng_state);
    }
    if (restarts)
        *restarts = ps.restarts();
    return 0;
}
