// returns_scan.cu — GAE / n-step return scan and advantage normalisation (HBM-bound stages).
//
// Reference behaviour: RolloutStorage::compute_returns (src/storage.cpp:113-144) and the prologue of
// PPO::update (src/algorithms/ppo.cpp:51-57).
//
// Design: the recurrence is sequential in time and independent across environments, and every tensor is
// time-major [T, N], so a warp reading one time step for 32 (or 128 with float4) neighbouring environments
// touches whole 128-byte lines.  Each thread owns VEC environments and walks backwards in time; the loads
// of UNROLL steps are issued before the dependent chain consumes them, which keeps 3*UNROLL lines per warp
// in flight.  Arithmetic uses __fmul_rn/__fadd_rn in the reference's operator order so results are
// bit-identical to the CPU libtorch path (no FMA contraction).
#include <mutex>
#include <string>

#include "common.cuh"

namespace cpprl_b200
{
namespace
{
constexpr int kScanThreads = 128;
constexpr int kScanUnroll = 8;

template <int VEC>
struct Pack;
template <>
struct Pack<1>
{
    float v[1];
    __device__ static Pack load(const float *p)
    {
        Pack r;
        r.v[0] = ld_stream1(p);
        return r;
    }
    __device__ void store(float *p) const { *p = v[0]; }
};
template <>
struct Pack<4>
{
    float v[4];
    __device__ static Pack load(const float *p)
    {
        float4 q = ld_stream4(p);
        Pack r;
        r.v[0] = q.x, r.v[1] = q.y, r.v[2] = q.z, r.v[3] = q.w;
        return r;
    }
    __device__ void store(float *p) const { st_stream4(p, make_float4(v[0], v[1], v[2], v[3])); }
};

// One thread = VEC consecutive environments, all T steps.
template <int VEC, bool GAE>
__global__ void __launch_bounds__(kScanThreads)
    returns_scan_kernel(const float *__restrict__ rewards, float *__restrict__ value_preds,
                        const float *__restrict__ masks, const float *__restrict__ next_value,
                        float *__restrict__ returns, int64_t T, int64_t N, float gamma, float gamma_tau)
{
    const int64_t env = (static_cast<int64_t>(blockIdx.x) * kScanThreads + threadIdx.x) * VEC;
    if (env >= N)
        return;

    Pack<VEC> carry = Pack<VEC>::load(next_value + env); // V[t+1] (GAE) or returns[t+1] (n-step)
    Pack<VEC> gae;
#pragma unroll
    for (int i = 0; i < VEC; ++i)
        gae.v[i] = 0.f;
    if (GAE)
        carry.store(value_preds + T * N + env); // value_predictions[-1] = next_value   (storage.cpp:120)
    else
        carry.store(returns + T * N + env); // returns[-1] = next_value             (storage.cpp:135)

    for (int64_t t_hi = T; t_hi > 0; t_hi -= kScanUnroll)
    {
        Pack<VEC> r[kScanUnroll], v[kScanUnroll], m[kScanUnroll];
        // issue every load of this chunk before the dependent chain starts
#pragma unroll
        for (int u = 0; u < kScanUnroll; ++u)
        {
            const int64_t t = t_hi - 1 - u;
            if (t >= 0)
            {
                r[u] = Pack<VEC>::load(rewards + t * N + env);
                m[u] = Pack<VEC>::load(masks + (t + 1) * N + env);
                if (GAE)
                    v[u] = Pack<VEC>::load(value_preds + t * N + env);
            }
        }
#pragma unroll
        for (int u = 0; u < kScanUnroll; ++u)
        {
            const int64_t t = t_hi - 1 - u;
            if (t >= 0)
            {
                Pack<VEC> out;
#pragma unroll
                for (int i = 0; i < VEC; ++i)
                {
                    if (GAE)
                    {
                        // delta = r + gamma*V[t+1]*m[t+1] - V[t]                 (storage.cpp:124-128)
                        float d = __fmul_rn(__fmul_rn(gamma, carry.v[i]), m[u].v[i]);
                        d = __fsub_rn(__fadd_rn(r[u].v[i], d), v[u].v[i]);
                        // gae = delta + gamma*tau*m[t+1]*gae                     (storage.cpp:129)
                        gae.v[i] = __fadd_rn(d, __fmul_rn(__fmul_rn(gamma_tau, m[u].v[i]), gae.v[i]));
                        out.v[i] = __fadd_rn(gae.v[i], v[u].v[i]); //              (storage.cpp:130)
                        carry.v[i] = v[u].v[i];
                    }
                    else
                    {
                        // returns[t] = returns[t+1]*gamma*m[t+1] + r[t]          (storage.cpp:138-141)
                        out.v[i] = __fadd_rn(__fmul_rn(__fmul_rn(carry.v[i], gamma), m[u].v[i]), r[u].v[i]);
                        carry.v[i] = out.v[i];
                    }
                }
                out.store(returns + t * N + env);
            }
        }
    }
}

// ---- advantages -----------------------------------------------------------------------------
constexpr int kAdvThreads = 256;

// sums[0] = sum(ret - v), sums[1] = sum((ret - v)^2) over the first n elements, in double.  Every CTA writes its
// partial pair to ws->partials[blockIdx]; the LAST CTA to finish adds the partials in index order (one warp: lane l
// takes l, l + 32, ... then a fixed shuffle tree), so the result does not depend on the order the CTAs ran in -- the
// update gives the same bits every run.  The arrival counter cleans itself for the next launch.
struct AdvWorkspace
{
    unsigned int counter, pad;
    double partials[1][2]; // [blocks][2]
};

__global__ void __launch_bounds__(kAdvThreads)
    advantage_moments_kernel(const float *__restrict__ returns, const float *__restrict__ value_preds, int64_t n,
                             double *__restrict__ sums, AdvWorkspace *__restrict__ ws)
{
    double s = 0.0, ss = 0.0;
    const int64_t stride = static_cast<int64_t>(gridDim.x) * kAdvThreads;
    for (int64_t i = static_cast<int64_t>(blockIdx.x) * kAdvThreads + threadIdx.x; i < n; i += stride)
    {
        const float a = __fsub_rn(returns[i], value_preds[i]);
        s += a;
        ss += static_cast<double>(a) * a;
    }
    s = warp_sum(s);
    ss = warp_sum(ss);
    __shared__ double sh[2][kAdvThreads / 32];
    __shared__ bool is_last;
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    if (l == 0)
        sh[0][w] = s, sh[1][w] = ss;
    __syncthreads();
    if (w != 0)
        return;
    s = l < kAdvThreads / 32 ? sh[0][l] : 0.0;
    ss = l < kAdvThreads / 32 ? sh[1][l] : 0.0;
    s = warp_sum(s);
    ss = warp_sum(ss);
    if (l == 0)
    {
        __stcg(&ws->partials[blockIdx.x][0], s);
        __stcg(&ws->partials[blockIdx.x][1], ss);
        __threadfence();
        is_last = atomicAdd(&ws->counter, 1u) == gridDim.x - 1;
    }
    __syncwarp();
    if (!is_last)
        return;
    __threadfence();
    s = 0.0, ss = 0.0;
    for (unsigned int b = l; b < gridDim.x; b += 32)
    {
        s += __ldcg(&ws->partials[b][0]);
        ss += __ldcg(&ws->partials[b][1]);
    }
    s = warp_sum(s);
    ss = warp_sum(ss);
    if (l == 0)
    {
        sums[0] = s;
        sums[1] = ss;
        ws->counter = 0u;
    }
}

// adv = ret - v, optionally (adv - mean) / (std + 1e-5); moments[2], moments[3] <- mean, std.
__global__ void __launch_bounds__(kAdvThreads)
    advantage_write_kernel(const float *__restrict__ returns, const float *__restrict__ value_preds, int64_t n,
                           int normalize, float *__restrict__ adv, double *__restrict__ moments)
{
    const double cnt = static_cast<double>(n);
    const double mean_d = moments[0] / cnt;
    double var_d = (moments[1] - moments[0] * mean_d) / (cnt - 1.0); // Bessel (ppo.cpp:57, Tensor::std())
    if (var_d < 0.0)
        var_d = 0.0;
    const float mean = static_cast<float>(mean_d);
    const float denom = __fadd_rn(static_cast<float>(sqrt(var_d)), 1e-5f);
    if (blockIdx.x == 0 && threadIdx.x == 0)
    {
        moments[2] = mean_d;
        moments[3] = sqrt(var_d);
    }
    const int64_t stride = static_cast<int64_t>(gridDim.x) * kAdvThreads;
    for (int64_t i = static_cast<int64_t>(blockIdx.x) * kAdvThreads + threadIdx.x; i < n; i += stride)
    {
        float a = __fsub_rn(returns[i], value_preds[i]);
        if (normalize)
            a = __fdiv_rn(__fsub_rn(a, mean), denom);
        adv[i] = a;
    }
}
} // namespace
} // namespace cpprl_b200

namespace cpprl_b200
{
static int64_t advantage_blocks(int64_t n)
{
    int64_t blocks = (n + kAdvThreads * 4 - 1) / (kAdvThreads * 4);
    const int64_t cap = static_cast<int64_t>(sm_count()) * 8;
    if (blocks > cap)
        blocks = cap;
    return blocks < 1 ? 1 : blocks;
}

// bytes of the partial-sum workspace launch_advantage_moments needs (zeroed ONCE by its owner: the arrival counter
// returns to zero at the end of every launch)
size_t advantage_workspace_bytes() { return sizeof(AdvWorkspace) + static_cast<size_t>(sm_count()) * 8 * 2 * sizeof(double); }

// sums[0] = sum(adv), sums[1] = sum(adv^2), deterministic (see advantage_moments_kernel).
int launch_advantage_moments(const float *d_returns, const float *d_value_preds, int64_t n, double *d_sums, void *d_ws,
                             cudaStream_t s)
{
    advantage_moments_kernel<<<static_cast<unsigned>(advantage_blocks(n)), kAdvThreads, 0, s>>>(
        d_returns, d_value_preds, n, d_sums, static_cast<AdvWorkspace *>(d_ws));
    CPPRL_LAUNCH_CHECK("advantage_moments_kernel");
    return 0;
}
} // namespace cpprl_b200

using namespace cpprl_b200;

extern "C" int cpprl_b200_compute_returns(const float *d_rewards, float *d_value_preds, const float *d_masks,
                                          const float *d_next_value, float *d_returns, int64_t T, int64_t N,
                                          int use_gae, float gamma, float tau, void *stream)
{
    if (!d_rewards || !d_value_preds || !d_masks || !d_next_value || !d_returns)
        return fail(CPPRL_B200_EINVAL, "compute_returns: null pointer");
    if (T < 0 || N < 0)
        return fail(CPPRL_B200_EINVAL, "compute_returns: negative size");
    if (int rc = require_device())
        return rc;
    if (N == 0)
        return 0;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const float gamma_tau = gamma * tau; // float*float, rounded once before touching a tensor (storage.cpp:129)
    auto aligned16 = [](const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; };
    const bool vec4 = (N % 4 == 0) && aligned16(d_rewards) && aligned16(d_value_preds) && aligned16(d_masks) &&
                      aligned16(d_next_value) && aligned16(d_returns);
    KernelTimer timer("returns", s);
    if (vec4)
    {
        const unsigned grid = static_cast<unsigned>((N / 4 + kScanThreads - 1) / kScanThreads);
        if (use_gae)
            returns_scan_kernel<4, true><<<grid, kScanThreads, 0, s>>>(d_rewards, d_value_preds, d_masks, d_next_value,
                                                                        d_returns, T, N, gamma, gamma_tau);
        else
            returns_scan_kernel<4, false><<<grid, kScanThreads, 0, s>>>(d_rewards, d_value_preds, d_masks, d_next_value,
                                                                         d_returns, T, N, gamma, gamma_tau);
    }
    else
    {
        const unsigned grid = static_cast<unsigned>((N + kScanThreads - 1) / kScanThreads);
        if (use_gae)
            returns_scan_kernel<1, true><<<grid, kScanThreads, 0, s>>>(d_rewards, d_value_preds, d_masks, d_next_value,
                                                                        d_returns, T, N, gamma, gamma_tau);
        else
            returns_scan_kernel<1, false><<<grid, kScanThreads, 0, s>>>(d_rewards, d_value_preds, d_masks, d_next_value,
                                                                         d_returns, T, N, gamma, gamma_tau);
    }
    CPPRL_LAUNCH_CHECK("returns_scan_kernel");
    return 0;
}

extern "C" int cpprl_b200_advantages(const float *d_returns, const float *d_value_preds, int64_t T, int64_t N,
                                     int normalize, float *d_advantages, double *d_moments, void *stream)
{
    if (!d_returns || !d_value_preds || !d_advantages || !d_moments)
        return fail(CPPRL_B200_EINVAL, "advantages: null pointer");
    if (T <= 0 || N <= 0)
        return fail(CPPRL_B200_EINVAL, "advantages: non-positive size");
    if (int rc = require_device())
        return rc;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const int64_t n = T * N;
    KernelTimer timer("advantages", s);
    CPPRL_CUDA_TRY(cudaMemsetAsync(d_moments, 0, 4 * sizeof(double), s));
    // the caller's d_moments holds four doubles only: the per-CTA partials go to a stream-ordered temporary
    void *ws = nullptr;
    CPPRL_CUDA_TRY(cudaMallocAsync(&ws, advantage_workspace_bytes(), s));
    CPPRL_CUDA_TRY(cudaMemsetAsync(ws, 0, sizeof(AdvWorkspace), s));
    const int64_t blocks = advantage_blocks(n);
    int rc = launch_advantage_moments(d_returns, d_value_preds, n, d_moments, ws, s);
    CPPRL_CUDA_TRY(cudaFreeAsync(ws, s));
    if (rc)
        return rc;
    advantage_write_kernel<<<static_cast<unsigned>(blocks), kAdvThreads, 0, s>>>(d_returns, d_value_preds, n, normalize,
                                                                                 d_advantages, d_moments);
    CPPRL_LAUNCH_CHECK("advantage_write_kernel");
    return 0;
}

// ---- cpprl_b200_advantages_noalloc -------------------------------------------------------------------------------
// The same two kernels as cpprl_b200_advantages for callers that cannot allocate (CUDA graph capture, serving loops,
// harnesses that forbid cudaMalloc): the per-CTA partial sums and the arrival counter of advantage_moments_kernel live
// in a file-scope device array instead of a stream-ordered temporary; it
// has no counterpart in the reference.
//   * Limit: T * N <= CPPRL_B200_ADVANTAGES_NOALLOC_MAX_ROWS (4 Mi rows) -- and the launch shape the existing entry
//     point would use must fit the array's kNoAllocMaxBlocks partial pairs (it does on every part with <= 256 SMs).
//   * Every GPU that loads the library holds the array (32 KB + 8 B) until the process exits.
//   * Every call shares it: the arrival counter is cleared on the stream at the start of each call (the partials are
//     overwritten before they are read), and calls that OVERLAP on one GPU (different streams) race on it.
//   * No cudaMalloc / cudaFree / cudaMallocAsync, no event creation, no library that allocates.
namespace
{
constexpr int kNoAllocMaxBlocks = 2048; // partial {sum, sum of squares} pairs the array holds
static_assert(CPPRL_B200_ADVANTAGES_NOALLOC_MAX_ROWS == (1ll << 22), "header and kernel file disagree on the row limit");
// AdvWorkspace layout: {counter, pad} then kNoAllocMaxBlocks pairs of doubles
static __device__ __align__(256) unsigned char g_adv_noalloc_scratch[sizeof(AdvWorkspace) - 2 * sizeof(double) + kNoAllocMaxBlocks * 2 * sizeof(double)];

// In host code `g_adv_noalloc_scratch` is only a stand-in for the symbol: its device address comes from
// cudaGetSymbolAddress, once per device, with that device current.
std::mutex g_noalloc_mu;
void *g_noalloc_ptr[64] = {};

int noalloc_scratch(void **out)
{
    int dev = 0;
    CPPRL_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64)
        return fail(CPPRL_B200_EUNSUPPORTED, "advantages_noalloc: device ordinal above 63");
    std::lock_guard<std::mutex> lock(g_noalloc_mu);
    if (!g_noalloc_ptr[dev])
    {
        void *p = nullptr;
        CPPRL_CUDA_TRY(cudaGetSymbolAddress(&p, g_adv_noalloc_scratch));
        if (!p)
            return fail(CPPRL_B200_ECUDA, "advantages_noalloc: cudaGetSymbolAddress returned a null address");
        g_noalloc_ptr[dev] = p;
    }
    *out = g_noalloc_ptr[dev];
    return 0;
}
} // namespace

extern "C" int cpprl_b200_advantages_noalloc(const float *d_returns, const float *d_value_preds, int64_t T, int64_t N,
                                             int normalize, float *d_advantages, double *d_moments, void *stream)
{
    if (!d_returns || !d_value_preds || !d_advantages || !d_moments)
        return fail(CPPRL_B200_EINVAL, "advantages_noalloc: null pointer");
    if (T <= 0 || N <= 0)
        return fail(CPPRL_B200_EINVAL, "advantages_noalloc: non-positive size");
    if (T > CPPRL_B200_ADVANTAGES_NOALLOC_MAX_ROWS || N > CPPRL_B200_ADVANTAGES_NOALLOC_MAX_ROWS ||
        T * N > CPPRL_B200_ADVANTAGES_NOALLOC_MAX_ROWS)
        return fail(CPPRL_B200_EUNSUPPORTED, "advantages_noalloc: T * N exceeds the static scratch's limit of " +
                                                 std::to_string(CPPRL_B200_ADVANTAGES_NOALLOC_MAX_ROWS) +
                                                 " rows; use cpprl_b200_advantages");
    if (int rc = require_device())
        return rc;
    const int64_t n = T * N;
    const int64_t blocks = advantage_blocks(n); // the launch shape of cpprl_b200_advantages: bit-identical results
    if (blocks > kNoAllocMaxBlocks)
        return fail(CPPRL_B200_EUNSUPPORTED, "advantages_noalloc: this device's launch shape needs more partial sums than the static scratch holds");
    void *ws = nullptr;
    if (int rc = noalloc_scratch(&ws))
        return rc;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    CPPRL_CUDA_TRY(cudaMemsetAsync(d_moments, 0, 4 * sizeof(double), s));
    CPPRL_CUDA_TRY(cudaMemsetAsync(ws, 0, sizeof(unsigned int) * 2, s)); // the arrival counter must start at zero
    if (int rc = launch_advantage_moments(d_returns, d_value_preds, n, d_moments, ws, s))
        return rc;
    advantage_write_kernel<<<static_cast<unsigned>(blocks), kAdvThreads, 0, s>>>(d_returns, d_value_preds, n, normalize,
                                                                                 d_advantages, d_moments);
    CPPRL_LAUNCH_CHECK("advantage_write_kernel");
    return 0;
}
