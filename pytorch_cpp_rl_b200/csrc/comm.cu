// comm.cu — creation / IPC export / peer mapping of the exchange area (see comm.cuh) and the permutation
// sharding kernel of env-sharded updates.
#include <algorithm>
#include <cstring>
#include <vector>

#include "comm.cuh"
#include "common.cuh"

struct cpprl_b200_comm
{
    int rank = 0, world = 1;
    int x_stride = 0;
    size_t bytes = 0;
    cpprl_b200::ExchangeArea *local = nullptr;
    cpprl_b200::ExchangeArea *area[cpprl_b200::kMaxRanks] = {};
    bool opened[cpprl_b200::kMaxRanks] = {};
    unsigned int step_seq = 0, update_seq = 0, rms_seq = 0;
};

namespace cpprl_b200
{
CommView comm_view(const cpprl_b200_comm *c)
{
    CommView v{};
    if (!c)
    {
        v.world = 1;
        return v;
    }
    v.world = c->world;
    v.rank = c->rank;
    v.x_stride = c->x_stride;
    for (int r = 0; r < c->world; ++r)
        v.area[r] = c->area[r];
    return v;
}
unsigned int *comm_step_seq(cpprl_b200_comm *c) { return &c->step_seq; }
unsigned int *comm_update_seq(cpprl_b200_comm *c) { return &c->update_seq; }
unsigned int *comm_rms_seq(cpprl_b200_comm *c) { return &c->rms_seq; }
double *comm_local_adv(cpprl_b200_comm *c) { return c->local->adv; }
double *comm_local_rms(cpprl_b200_comm *c) { return c->local->rms[0]; }

namespace
{
// Stable compaction of the GLOBAL permutation to the rows this rank owns, minibatch by minibatch.
// Global flat id g = t*N_global + env (feed_forward_generator.cpp:51-76); the rank owns envs [env0, env0 + n_local)
// and stores them locally as t*n_local + (env - env0).  Every minibatch is cut into `chunks` pieces, one CTA each:
// pass 1 counts the owned ids of every piece, pass 2 starts each piece at the sum of the counts before it — order is
// preserved, so the result does not depend on the launch shape (one CTA per minibatch took 260 us per epoch at 8 GPUs).
__device__ __forceinline__ long long shard_local_id(long long g, long long n_global, long long env0, long long n_local)
{
    const long long t = g / n_global, env = g - t * n_global;
    return (env >= env0 && env < env0 + n_local) ? t * n_local + (env - env0) : -1;
}

__global__ void __launch_bounds__(256)
    shard_count_kernel(const long long *__restrict__ perm, long long B, int chunks, long long n_global, long long env0,
                       long long n_local, int *__restrict__ chunk_counts)
{
    const int mb = blockIdx.x / chunks, c = blockIdx.x - mb * chunks;
    const long long len = (B + chunks - 1) / chunks, lo = c * len, hi = lo + len < B ? lo + len : B;
    const long long *src = perm + static_cast<long long>(mb) * B;
    int mine = 0;
    for (long long i = lo + threadIdx.x; i < hi; i += 256)
        mine += shard_local_id(src[i], n_global, env0, n_local) >= 0;
    __shared__ int warp_tot[8];
    for (int o = 16; o > 0; o >>= 1)
        mine += __shfl_down_sync(0xffffffffu, mine, o);
    if ((threadIdx.x & 31) == 0)
        warp_tot[threadIdx.x >> 5] = mine;
    __syncthreads();
    if (threadIdx.x == 0)
    {
        int tot = 0;
        for (int w = 0; w < 8; ++w)
            tot += warp_tot[w];
        chunk_counts[blockIdx.x] = tot;
    }
}

__global__ void __launch_bounds__(256)
    shard_write_kernel(const long long *__restrict__ perm, long long B, int chunks, long long n_global, long long env0,
                       long long n_local, const int *__restrict__ chunk_counts, long long *__restrict__ out,
                       int *__restrict__ counts)
{
    const int mb = blockIdx.x / chunks, c = blockIdx.x - mb * chunks;
    const long long len = (B + chunks - 1) / chunks, lo = c * len, hi = lo + len < B ? lo + len : B;
    const long long *src = perm + static_cast<long long>(mb) * B;
    long long *dst = out + static_cast<long long>(mb) * B;
    __shared__ int warp_tot[8];
    __shared__ int base;
    if (threadIdx.x == 0)
    {
        int b = 0;
        for (int k = 0; k < c; ++k)
            b += chunk_counts[mb * chunks + k];
        base = b;
        if (c == chunks - 1)
            counts[mb] = b + chunk_counts[blockIdx.x];
    }
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (long long i0 = lo; i0 < hi; i0 += 256)
    {
        const long long i = i0 + threadIdx.x;
        const long long local = i < hi ? shard_local_id(src[i], n_global, env0, n_local) : -1;
        const unsigned int ballot = __ballot_sync(0xffffffffu, local >= 0);
        const int before = __popc(ballot & ((1u << lane) - 1u));
        if (lane == 0)
            warp_tot[warp] = __popc(ballot);
        __syncthreads();
        int offset = base;
        for (int w = 0; w < warp; ++w)
            offset += warp_tot[w];
        if (local >= 0)
            dst[offset + before] = local;
        __syncthreads();
        if (threadIdx.x == 0)
        {
            int tot = 0;
            for (int w = 0; w < 8; ++w)
                tot += warp_tot[w];
            base += tot;
        }
        __syncthreads();
    }
}
} // namespace

int shard_permutation_chunks(int64_t B) { return static_cast<int>(std::min<int64_t>(64, std::max<int64_t>(1, (B + 4095) / 4096))); }

// d_counts: n_mb totals followed by n_mb * shard_permutation_chunks(B) piece counts (scratch)
int launch_shard_permutation(const int64_t *d_perm, int64_t n_mb, int64_t B, int64_t n_global, int64_t env0,
                             int64_t n_local, int64_t *d_out, int *d_counts, cudaStream_t s)
{
    const int chunks = shard_permutation_chunks(B);
    const unsigned grid = static_cast<unsigned>(n_mb * chunks);
    int *piece = d_counts + n_mb;
    shard_count_kernel<<<grid, 256, 0, s>>>(reinterpret_cast<const long long *>(d_perm), B, chunks, n_global, env0, n_local, piece);
    CPPRL_LAUNCH_CHECK("shard_count_kernel");
    shard_write_kernel<<<grid, 256, 0, s>>>(reinterpret_cast<const long long *>(d_perm), B, chunks, n_global, env0, n_local, piece,
                                            reinterpret_cast<long long *>(d_out), d_counts);
    CPPRL_LAUNCH_CHECK("shard_write_kernel");
    return 0;
}
} // namespace cpprl_b200

using namespace cpprl_b200;

extern "C" int cpprl_b200_comm_create(cpprl_b200_comm **out, int rank, int world, int64_t max_params)
{
    if (!out || world < 1 || world > kMaxRanks || rank < 0 || rank >= world || max_params <= 0)
        return fail(CPPRL_B200_EINVAL, "comm_create: need 0 <= rank < world <= 8 and max_params > 0");
    if (int rc = require_device())
        return rc;
    auto *c = new cpprl_b200_comm();
    c->rank = rank;
    c->world = world;
    // every CTA's record is its slice of the parameters (rounded to 4, at least 4) plus the statistics slots
    c->x_stride = static_cast<int>(((max_params + 3) / 4) * 4 + (4 + 8) * kMaxCtas);
    c->bytes = sizeof(ExchangeArea) + sizeof(LLWord) * 2 * static_cast<size_t>(world) * static_cast<size_t>(c->x_stride);
    if (cudaMalloc(&c->local, c->bytes) != cudaSuccess || cudaMemset(c->local, 0, c->bytes) != cudaSuccess)
    {
        delete c;
        return fail(CPPRL_B200_ECUDA, "comm_create: cannot allocate the exchange area");
    }
    c->area[rank] = c->local;
    *out = c;
    return 0;
}

extern "C" int cpprl_b200_comm_export(cpprl_b200_comm *c, void *handle_out)
{
    if (!c || !handle_out)
        return fail(CPPRL_B200_EINVAL, "comm_export: null argument");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    cudaIpcMemHandle_t h;
    CPPRL_CUDA_TRY(cudaIpcGetMemHandle(&h, c->local));
    std::memcpy(handle_out, &h, sizeof(h));
    return 0;
}

extern "C" int cpprl_b200_comm_connect(cpprl_b200_comm *c, const void *handles)
{
    if (!c || (!handles && c->world > 1))
        return fail(CPPRL_B200_EINVAL, "comm_connect: null argument");
    for (int r = 0; r < c->world; ++r)
    {
        if (r == c->rank || c->area[r])
            continue;
        cudaIpcMemHandle_t h;
        std::memcpy(&h, static_cast<const char *>(handles) + 64 * r, sizeof(h));
        void *p = nullptr;
        cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess)
            return fail(CPPRL_B200_ENCCL, std::string("comm_connect: cannot map the exchange area of rank ") +
                                              std::to_string(r) + ": " + cudaGetErrorString(e));
        c->area[r] = static_cast<ExchangeArea *>(p);
        c->opened[r] = true;
    }
    return 0;
}

extern "C" void cpprl_b200_comm_destroy(cpprl_b200_comm *c)
{
    if (!c)
        return;
    for (int r = 0; r < c->world; ++r)
        if (c->opened[r])
            cudaIpcCloseMemHandle(c->area[r]);
    cudaFree(c->local);
    delete c;
}

extern "C" int cpprl_b200_comm_check(cpprl_b200_comm *c, void *stream)
{
    if (!c)
        return fail(CPPRL_B200_EINVAL, "comm_check: null communicator");
    // error words kernels raise in this rank's own exchange area when a peer's record never arrived
    int rms_error = 0;
    CPPRL_CUDA_TRY(cudaStreamSynchronize(static_cast<cudaStream_t>(stream)));
    CPPRL_CUDA_TRY(cudaMemcpy(&rms_error, &c->local->rms_error, sizeof(int), cudaMemcpyDeviceToHost));
    if (rms_error)
        return fail(CPPRL_B200_ENCCL, "rms_update_sharded: a peer rank never published its sums (bounded wait expired); "
                                      "the normaliser was left untouched");
    return 0;
}

extern "C" int cpprl_b200_comm_info(const cpprl_b200_comm *c, int *rank, int *world)
{
    if (!c)
        return fail(CPPRL_B200_EINVAL, "comm_info: null communicator");
    if (rank)
        *rank = c->rank;
    if (world)
        *world = c->world;
    return 0;
}
