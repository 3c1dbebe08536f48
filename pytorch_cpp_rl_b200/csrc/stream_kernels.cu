// stream_kernels.cu — the remaining bandwidth-bound stages: minibatch row gather, observation
// normaliser (column moments + Chan merge, normalise/clamp) and flat-buffer Adam / RMSprop steps.
#include "comm.cuh"
#include "common.cuh"

struct cpprl_b200_comm;
namespace cpprl_b200
{
CommView comm_view(const cpprl_b200_comm *c);
unsigned int *comm_rms_seq(cpprl_b200_comm *c);
} // namespace cpprl_b200

namespace cpprl_b200
{
namespace
{
// ---- gather ---------------------------------------------------------------------------------
// dst[i, :] = src[idx[i], :].  One thread per 16-byte (or 4-byte) word of the destination, so a warp
// writes full lines and reads contiguous pieces of source rows.
template <typename WordT>
__global__ void __launch_bounds__(256)
    gather_rows_kernel(const WordT *__restrict__ src, int64_t n_src_rows, int64_t words_per_row,
                       const int64_t *__restrict__ idx, int64_t n_idx, WordT *__restrict__ dst)
{
    const int64_t total = n_idx * words_per_row;
    const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
    for (int64_t w = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; w < total; w += stride)
    {
        const int64_t i = w / words_per_row;
        const int64_t c = w - i * words_per_row;
        int64_t r = idx[i];
        if (r < 0)
            r += n_src_rows; // torch advanced indexing accepts negative ids
        dst[w] = src[r * words_per_row + c];
    }
}

// ---- running mean / std ------------------------------------------------------------------------
constexpr int kRmsThreads = 256;

// Block b leaves its column sums and column sums of squares (double) in ws[b][0..2D); the last block to finish adds the
// blocks' partials in a FIXED order (slices of blocks by block index, then the slices in order: the same bits every run,
// where atomics would add in arrival order), leaves the totals in ws[0][..] and performs the Chan merge
// (running_mean_std.cpp:30-43).  The arrival counter sits behind the kRmsMaxBlocks partial rows (caller memset).
constexpr int kRmsMaxBlocks = 1024;
__global__ void __launch_bounds__(kRmsThreads)
    rms_update_kernel(const float *__restrict__ x, int64_t M, int D, double *__restrict__ ws,
                      unsigned int *__restrict__ done_counter, float *__restrict__ count,
                      float *__restrict__ mean, float *__restrict__ var, const CommView comm, unsigned int seq,
                      int *__restrict__ comm_error)
{
    extern __shared__ double sh[]; // [2][lanes_used]
    // threads 0..lanes_used-1 each stay on one column: lanes_used is a multiple of D
    const int lanes_used = (kRmsThreads / D) * D;
    const int tid = threadIdx.x;
    double s = 0.0, ss = 0.0;
    if (tid < lanes_used)
    {
        const int64_t total = M * D;
        const int64_t stride = static_cast<int64_t>(gridDim.x) * lanes_used;
        for (int64_t i = static_cast<int64_t>(blockIdx.x) * lanes_used + tid; i < total; i += stride)
        {
            const float v = x[i];
            s += v;
            ss += static_cast<double>(v) * v;
        }
        sh[tid] = s;
        sh[lanes_used + tid] = ss;
    }
    __syncthreads();
    if (tid < D)
    {
        double cs = 0.0, css = 0.0;
        for (int j = tid; j < lanes_used; j += D)
        {
            cs += sh[j];
            css += sh[lanes_used + j];
        }
        __stcg(&ws[static_cast<size_t>(blockIdx.x) * 2 * D + tid], cs);
        __stcg(&ws[static_cast<size_t>(blockIdx.x) * 2 * D + D + tid], css);
    }
    __threadfence();
    __syncthreads();
    __shared__ bool is_last;
    if (tid == 0)
        is_last = atomicAdd(done_counter, 1u) == gridDim.x - 1;
    __syncthreads();
    if (!is_last)
        return;
    __threadfence();
    {
        const int cols = 2 * D, nblocks = static_cast<int>(gridDim.x);
        const int slices = cols <= kRmsThreads ? kRmsThreads / cols : 1; // (cols > kRmsThreads: one slice, two column passes)
        for (int c0 = 0; c0 < cols; c0 += kRmsThreads)
        {
            const int c = c0 + (slices > 1 ? tid % cols : tid), j = slices > 1 ? tid / cols : 0;
            double acc = 0.0;
            if (c < cols && j < slices)
                for (int b = j; b < nblocks; b += slices)
                    acc += __ldcg(&ws[static_cast<size_t>(b) * cols + c]);
            if (c < cols && j < slices)
                sh[j * cols + c] = acc; // (slices * cols <= 2 * lanes_used doubles of dynamic shared memory)
        }
        __syncthreads();
        for (int c = tid; c < cols; c += kRmsThreads)
        {
            double total = 0.0;
            for (int j = 0; j < slices; ++j)
                total += sh[j * cols + c];
            ws[c] = total; // (every partial of block 0's row has been read: the barrier above)
        }
        __syncthreads();
    }
    double rows_total = static_cast<double>(M);
    __shared__ double sh_rows;
    if (comm.world > 1)
    {
        // env-sharded update: the batch is the union of every rank's rows.  Publish this rank's column sums and row
        // count, wait for the peers', and add them in rank order (identical merge on every rank).
        ExchangeArea *mine = comm.area[comm.rank];
        const int par = static_cast<int>(seq & 1u);
        __shared__ int peer_missing;
        if (tid == 0)
            peer_missing = 0;
        for (int c = tid; c < 2 * D; c += kRmsThreads)
            mine->rms[par][c] = __ldcg(&ws[c]);
        if (tid == 0)
            mine->rms[par][2 * D] = static_cast<double>(M);
        __threadfence_system();
        __syncthreads();
        if (tid == 0)
            flag_publish(&mine->rms_flag, seq);
        if (tid < comm.world && tid != comm.rank && !flag_wait(&comm.area[tid]->rms_flag, seq))
            peer_missing = 1;
        __syncthreads();
        if (peer_missing)
        {
            // a peer never published: merging stale or zero sums would silently corrupt the normaliser, so the
            // statistics are left untouched and the host turns the flag into CPPRL_B200_ENCCL
            if (tid == 0)
            {
                mine->rms_error = 1;
                if (comm_error)
                    *comm_error = 1;
            }
            return;
        }
        for (int c = tid; c < 2 * D; c += kRmsThreads)
        {
            double sum = 0.0;
            for (int r = 0; r < comm.world; ++r)
                sum += r == comm.rank ? __ldcg(&ws[c]) : peer_load(&comm.area[r]->rms[par][c]);
            ws[c] = sum;
        }
        if (tid == 0)
        {
            double rows = 0.0;
            for (int r = 0; r < comm.world; ++r)
                rows += r == comm.rank ? static_cast<double>(M) : peer_load(&comm.area[r]->rms[par][2 * D]);
            sh_rows = rows;
        }
        __syncthreads();
        rows_total = sh_rows;
    }
    const float old_count = count[0];
    const float batch_count = static_cast<float>(rows_total);
    const float total_count = __fadd_rn(old_count, batch_count);
    for (int c = tid; c < D; c += kRmsThreads)
    {
        const double sum = __ldcg(&ws[c]), sumsq = __ldcg(&ws[D + c]);
        const double bm_d = sum / rows_total;
        double bv_d = sumsq / rows_total - bm_d * bm_d; // population variance (var(0,false))
        if (bv_d < 0.0)
            bv_d = 0.0;
        const float batch_mean = static_cast<float>(bm_d), batch_var = static_cast<float>(bv_d);
        const float delta = __fsub_rn(batch_mean, mean[c]);
        // mean += delta * batch_count / total_count
        const float new_mean = __fadd_rn(mean[c], __fdiv_rn(__fmul_rn(delta, batch_count), total_count));
        const float m_a = __fmul_rn(var[c], old_count);
        const float m_b = __fmul_rn(batch_var, batch_count);
        const float corr =
            __fdiv_rn(__fmul_rn(__fmul_rn(__fmul_rn(delta, delta), old_count), batch_count), total_count);
        const float m2 = __fadd_rn(__fadd_rn(m_a, m_b), corr);
        mean[c] = new_mean;
        var[c] = __fdiv_rn(m2, total_count);
    }
    __syncthreads();
    if (tid == 0)
        count[0] = total_count;
}

__global__ void __launch_bounds__(256)
    normalize_obs_kernel(const float *__restrict__ x, int64_t total, int D, const float *__restrict__ mean,
                         const float *__restrict__ var, const float *__restrict__ clip, float *__restrict__ out)
{
    const float c = clip[0];
    const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
    for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += stride)
    {
        const int col = static_cast<int>(i % D);
        // (x - mean) / sqrt(var + 1e-8), clamp                     (observation_normalizer.cpp:71-73)
        const float z = __fdiv_rn(__fsub_rn(x[i], mean[col]), __fsqrt_rn(__fadd_rn(var[col], 1e-8f)));
        out[i] = fminf(fmaxf(z, -c), c);
    }
}

// ---- optimisers --------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
    adam_step_kernel(float *__restrict__ p, const float *__restrict__ g, float *__restrict__ m,
                     float *__restrict__ v, int64_t n, float step_size, float bc2_sqrt, float beta1, float beta2,
                     float eps)
{
    const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n)
        return;
    const float gi = g[i];
    const float mi = __fadd_rn(__fmul_rn(m[i], beta1), __fmul_rn(gi, 1.f - beta1));
    const float vi = __fadd_rn(__fmul_rn(v[i], beta2), __fmul_rn(__fmul_rn(gi, gi), 1.f - beta2));
    const float denom = __fadd_rn(__fdiv_rn(__fsqrt_rn(vi), bc2_sqrt), eps);
    m[i] = mi;
    v[i] = vi;
    p[i] = __fsub_rn(p[i], __fmul_rn(step_size, __fdiv_rn(mi, denom)));
}

__global__ void __launch_bounds__(256)
    rmsprop_step_kernel(float *__restrict__ p, const float *__restrict__ g, float *__restrict__ sq, int64_t n,
                        float lr, float alpha, float eps)
{
    const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n)
        return;
    const float gi = g[i];
    const float si = __fadd_rn(__fmul_rn(sq[i], alpha), __fmul_rn(__fmul_rn(gi, gi), 1.f - alpha));
    sq[i] = si;
    p[i] = __fsub_rn(p[i], __fmul_rn(lr, __fdiv_rn(gi, __fadd_rn(__fsqrt_rn(si), eps))));
}
} // namespace
} // namespace cpprl_b200

using namespace cpprl_b200;

extern "C" int cpprl_b200_gather_rows(const void *d_src, int64_t n_src_rows, int64_t row_bytes, const int64_t *d_idx,
                                      int64_t n_idx, void *d_dst, void *stream)
{
    if (n_idx == 0)
        return 0; // empty minibatch: nothing to copy (pointers of empty tensors may be null)
    if (!d_src || !d_idx || !d_dst)
        return fail(CPPRL_B200_EINVAL, "gather_rows: null pointer");
    if (row_bytes <= 0 || row_bytes % 4 != 0 || n_idx < 0 || n_src_rows <= 0)
        return fail(CPPRL_B200_EINVAL, "gather_rows: row_bytes must be a positive multiple of 4");
    if (int rc = require_device())
        return rc;
    if (n_idx == 0)
        return 0;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const bool wide = row_bytes % 16 == 0 && (reinterpret_cast<uintptr_t>(d_src) & 15u) == 0 &&
                      (reinterpret_cast<uintptr_t>(d_dst) & 15u) == 0;
    const int64_t wpr = wide ? row_bytes / 16 : row_bytes / 4;
    const int64_t total = n_idx * wpr;
    int64_t blocks = (total + 255) / 256;
    const int64_t cap = static_cast<int64_t>(sm_count()) * 16;
    if (blocks > cap)
        blocks = cap;
    KernelTimer timer("gather", s);
    if (wide)
        gather_rows_kernel<uint4><<<static_cast<unsigned>(blocks), 256, 0, s>>>(
            static_cast<const uint4 *>(d_src), n_src_rows, wpr, d_idx, n_idx, static_cast<uint4 *>(d_dst));
    else
        gather_rows_kernel<uint32_t><<<static_cast<unsigned>(blocks), 256, 0, s>>>(
            static_cast<const uint32_t *>(d_src), n_src_rows, wpr, d_idx, n_idx, static_cast<uint32_t *>(d_dst));
    CPPRL_LAUNCH_CHECK("gather_rows_kernel");
    return 0;
}

extern "C" int64_t cpprl_b200_rms_workspace_bytes(int64_t /*M*/, int64_t D)
{
    return D <= 0 ? 0 : static_cast<int64_t>(sizeof(double)) * 2 * D * kRmsMaxBlocks + 16;
}

namespace cpprl_b200
{
int launch_rms_update(const float *d_x, int64_t M, int64_t D, float *d_count, float *d_mean, float *d_var,
                      void *d_workspace, cpprl_b200_comm *comm, cudaStream_t s, int *d_comm_error)
{
    if (!d_x || !d_count || !d_mean || !d_var || !d_workspace)
        return fail(CPPRL_B200_EINVAL, "rms_update: null pointer");
    if (M <= 0 || D <= 0)
        return fail(CPPRL_B200_EINVAL, "rms_update: non-positive size");
    if (D > kRmsThreads || D > kXMaxDim)
        return fail(CPPRL_B200_EUNSUPPORTED, "rms_update: D > 256 is not implemented");
    if (int rc = require_device())
        return rc;
    KernelTimer timer("rms", s);
    double *ws = static_cast<double *>(d_workspace);
    unsigned int *counter = reinterpret_cast<unsigned int *>(ws + static_cast<size_t>(2) * D * kRmsMaxBlocks);
    CPPRL_CUDA_TRY(cudaMemsetAsync(counter, 0, 16, s)); // (the partial rows are written before they are read)
    const int lanes_used = (kRmsThreads / static_cast<int>(D)) * static_cast<int>(D);
    int64_t blocks = (M * D + lanes_used * 8 - 1) / (static_cast<int64_t>(lanes_used) * 8);
    const int64_t cap = std::min<int64_t>(static_cast<int64_t>(sm_count()) * 4, kRmsMaxBlocks);
    if (blocks > cap)
        blocks = cap;
    if (blocks < 1)
        blocks = 1;
    const CommView view = comm_view(comm);
    const unsigned int seq = view.world > 1 ? ++*comm_rms_seq(comm) : 0u;
    rms_update_kernel<<<static_cast<unsigned>(blocks), kRmsThreads, sizeof(double) * 2 * lanes_used, s>>>(
        d_x, M, static_cast<int>(D), ws, counter, d_count, d_mean, d_var, view, seq, d_comm_error);
    CPPRL_LAUNCH_CHECK("rms_update_kernel");
    return 0;
}
} // namespace cpprl_b200

extern "C" int cpprl_b200_rms_update(const float *d_x, int64_t M, int64_t D, float *d_count, float *d_mean,
                                     float *d_var, void *d_workspace, void *stream)
{
    return launch_rms_update(d_x, M, D, d_count, d_mean, d_var, d_workspace, nullptr, static_cast<cudaStream_t>(stream), nullptr);
}

extern "C" int cpprl_b200_rms_update_sharded(cpprl_b200_comm *comm, const float *d_x, int64_t M, int64_t D,
                                             float *d_count, float *d_mean, float *d_var, void *d_workspace, void *stream)
{
    if (!comm)
        return fail(CPPRL_B200_EINVAL, "rms_update_sharded: null communicator");
    // a missing peer raises the communicator's error word: cpprl_b200_comm_check reports it as CPPRL_B200_ENCCL
    return launch_rms_update(d_x, M, D, d_count, d_mean, d_var, d_workspace, comm, static_cast<cudaStream_t>(stream), nullptr);
}

extern "C" int cpprl_b200_normalize_observations(const float *d_x, int64_t M, int64_t D, const float *d_mean,
                                                 const float *d_var, const float *d_clip, float *d_out, void *stream)
{
    if (!d_x || !d_mean || !d_var || !d_clip || !d_out)
        return fail(CPPRL_B200_EINVAL, "normalize_observations: null pointer");
    if (M < 0 || D <= 0)
        return fail(CPPRL_B200_EINVAL, "normalize_observations: bad size");
    if (int rc = require_device())
        return rc;
    if (M == 0)
        return 0;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const int64_t total = M * D;
    int64_t blocks = (total + 255) / 256;
    const int64_t cap = static_cast<int64_t>(sm_count()) * 16;
    if (blocks > cap)
        blocks = cap;
    normalize_obs_kernel<<<static_cast<unsigned>(blocks), 256, 0, s>>>(d_x, total, static_cast<int>(D), d_mean, d_var,
                                                                       d_clip, d_out);
    CPPRL_LAUNCH_CHECK("normalize_obs_kernel");
    return 0;
}

extern "C" int cpprl_b200_adam_step(float *d_params, const float *d_grads, float *d_exp_avg, float *d_exp_avg_sq,
                                    int64_t n, int64_t step, float lr, float beta1, float beta2, float eps, void *stream)
{
    if (!d_params || !d_grads || !d_exp_avg || !d_exp_avg_sq)
        return fail(CPPRL_B200_EINVAL, "adam_step: null pointer");
    if (n < 0 || step < 1)
        return fail(CPPRL_B200_EINVAL, "adam_step: n must be >= 0 and step >= 1");
    if (int rc = require_device())
        return rc;
    if (n == 0)
        return 0;
    // bias corrections as the host doubles libtorch computes
    const double bc1 = 1.0 - std::pow(static_cast<double>(beta1), static_cast<double>(step));
    const double bc2 = 1.0 - std::pow(static_cast<double>(beta2), static_cast<double>(step));
    const float step_size = static_cast<float>(static_cast<double>(lr) / bc1);
    const float bc2_sqrt = static_cast<float>(std::sqrt(bc2));
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    adam_step_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, s>>>(d_params, d_grads, d_exp_avg, d_exp_avg_sq, n,
                                                                            step_size, bc2_sqrt, beta1, beta2, eps);
    CPPRL_LAUNCH_CHECK("adam_step_kernel");
    return 0;
}

extern "C" int cpprl_b200_rmsprop_step(float *d_params, const float *d_grads, float *d_square_avg, int64_t n, float lr,
                                       float alpha, float eps, void *stream)
{
    if (!d_params || !d_grads || !d_square_avg)
        return fail(CPPRL_B200_EINVAL, "rmsprop_step: null pointer");
    if (n < 0)
        return fail(CPPRL_B200_EINVAL, "rmsprop_step: n must be >= 0");
    if (int rc = require_device())
        return rc;
    if (n == 0)
        return 0;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    rmsprop_step_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, s>>>(d_params, d_grads, d_square_avg, n, lr,
                                                                               alpha, eps);
    CPPRL_LAUNCH_CHECK("rmsprop_step_kernel");
    return 0;
}
