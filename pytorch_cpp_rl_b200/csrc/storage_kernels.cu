// storage_kernels.cu — the write side of the rollout buffer: RolloutStorage::insert,
// after_update, set_first_observation and the environment-axis concatenation of per-worker storages, each as ONE
// launch of a fused multi-segment copy instead of one copy kernel per field (7 for insert, 3 for after_update,
// 8 per part for the concatenation).  Pure byte movement: HBM-bound for image observations
// (cfg5 insert: 256 x 28 224 x 4 B = 28.9 MB read + written), launch-latency-bound for the MLP configs.
#include "common.cuh"

namespace cpprl_b200
{
namespace
{
constexpr int kMaxSegments = 8;
constexpr int kCopyThreads = 256;
constexpr int kCopyUnroll = 4;

// rows x row_bytes, both sides pitched (a single-row segment is a flat copy)
struct CopySegment
{
    const char *src;
    char *dst;
    int64_t row_bytes, rows, src_pitch, dst_pitch;
    int vec16; // every address, pitch and row length is a multiple of 16 bytes
};

struct CopyBatch
{
    CopySegment seg[kMaxSegments];
    int block_begin[kMaxSegments + 1]; // segment i owns blocks [block_begin[i], block_begin[i + 1])
    int count;
};

template <typename WordT>
__device__ __forceinline__ void copy_segment(const CopySegment &sg, int64_t first, int64_t stride)
{
    const int64_t words_per_row = sg.row_bytes / static_cast<int64_t>(sizeof(WordT));
    const int64_t total = words_per_row * sg.rows;
    // kCopyUnroll independent loads in flight per thread before the first store
    for (int64_t w = first; w < total; w += stride * kCopyUnroll)
    {
        WordT v[kCopyUnroll];
        int64_t off[kCopyUnroll];
#pragma unroll
        for (int u = 0; u < kCopyUnroll; ++u)
        {
            const int64_t wu = w + u * stride;
            if (wu < total)
            {
                const int64_t r = sg.rows == 1 ? 0 : wu / words_per_row;
                const int64_t c = (wu - r * words_per_row) * static_cast<int64_t>(sizeof(WordT));
                off[u] = r * sg.dst_pitch + c;
                v[u] = __ldcs(reinterpret_cast<const WordT *>(sg.src + r * sg.src_pitch + c));
            }
        }
#pragma unroll
        for (int u = 0; u < kCopyUnroll; ++u)
            if (w + u * stride < total)
                __stcs(reinterpret_cast<WordT *>(sg.dst + off[u]), v[u]);
    }
}

// Every segment owns a contiguous range of blocks sized to its length (a [N] column gets one block, the frames
// get the rest of the machine), so no block is launched only to leave.
__global__ void __launch_bounds__(kCopyThreads) storage_copy_kernel(const __grid_constant__ CopyBatch batch)
{
    int k = 0;
    while (k + 1 < batch.count && static_cast<int>(blockIdx.x) >= batch.block_begin[k + 1])
        ++k;
    const CopySegment &sg = batch.seg[k];
    const int64_t first = static_cast<int64_t>(blockIdx.x - batch.block_begin[k]) * kCopyThreads + threadIdx.x;
    const int64_t stride = static_cast<int64_t>(batch.block_begin[k + 1] - batch.block_begin[k]) * kCopyThreads;
    if (sg.vec16)
        copy_segment<uint4>(sg, first, stride);
    else
        copy_segment<uint32_t>(sg, first, stride);
}

bool aligned16(const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

void add_segment(CopyBatch &b, const void *src, void *dst, int64_t row_bytes, int64_t rows, int64_t src_pitch,
                 int64_t dst_pitch)
{
    if (row_bytes <= 0 || rows <= 0)
        return;
    if (src_pitch == row_bytes && dst_pitch == row_bytes)
    {
        // dense on both sides: one flat row
        row_bytes *= rows;
        rows = 1;
        src_pitch = dst_pitch = row_bytes;
    }
    CopySegment &sg = b.seg[b.count++];
    sg.src = static_cast<const char *>(src);
    sg.dst = static_cast<char *>(dst);
    sg.row_bytes = row_bytes;
    sg.rows = rows;
    sg.src_pitch = src_pitch;
    sg.dst_pitch = dst_pitch;
    sg.vec16 = aligned16(src) && aligned16(dst) && row_bytes % 16 == 0 && (rows == 1 || (src_pitch % 16 == 0 && dst_pitch % 16 == 0));
}

int launch_copy(CopyBatch &b, const char *family, cudaStream_t s)
{
    if (b.count == 0)
        return 0;
    const int64_t per_block = static_cast<int64_t>(kCopyThreads) * kCopyUnroll;
    const int64_t cap = static_cast<int64_t>(sm_count()) * 8; // 8 CTAs of 256 threads = a full SM
    b.block_begin[0] = 0;
    for (int i = 0; i < b.count; ++i)
    {
        const int64_t words = b.seg[i].row_bytes * b.seg[i].rows / (b.seg[i].vec16 ? 16 : 4);
        int64_t blocks = (words + per_block - 1) / per_block;
        blocks = blocks > cap ? cap : (blocks < 1 ? 1 : blocks);
        b.block_begin[i + 1] = b.block_begin[i] + static_cast<int>(blocks);
    }
    KernelTimer timer(family, s);
    storage_copy_kernel<<<static_cast<unsigned>(b.block_begin[b.count]), kCopyThreads, 0, s>>>(b);
    CPPRL_LAUNCH_CHECK("storage_copy_kernel");
    return 0;
}

int check_storage(const cpprl_b200_storage *st, const char *who)
{
    if (!st)
        return fail(CPPRL_B200_EINVAL, std::string(who) + ": null storage description");
    if (st->T <= 0 || st->N <= 0 || st->obs_dim <= 0 || st->hidden_dim <= 0 || st->action_cols <= 0)
        return fail(CPPRL_B200_EINVAL, std::string(who) + ": non-positive size");
    if (st->action_elem_bytes != 4 && st->action_elem_bytes != 8)
        return fail(CPPRL_B200_EINVAL, std::string(who) + ": action_elem_bytes must be 4 (fp32) or 8 (int64)");
    if (!st->d_observations || !st->d_hidden_states || !st->d_rewards || !st->d_value_preds || !st->d_returns ||
        !st->d_action_log_probs || !st->d_actions || !st->d_masks)
        return fail(CPPRL_B200_EINVAL, std::string(who) + ": null rollout buffer");
    return require_device();
}
} // namespace
} // namespace cpprl_b200

using namespace cpprl_b200;

extern "C" int cpprl_b200_storage_insert(const cpprl_b200_storage *st, int64_t step, const float *d_observation,
                                         const float *d_hidden_state, const void *d_action,
                                         const float *d_action_log_prob, const float *d_value_prediction,
                                         const float *d_reward, const float *d_mask, void *stream)
{
    if (int rc = check_storage(st, "storage_insert"))
        return rc;
    if (step < 0 || step >= st->T)
        return fail(CPPRL_B200_EINVAL, "storage_insert: step outside [0, T)");
    if (!d_observation || !d_hidden_state || !d_action || !d_action_log_prob || !d_value_prediction || !d_reward || !d_mask)
        return fail(CPPRL_B200_EINVAL, "storage_insert: null input");
    const int64_t N = st->N, f = sizeof(float);
    const int64_t obs_row = N * st->obs_dim * f, hid_row = N * st->hidden_dim * f;
    const int64_t act_row = N * st->action_cols * st->action_elem_bytes, one_row = N * f;
    CopyBatch b{};
    // storage.cpp:183-191: observation, hidden state and mask land on row step+1, the rest on row step
    add_segment(b, d_observation, reinterpret_cast<char *>(st->d_observations) + (step + 1) * obs_row, obs_row, 1, obs_row, obs_row);
    add_segment(b, d_hidden_state, reinterpret_cast<char *>(st->d_hidden_states) + (step + 1) * hid_row, hid_row, 1, hid_row, hid_row);
    add_segment(b, d_action, static_cast<char *>(st->d_actions) + step * act_row, act_row, 1, act_row, act_row);
    add_segment(b, d_action_log_prob, reinterpret_cast<char *>(st->d_action_log_probs) + step * one_row, one_row, 1, one_row, one_row);
    add_segment(b, d_value_prediction, reinterpret_cast<char *>(st->d_value_preds) + step * one_row, one_row, 1, one_row, one_row);
    add_segment(b, d_reward, reinterpret_cast<char *>(st->d_rewards) + step * one_row, one_row, 1, one_row, one_row);
    add_segment(b, d_mask, reinterpret_cast<char *>(st->d_masks) + (step + 1) * one_row, one_row, 1, one_row, one_row);
    return launch_copy(b, "storage", static_cast<cudaStream_t>(stream));
}

extern "C" int cpprl_b200_storage_after_update(const cpprl_b200_storage *st, void *stream)
{
    if (int rc = check_storage(st, "storage_after_update"))
        return rc;
    const int64_t N = st->N, f = sizeof(float), T = st->T;
    const int64_t obs_row = N * st->obs_dim * f, hid_row = N * st->hidden_dim * f, one_row = N * f;
    CopyBatch b{};
    // storage.cpp:106-111: row 0 <- row T of observations, hidden states and masks
    add_segment(b, reinterpret_cast<char *>(st->d_observations) + T * obs_row, st->d_observations, obs_row, 1, obs_row, obs_row);
    add_segment(b, reinterpret_cast<char *>(st->d_hidden_states) + T * hid_row, st->d_hidden_states, hid_row, 1, hid_row, hid_row);
    add_segment(b, reinterpret_cast<char *>(st->d_masks) + T * one_row, st->d_masks, one_row, 1, one_row, one_row);
    return launch_copy(b, "storage", static_cast<cudaStream_t>(stream));
}

extern "C" int cpprl_b200_storage_set_first_observation(const cpprl_b200_storage *st, const float *d_observation,
                                                        void *stream)
{
    if (int rc = check_storage(st, "storage_set_first_observation"))
        return rc;
    if (!d_observation)
        return fail(CPPRL_B200_EINVAL, "storage_set_first_observation: null input");
    const int64_t obs_row = st->N * st->obs_dim * static_cast<int64_t>(sizeof(float));
    CopyBatch b{};
    add_segment(b, d_observation, st->d_observations, obs_row, 1, obs_row, obs_row);
    return launch_copy(b, "storage", static_cast<cudaStream_t>(stream));
}

extern "C" int cpprl_b200_storage_concat(const cpprl_b200_storage *dst, const cpprl_b200_storage *part,
                                         int64_t env_offset, void *stream)
{
    if (int rc = check_storage(dst, "storage_concat (destination)"))
        return rc;
    if (int rc = check_storage(part, "storage_concat (part)"))
        return rc;
    if (part->T != dst->T || part->obs_dim != dst->obs_dim || part->hidden_dim != dst->hidden_dim ||
        part->action_cols != dst->action_cols || part->action_elem_bytes != dst->action_elem_bytes)
        return fail(CPPRL_B200_EINVAL, "storage_concat: part and destination differ in a dimension other than the environment count");
    if (env_offset < 0 || env_offset + part->N > dst->N)
        return fail(CPPRL_B200_EINVAL, "storage_concat: environments [env_offset, env_offset + part.N) fall outside the destination");
    const int64_t T = dst->T, f = sizeof(float);
    CopyBatch b{};
    // torch::cat(parts, 1) of every field (storage.cpp:51-104): time row t of the part becomes columns
    // [env_offset, env_offset + part.N) of time row t of the destination
    auto field = [&](const void *src, void *dstp, int64_t cols_bytes, int64_t rows) {
        add_segment(b, src, static_cast<char *>(dstp) + env_offset * cols_bytes, part->N * cols_bytes, rows,
                    part->N * cols_bytes, dst->N * cols_bytes);
    };
    field(part->d_observations, dst->d_observations, dst->obs_dim * f, T + 1);
    field(part->d_hidden_states, dst->d_hidden_states, dst->hidden_dim * f, T + 1);
    field(part->d_rewards, dst->d_rewards, f, T);
    field(part->d_value_preds, dst->d_value_preds, f, T + 1);
    field(part->d_returns, dst->d_returns, f, T + 1);
    field(part->d_action_log_probs, dst->d_action_log_probs, f, T);
    field(part->d_actions, dst->d_actions, dst->action_cols * dst->action_elem_bytes, T);
    field(part->d_masks, dst->d_masks, f, T + 1);
    return launch_copy(b, "storage", static_cast<cudaStream_t>(stream));
}
