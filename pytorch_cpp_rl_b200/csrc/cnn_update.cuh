// cnn_update.cuh — argument blocks of the CnnBase update path (src/model/cnn_base.cpp:9-47).
#pragma once
#include <stdint.h>

#include "mlp_update.cuh"

namespace cpprl_b200
{
// Geometry the reference hard-codes: 84x84 frames -> conv 8x8/4 (32) -> 20x20 -> conv 4x4/2 (64) -> 9x9 ->
// conv 3x3/1 (32) -> 7x7 -> flatten 1568 -> linear (hidden) (cnn_base.cpp:13-21).
constexpr int kCnnIn = 84, kCnnO1 = 20, kCnnP1 = 400, kCnnC1 = 32, kCnnO2 = 9, kCnnP2 = 81, kCnnC2 = 64, kCnnO3 = 7,
              kCnnP3 = 49, kCnnC3 = 32, kCnnFlat = 1568, kCnnHid = 512;
constexpr int kCnnK2 = 4 * 4 * kCnnC1, kCnnK3 = 3 * 3 * kCnnC2; // 512, 576

// Offsets (floats) inside the flat parameter buffer, reference registration order:
// main.0.{weight[32,C,8,8],bias} main.2.{weight[64,32,4,4],bias} main.4.{weight[32,64,3,3],bias}
// main.7.{weight[512,1568],bias} critic_linear.{weight[1,512],bias} then the head.
struct CnnParamOffsets
{
    int w1, b1, w2, b2, w3, b3, wfc, bfc, wv, bv, slog, wo, bo, P;
};

// Offsets (floats) inside the weight-image buffer: every GEMM reads its B operand contiguous along n.
struct CnnImageLayout
{
    int w1;  // [K1 = C*64][32]      k = (c, ky, kx)           conv1 forward (obs are NCHW)
    int w2;  // [512][64]            k = (ky, kx, c)           conv2 forward (activations are NHWC)
    int w3;  // [576][32]            k = (ky, kx, c)           conv3 forward
    int wfc; // [1568][512]          k = (y, x, c)             linear forward
    int wfd; // [512][1568]                                    linear data gradient
    int w3d; // [288 = (tap, n)][64]                           conv3 data gradient
    int w2d; // [4 parity classes][256 = (a, b, n)][32]        conv2 data gradient (stride 2: taps by parity)
    int total;
    int grad_total; // the gradient image only has the first four blocks
};

struct CnnWorkspace
{
    float *img, *gimg, *gflat; // weight images, gradient images (w1, w2, w3, wfc), flat gradients (biases, heads)
    float *act1, *act2, *act3, *h, *dz1, *dz2, *dz3, *dh;
    float *sums;  // [8] minibatch partial sums: value, action, entropy, kl, clip
    float *consts; // [4] optimiser step constants written by the decide kernel
    int chunk;    // rows per pass
};

struct CnnUpdateArgs
{
    float *params;
    CnnParamOffsets po;
    CnnImageLayout il;
    int C, A, head;
    const float *obs;      // [rows(+N), C, 84, 84] raw pixels 0..255 (divided by 255 in the load, cnn_base.cpp:37)
    // non-zero: cnn_scan_observations found every observation of this update on the TF32 grid (raw 0..255 pixels
    // are); conv1's GEMMs then run the instantiation without the A_lo x B_hi product (cnn_gemm_tc.cuh)
    int obs_exact;
    const void *actions;
    const float *value_preds, *returns, *old_logp;
    long long rows;
    const long long *perm; // row ids of THIS minibatch (device), nullable = identity from row_base
    long long row_base;
    int B;                 // rows of the minibatch (held by this rank)
    const double *adv_sums;
    int algo;
    float clip_param, clip_lo, clip_hi, c_actor, c_value, c_entropy;
    double kl_limit;
    int opt;
    float *opt_m, *opt_v;
    double lr;
    float beta1, beta2, eps, alpha;
    UpdateState *state;
    float *out_value, *out_logp, *out_entropy_sum; // forward evaluation
    float *out_dist;       // nullable [B][A]: normalised log-probabilities / means (see mlp_update.cuh)
    // env-sharded update (world > 1): B rows of this rank out of a minibatch of B_norm rows box-wide; gradients and
    // partial sums are added over the ranks (exchange sequence number xseq) before the optimiser step
    int B_norm;
    CommView comm;
    unsigned int xseq;
};

CnnParamOffsets cnn_param_offsets(int C, int A, int head);
CnnImageLayout cnn_image_layout(int C);
size_t cnn_workspace_floats(int C, int A, int chunk, bool training);
void cnn_workspace_carve(CnnWorkspace &ws, float *base, int C, int A, int chunk, bool training);
// *exact = 1 if none of the n floats at obs has bits below the TF32 mantissa (flag: one int of device scratch; the
// call synchronises the stream once to bring the answer to the host: the kernel variant is a host-side choice)
int cnn_scan_observations(const float *obs, long long n, int *flag, int *exact, cudaStream_t s);
inline int *cnn_inexact_flag(const CnnWorkspace &ws) { return reinterpret_cast<int *>(ws.consts + 8); } // padding of consts
int cnn_build_images(const CnnUpdateArgs &a, const CnnWorkspace &ws, cudaStream_t s);
int cnn_minibatch(const CnnUpdateArgs &a, const CnnWorkspace &ws, cudaStream_t s); // forward + backward + decide + optimiser
int cnn_forward(const CnnUpdateArgs &a, const CnnWorkspace &ws, cudaStream_t s);   // evaluate_actions
int cnn_adv_exchange(const CommView &comm, unsigned int seq, double *out, UpdateState *st, cudaStream_t s);
long long cnn_exchange_floats(int C, int A);
int cnn_debug_tc_gemm(int variant, const float *A, const float *B, float *C, int M, int N, int K, cudaStream_t s);
} // namespace cpprl_b200
