// Bandwidth-bound kernels of the training step: the loss functions compiled by
// PyraPose/bin/train.py:95-102 (focal + orthogonal_l1, PyraPose/losses.py:22-68, 321-408) with their gradients, the
// adjoints of the engine's elementwise ops, bias gradients, and Keras' Adam(lr, clipnorm) update.
// Internal interface between train_kernels.cu and train.cu.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace ppb {

// Geometry of a head output: `n_levels` pyramid levels stored as row segments of one padded-flat matrix.
struct HeadGeom {
  int n_levels;
  int h[3], w[3];
  int row_begin[3];  // first matrix row of the level (level planes are [batch, h+2, w+2])
  int loc_base[3];   // first location index of the level inside one image
  int locs;          // locations per image over all levels
};

// Slots of the loss accumulator (double[8]) and of the positive counters (int[4]).
enum LossSlot { kLossReg = 0, kLossCls = 1, kLossMask = 2, kLossL2 = 3, kGradSumSq = 4 };

// counts[slot] = number of rows of y_true [rows, width] whose last column equals 1 (the anchor state).
cudaError_t launch_count_positive(const float* y_true, long long rows, int width, int* counts, int slot,
                                  cudaStream_t st);

// focal(alpha, gamma) on a sigmoid head (PyraPose/losses.py:22-68).
//   prob   f32 [batch, locs*A, C]        (the model output)
//   y_true f32 [batch, locs*A, C+1]      (one-hot labels + anchor state: -1 ignore, 0 background, 1 object)
//   grad   bf16 padded-flat [rows, ld]   dLoss/dlogit at column a*C + c of the location's row (border rows untouched)
//   loss   double accumulators; the kernel adds  sum(focal) / max(1, counts[slot])  to loss[slot]
cudaError_t launch_focal_loss_grad(const float* prob, const float* y_true, __nv_bfloat16* grad, int ld, HeadGeom g,
                                   int batch, int A, int C, float alpha, float gamma, const int* counts, int slot,
                                   double* loss, cudaStream_t st);

// orthogonal_l1(weight, sigma) on the raw 16-value regression head (PyraPose/losses.py:321-408).
//   pred f32 [batch, locs*A, 16], y_true f32 [batch, locs*A, 17]; grad column a*16 + j.
cudaError_t launch_orth_l1_loss_grad(const float* pred, const float* y_true, __nv_bfloat16* grad, int ld, HeadGeom g,
                                     int batch, int A, float weight, float sigma, const int* counts, int slot,
                                     double* loss, cudaStream_t st);

// out = relu_mask( a + b + c + downsum(d) + unplanes(p) ), every operand optional except that one must be given.
//   a, b, c : bf16 padded-flat [B, H+2, W+2, C]
//   d       : bf16 padded-flat [B, dH+2, dW+2, C]; adjoint of UpsampleLike (nearest, out = H x W from d's... see .cu)
//   p       : bf16 parity planes [n_planes][B, pH+2, pW+2, C] (adjoint of launch_parity_planes)
//   mask    : bf16 padded-flat like out; result zeroed where mask <= 0
cudaError_t launch_grad_combine(const __nv_bfloat16* a, const __nv_bfloat16* b, const __nv_bfloat16* c,
                                const __nv_bfloat16* d, int dH, int dW, const __nv_bfloat16* p, int pH, int pW,
                                int n_planes, const __nv_bfloat16* mask, __nv_bfloat16* out, int B, int H, int W, int C,
                                cudaStream_t st);

// dbias[c] += sum over rows of g[row, c]   (g bf16 [rows, ld], c < cout; dbias fp32, caller zeroes it)
cudaError_t launch_bias_grad(const __nv_bfloat16* g, long long rows, int ld, int cout, float* dbias, cudaStream_t st);

// One trainable tensor of the optimizer. Kernels ("kind" 0) are stored [cout][k] with k = tap*cin + ci; the raw Keras
// gradient is grad * scale[co] (frozen BatchNorm folded into the forward weights).
struct ParamDesc {
  long long offset;    // into the flat grad / master / m / v buffers (elements)
  long long count;     // cout * k (kernel) or cout (bias)
  int kind;            // 0 kernel, 1 bias
  int cout, cin, n_taps;
  int k_total;         // columns of the forward bf16 matrix (= n_taps * cin)
  int g_ld;            // cout rounded up to 64: column block of one tap in the dgrad matrix
  const float* scale;  // [cout] BatchNorm scale or null
  float l2;            // kernel_regularizer l2 factor (loss += l2 * sum w^2), 0 if none
  __nv_bfloat16* w_fwd;  // [cout_pad][k_total]
  __nv_bfloat16* w_dgrad;  // [cin][n_taps * g_ld] or null (layer needs no data gradient)
  float* bias_fwd;       // kind 1: the forward fp32 bias
  int tap_pos[9];        // position of forward tap t inside the dgrad matrix (stride-2 layers group taps by plane)
};

// Adds the L2 regulariser: grad += 2*l2*w, loss[kLossL2] += l2 * sum w^2 (layers with l2 > 0 only), then accumulates
// sum((grad*scale)^2) over every parameter into loss[kGradSumSq]. `descs` lives in device memory.
cudaError_t launch_grad_prepare(const ParamDesc* descs, int n_descs, const int* chunk_desc, const int* chunk_index,
                                int n_chunks, float* grads, const float* master, float grad_scale, double* loss,
                                cudaStream_t st);

// keras.optimizers.Adam(lr, beta_1, beta_2, epsilon, clipnorm) update (Keras 2.3.1 optimizers.py): global-norm clip
// (norm = sqrt(loss[kGradSumSq])), m/v moments, bias-corrected step; rewrites the bf16 forward and dgrad weights.
cudaError_t launch_adam_update(const ParamDesc* descs, int n_descs, const int* chunk_desc, const int* chunk_index,
                               int n_chunks, const float* grads, float* master, float* m, float* v, float lr_t,
                               float beta1, float beta2, float eps, float clipnorm, const double* loss,
                               cudaStream_t st);

constexpr int kOptChunk = 4096;  // elements per optimizer chunk

}  // namespace ppb
