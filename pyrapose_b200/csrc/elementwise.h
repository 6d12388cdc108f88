// Bandwidth-bound helper kernels around the implicit-GEMM convolutions (internal interface).
// All activations are bf16 "padded-flat" [B, H+2, W+2, C] with a zero border unless stated otherwise.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>

namespace ppb {

// Stem input transform: fp32 NHWC image [B,H,W,3] -> bf16 space-to-depth grid S [B, Ho+3, Wo+3, 16] with
//   S[b, Y, X, (py*2+px)*3 + c] = image[b, 2*(Y-2)+py, 2*(X-2)+px, c]   (zero outside the image; channels 12..15 zero)
// so that ZeroPadding2D(3) + 7x7/2 VALID (keras-resnet conv1; PyraPose/models/resnet.py:87) becomes four row-shifted
// GEMM taps (dy = 0..3) whose A rows are the 64 contiguous values S[b, oy+dy, ox..ox+3, :] — read by TMA through a
// tensor map whose row stride (32 B) is smaller than its row extent (128 B). No im2col matrix is materialised.
cudaError_t launch_stem_s2d(const float* image, __nv_bfloat16* out, int B, int H, int W, int Ho, int Wo,
                            cudaStream_t st);
// Same from raw uint8 BGR pixels with preprocess_image(mode='caffe') fused in (PyraPose/utils/image.py:35-62):
// value = float32(pixel) - mean[c]. Input staging of SURVEY.md §8f-2: a quarter of the host->device bytes.
cudaError_t launch_stem_s2d_u8(const unsigned char* image, const float mean[3], __nv_bfloat16* out, int B, int H, int W,
                               int Ho, int Wo, cudaStream_t st);

// MaxPooling2D(3, strides 2, 'same') (keras-resnet pool1): input pixel grid [B,in_hs,in_ws,C] of which the top-left
// Hi x Wi region is valid -> padded-flat [B,Ho+2,Wo+2,C].
cudaError_t launch_maxpool3x3s2(const __nv_bfloat16* in, __nv_bfloat16* out, int B, int Hi, int Wi, int in_hs,
                                int in_ws, int C, int Ho, int Wo, cudaStream_t st);

// Parity planes for stride-2 convolutions: plane (ph,pw)[b,i,j] = in[b, 2i+ph, 2j+pw] (zero outside), each plane
// padded-flat [B,Ho+2,Wo+2,C], planes stored back to back (n_planes = 1: plain subsampling for 1x1/2 convs;
// 4: space-to-depth for 3x3/2 'same' convs).
cudaError_t launch_parity_planes(const __nv_bfloat16* in, __nv_bfloat16* out, int B, int H, int W, int C, int Ho,
                                 int Wo, int n_planes, cudaStream_t st);

// out = a + b (+ c) (+ nearest-resize(u) to a's size); keras Add + UpsampleLike
// (PyraPose/models/retinanet.py:196-211, layers/_misc.py:96-115). b, c, u may be null.
cudaError_t launch_add(const __nv_bfloat16* a, const __nv_bfloat16* b, const __nv_bfloat16* c,
                       const __nv_bfloat16* u, int uH, int uW, __nv_bfloat16* out, int B, int H, int W, int C,
                       cudaStream_t st);

}  // namespace ppb

// ---- fp32 path (Engine precision 1, ConvParams::split): every activation row holds 3*C bf16 values, the three parts
// p0 = bf16(v), p1 = bf16(v - p0), p2 = bf16(v - p0 - p1) of each channel in columns [0, C), [C, 2C), [2C, 3C). The
// copy-like kernel (parity planes) runs unchanged on 3*C channels; the arithmetic ones have split forms that compute on
// the fp32 sum of the parts and split the result again.
namespace ppb {

// Stem in full fp32 on the CUDA cores (1.45 GFLOP per image, not a tensor-core job at K = 147):
// ZeroPadding2D(3) + 7x7/2 VALID + folded BatchNorm + ReLU from the fp32 (or raw uint8 - mean) image into the stem's
// pixel grid [B, Ho+3, Wo+3, 3*64]. w = [7][7][3][64] fp32 (BatchNorm scale folded in), bias = [64].
cudaError_t launch_stem_conv_split(const void* image, int is_u8, const float mean[3], const float* w, const float* bias,
                                   __nv_bfloat16* out, int B, int H, int W, int Ho, int Wo, cudaStream_t st);
cudaError_t launch_maxpool3x3s2_split(const __nv_bfloat16* in, __nv_bfloat16* out, int B, int Hi, int Wi, int in_hs,
                                      int in_ws, int C, int Ho, int Wo, cudaStream_t st);
cudaError_t launch_add_split(const __nv_bfloat16* a, const __nv_bfloat16* b, const __nv_bfloat16* c,
                             const __nv_bfloat16* u, int uH, int uW, __nv_bfloat16* out, int B, int H, int W, int C,
                             cudaStream_t st);

}  // namespace ppb
