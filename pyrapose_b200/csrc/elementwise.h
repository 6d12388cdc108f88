// Bandwidth-bound helper kernels around the implicit-GEMM convolutions (internal interface).
// All activations are bf16 "padded-flat" [B, H+2, W+2, C] with a zero border unless stated otherwise.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>

namespace ppb {

// Stem im2col: fp32 NHWC image [B,H,W,3] -> bf16 [B*Ho*Wo, 192] for ZeroPadding2D(3) + 7x7/2 VALID
// (keras-resnet conv1; PyraPose/models/resnet.py:87). K index = (r*7+s)*3 + c, columns 147..191 zero.
cudaError_t launch_stem_im2col(const float* image, __nv_bfloat16* out, int B, int H, int W, int Ho, int Wo,
                               cudaStream_t st);

// MaxPooling2D(3, strides 2, 'same') (keras-resnet pool1): plain [B,Hi,Wi,C] -> padded-flat [B,Ho+2,Wo+2,C].
cudaError_t launch_maxpool3x3s2(const __nv_bfloat16* in, __nv_bfloat16* out, int B, int Hi, int Wi, int C, int Ho,
                                int Wo, cudaStream_t st);

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
