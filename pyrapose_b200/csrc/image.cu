// Input staging right before the hot path: what the reference's generator and evaluators do
// on the host for every image before predict_on_batch —
//   preprocess_image(mode='caffe')   PyraPose/utils/image.py:35-62    float32, BGR mean (103.939, 116.779, 123.68) removed
//   resize_image(min 480 / max 640)  PyraPose/utils/image.py:281-323  cv2.resize(img, None, fx=scale, fy=scale), bilinear,
//                                                                     on the float32 mean-subtracted image
//   compute_inputs                   PyraPose/preprocessing/generator.py:320-336  copy to the upper-left corner of a zero batch
// — as ONE kernel per image that reads the raw uint8 (or float32) pixels and writes the image's slot of the batch
// tensor the engine is bound to: resized, mean-subtracted, zero-padded.
//
// Arithmetic = OpenCV's float32 INTER_LINEAR path (resize.cpp: hresize then vresize, coefficients computed in float
// from the double scale): fx = (dx + 0.5) * scale_x - 0.5 in double; sx = floor(fx); a1 = float(fx - sx) (a1 = 0 and the edge
// pixel when sx falls outside), row value = S[sx] * a0 + S[sx + 1] * a1, pixel = row0 * b0 + row1 * b1. Compiled with
// -fmad=false, every product and sum rounded separately; OpenCV's SIMD kernels may fuse them, so parity with
// cv2.resize is stated as <= 2 ulp of the result (tests/test_image_gpu.py), not bit-exact.
// Bandwidth-bound: one thread per output pixel (3 channels), 4 x 3 source bytes read, 12 bytes written, consecutive
// threads on consecutive pixels of a row.
#include "../../include/pyrapose_b200.h"
#include "errors.h"
#include "launch_count.h"

namespace ppb {
namespace {

template <typename T>
__global__ void __launch_bounds__(256)
resize_preprocess_kernel(const T* __restrict__ src, int sh, int sw, float* __restrict__ dst, int dh, int dw, int pad_h,
                         int pad_w, double scale_x, double scale_y, float m0, float m1, float m2, int identity) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  const int y = blockIdx.y;
  if (x >= pad_w || y >= pad_h) return;
  float* o = dst + (static_cast<size_t>(y) * pad_w + x) * 3;
  if (x >= dw || y >= dh) {  // compute_inputs: the batch tensor is zero outside the image
    o[0] = 0.0f;
    o[1] = 0.0f;
    o[2] = 0.0f;
    return;
  }
  const float mean[3] = {m0, m1, m2};
  if (identity) {  // scale == 1: cv2.resize returns the pixels unchanged
    const T* s = src + (static_cast<size_t>(y) * sw + x) * 3;
#pragma unroll
    for (int c = 0; c < 3; ++c) o[c] = static_cast<float>(s[c]) - mean[c];
    return;
  }
  // OpenCV 4.13 keeps the source coordinate in double until the integer part is removed (a float32 coordinate near
  // 1240 would quantise the weight to 1.2e-4); measured against cv2.resize with an alternating 0/1 image.
  const double fxd = (x + 0.5) * scale_x - 0.5;
  int sx = static_cast<int>(floor(fxd));
  float fx = static_cast<float>(fxd - sx);
  if (sx < 0) {
    fx = 0.0f;
    sx = 0;
  }
  if (sx >= sw - 1) {
    fx = 0.0f;
    sx = sw - 1;
  }
  const double fyd = (y + 0.5) * scale_y - 0.5;
  int sy = static_cast<int>(floor(fyd));
  float fy = static_cast<float>(fyd - sy);
  if (sy < 0) {
    fy = 0.0f;
    sy = 0;
  }
  if (sy >= sh - 1) {
    fy = 0.0f;
    sy = sh - 1;
  }
  const int sx1 = min(sx + 1, sw - 1), sy1 = min(sy + 1, sh - 1);
  const float a0 = 1.0f - fx, a1 = fx, b0 = 1.0f - fy, b1 = fy;
  const T* r0 = src + static_cast<size_t>(sy) * sw * 3;
  const T* r1 = src + static_cast<size_t>(sy1) * sw * 3;
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const float p00 = static_cast<float>(r0[sx * 3 + c]) - mean[c], p01 = static_cast<float>(r0[sx1 * 3 + c]) - mean[c];
    const float p10 = static_cast<float>(r1[sx * 3 + c]) - mean[c], p11 = static_cast<float>(r1[sx1 * 3 + c]) - mean[c];
    const float h0 = p00 * a0 + p01 * a1;
    const float h1 = p10 * a0 + p11 * a1;
    o[c] = h0 * b0 + h1 * b1;
  }
}

}  // namespace
}  // namespace ppb

using namespace ppb;

extern "C" int ppb_resize_preprocess_f32(const void* src, int src_is_u8, int src_h, int src_w, float* dst, int dst_h,
                                         int dst_w, int pad_h, int pad_w, double fx, double fy,
                                         const float* mean3_host, void* stream) {
  PPB_REQUIRE(src_h > 0 && src_w > 0 && dst_h > 0 && dst_w > 0 && pad_h >= dst_h && pad_w >= dst_w,
              "ppb_resize_preprocess_f32: bad shape src %dx%d dst %dx%d pad %dx%d", src_h, src_w, dst_h, dst_w, pad_h,
              pad_w);
  PPB_REQUIRE(src != nullptr && dst != nullptr, "ppb_resize_preprocess_f32: null buffer");
  PPB_REQUIRE(fx > 0.0 && fy > 0.0, "ppb_resize_preprocess_f32: scale factors must be positive");
  const float m0 = mean3_host ? mean3_host[0] : 0.0f, m1 = mean3_host ? mean3_host[1] : 0.0f,
              m2 = mean3_host ? mean3_host[2] : 0.0f;
  const int identity = (dst_h == src_h && dst_w == src_w && fx == 1.0 && fy == 1.0) ? 1 : 0;
  const dim3 grid((pad_w + 255) / 256, pad_h);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  // cv::resize: scale_x = 1 / inv_scale_x with inv_scale_x = fx as passed (dsize is derived from it, not the reverse)
  if (src_is_u8)
    resize_preprocess_kernel<unsigned char><<<grid, 256, 0, st>>>(static_cast<const unsigned char*>(src), src_h, src_w,
                                                                  dst, dst_h, dst_w, pad_h, pad_w, 1.0 / fx, 1.0 / fy,
                                                                  m0, m1, m2, identity);
  else
    resize_preprocess_kernel<float><<<grid, 256, 0, st>>>(static_cast<const float*>(src), src_h, src_w, dst, dst_h,
                                                          dst_w, pad_h, pad_w, 1.0 / fx, 1.0 / fy, m0, m1, m2, identity);
  PPB_CUDA(cudaGetLastError());
  count_launch();
  return kOk;
}
