// Per-anchor decode kernels (stand-alone layer API). The engine fuses the same arithmetic into the
// regression head's epilogue (conv_igemm.cu); these entry points serve RegressBoxes3D / RegressBoxes /
// DenormBoxes3D / ClipBoxes when they are called as layers on caller-supplied tensors.
//
// Reference: PyraPose/backend/common.py:6-22 (bbox_transform_inv), :25-56 (box3D_transform_inv),
// :59-90 (box3D_denormalization); PyraPose/layers/_misc.py:259-280 (ClipBoxes).
// Bandwidth-bound: 16 B of anchor + 64 B of deltas read, 64 B written per anchor. Four threads share
// one anchor, each owning one float4 (x1-,y1-,x2-,y2-based values), so every global access is a fully
// coalesced 128-bit transaction. Compiled with -fmad=false (TF evaluates mul and add as separate ops).
#include "../../include/pyrapose_b200.h"
#include "errors.h"
#include "launch_count.h"

namespace ppb {
namespace {

struct Vec16 {
  float v[16];
};

// groups = values per anchor / 4 (4 for the 16 corner values, 1 for plain boxes).
// grid (ceil(anchors_per_image * GROUPS / (256 * kDecodeUnroll)), batch): the image index comes from blockIdx.y, so no
// 64-bit division or modulo is needed to find the (shared) anchor of a thread. Every thread owns kDecodeUnroll vectors
// 256 apart and issues all of its loads before the first use: 2 x kDecodeUnroll x 16 B in flight per thread is what it
// takes to cover the HBM latency at full occupancy.
constexpr int kDecodeUnroll = 4;

template <int GROUPS, bool DENORM>
__global__ void __launch_bounds__(256) transform_inv_kernel(const float4* __restrict__ anchors,
                                                            const float4* __restrict__ deltas,
                                                            float4* __restrict__ out, float4* __restrict__ hull,
                                                            unsigned anchors_per_image, int anchors_batched, Vec16 mean,
                                                            Vec16 stdv) {
  const unsigned t0 = blockIdx.x * (256u * kDecodeUnroll) + threadIdx.x;  // first vector index inside the image
  const int g = static_cast<int>(t0 % GROUPS);                             // 256 % GROUPS == 0: the same for every u
  const size_t img = static_cast<size_t>(blockIdx.y) * anchors_per_image;
  float4 box[kDecodeUnroll], d[kDecodeUnroll];
#pragma unroll
  for (int u = 0; u < kDecodeUnroll; ++u) {
    const unsigned t = t0 + u * 256u;
    const unsigned a = t / GROUPS;
    const bool valid = a < anchors_per_image;  // no early return: the hull reduction below shuffles across lanes
    box[u] = valid ? __ldg(anchors + (anchors_batched ? img + a : a)) : make_float4(0.f, 0.f, 0.f, 0.f);
    d[u] = valid ? __ldg(deltas + img * GROUPS + t) : make_float4(0.f, 0.f, 0.f, 0.f);
  }
  // pick this thread's four (mean, std) pairs with selects: indexing a by-value parameter array with a runtime
  // index would force the whole array into local memory
  float m[4], sd[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    m[j] = mean.v[j];
    sd[j] = stdv.v[j];
#pragma unroll
    for (int gg = 1; gg < GROUPS; ++gg) {
      m[j] = g == gg ? mean.v[gg * 4 + j] : m[j];
      sd[j] = g == gg ? stdv.v[gg * 4 + j] : sd[j];
    }
  }
#pragma unroll
  for (int u = 0; u < kDecodeUnroll; ++u) {
    const unsigned t = t0 + u * 256u;
    const unsigned a = t / GROUPS;
    const bool valid = a < anchors_per_image;
    const float w = box[u].z - box[u].x;   // common.py:33
    const float h = box[u].w - box[u].y;   // common.py:34
    float4 r;
    r.x = box[u].x + (d[u].x * sd[0] + m[0]) * w;
    r.y = box[u].y + (d[u].y * sd[1] + m[1]) * h;
    r.z = box[u].z + (d[u].z * sd[2] + m[2]) * w;
    r.w = box[u].w + (d[u].w * sd[3] + m[3]) * h;
    if (DENORM) {  // common.py:69-84
      r.x = r.x / 320.0f;
      r.y = r.y / 240.0f;
      r.z = r.z / 320.0f;
      r.w = r.w / 240.0f;
    }
    if (valid) out[img * GROUPS + t] = r;
    if (GROUPS == 4 && hull != nullptr) {
      // axis-aligned hull of the 8 corners: min/max over the four threads of this anchor
      float xmin = fminf(r.x, r.z), xmax = fmaxf(r.x, r.z), ymin = fminf(r.y, r.w), ymax = fmaxf(r.y, r.w);
#pragma unroll
      for (int off = 1; off < 4; off <<= 1) {
        xmin = fminf(xmin, __shfl_xor_sync(0xffffffffu, xmin, off));
        xmax = fmaxf(xmax, __shfl_xor_sync(0xffffffffu, xmax, off));
        ymin = fminf(ymin, __shfl_xor_sync(0xffffffffu, ymin, off));
        ymax = fmaxf(ymax, __shfl_xor_sync(0xffffffffu, ymax, off));
      }
      if (valid && g == 0) hull[img + a] = make_float4(xmin, ymin, xmax, ymax);
    }
  }
}

__global__ void __launch_bounds__(256) clip_boxes_kernel(const float4* __restrict__ boxes, float4* __restrict__ out,
                                                         long long total, float width, float height) {
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= total) return;
  float4 b = __ldg(boxes + i);
  b.x = fminf(fmaxf(b.x, 0.0f), width);
  b.y = fminf(fmaxf(b.y, 0.0f), height);
  b.z = fminf(fmaxf(b.z, 0.0f), width);
  b.w = fminf(fmaxf(b.w, 0.0f), height);
  out[i] = b;
}

Vec16 load16(const float* p, float fill, int n) {
  Vec16 v;
  for (int i = 0; i < 16; ++i) v.v[i] = (p != nullptr && i < n) ? p[i] : fill;
  return v;
}

// UpsampleLike (PyraPose/layers/_misc.py:96-115 -> tf.image.resize NEAREST, half-pixel centres): fp32 NHWC,
// out[b, y, x, :] = in[b, min(floor((y + 0.5) * ih / oh), ih - 1), min(floor((x + 0.5) * iw / ow), iw - 1), :].
// One thread per float4 of the output (C % 4 == 0) or per float; consecutive threads on consecutive channels.
template <typename V>
__global__ void __launch_bounds__(256) upsample_like_kernel(const V* __restrict__ in, V* __restrict__ out, int B, int ih,
                                                            int iw, int oh, int ow, int cv) {
  const long long t = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const long long total = static_cast<long long>(B) * oh * ow * cv;
  if (t >= total) return;
  const int c = static_cast<int>(t % cv);
  const long long pix = t / cv;
  const int x = static_cast<int>(pix % ow), y = static_cast<int>((pix / ow) % oh), b = static_cast<int>(pix / (static_cast<long long>(ow) * oh));
  int sy = static_cast<int>(floorf((static_cast<float>(y) + 0.5f) * (static_cast<float>(ih) / oh)));
  int sx = static_cast<int>(floorf((static_cast<float>(x) + 0.5f) * (static_cast<float>(iw) / ow)));
  sy = min(sy, ih - 1);
  sx = min(sx, iw - 1);
  out[t] = __ldg(in + ((static_cast<long long>(b) * ih + sy) * iw + sx) * cv + c);
}

}  // namespace
}  // namespace ppb

using namespace ppb;

extern "C" {

int ppb_box3d_transform_inv_f32(const float* anchors, const float* deltas, float* out, float* hull, int batch,
                                long long anchors_per_image, int anchors_batched, const float* mean_host,
                                const float* std_host, int denormalize, void* stream) {
  PPB_REQUIRE(batch >= 0 && anchors_per_image >= 0, "ppb_box3d_transform_inv_f32: bad sizes");
  const long long total = static_cast<long long>(batch) * anchors_per_image;
  if (total == 0) return kOk;
  PPB_REQUIRE(anchors != nullptr && deltas != nullptr && out != nullptr, "ppb_box3d_transform_inv_f32: null buffer");
  const Vec16 mean = load16(mean_host, 0.0f, 16), stdv = load16(std_host, 0.2f, 16);
  PPB_REQUIRE(anchors_per_image < (1ll << 29) && batch <= 65535, "ppb_box3d_transform_inv_f32: sizes out of range");
  const dim3 grid(static_cast<unsigned>((anchors_per_image * 4 + 256 * kDecodeUnroll - 1) / (256 * kDecodeUnroll)),
                  static_cast<unsigned>(batch));
  const unsigned api = static_cast<unsigned>(anchors_per_image);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (denormalize)
    transform_inv_kernel<4, true><<<grid, 256, 0, st>>>(
        reinterpret_cast<const float4*>(anchors), reinterpret_cast<const float4*>(deltas),
        reinterpret_cast<float4*>(out), reinterpret_cast<float4*>(hull), api, anchors_batched, mean, stdv);
  else
    transform_inv_kernel<4, false><<<grid, 256, 0, st>>>(
        reinterpret_cast<const float4*>(anchors), reinterpret_cast<const float4*>(deltas),
        reinterpret_cast<float4*>(out), reinterpret_cast<float4*>(hull), api, anchors_batched, mean, stdv);
  PPB_CUDA(cudaGetLastError());
  count_launch();
  return kOk;
}

int ppb_bbox_transform_inv_f32(const float* anchors, const float* deltas, float* out, int batch,
                               long long anchors_per_image, int anchors_batched, const float* mean_host,
                               const float* std_host, void* stream) {
  PPB_REQUIRE(batch >= 0 && anchors_per_image >= 0, "ppb_bbox_transform_inv_f32: bad sizes");
  const long long total = static_cast<long long>(batch) * anchors_per_image;
  if (total == 0) return kOk;
  PPB_REQUIRE(anchors != nullptr && deltas != nullptr && out != nullptr, "ppb_bbox_transform_inv_f32: null buffer");
  const Vec16 mean = load16(mean_host, 0.0f, 4), stdv = load16(std_host, 0.2f, 4);
  PPB_REQUIRE(anchors_per_image < (1ll << 31) && batch <= 65535, "ppb_bbox_transform_inv_f32: sizes out of range");
  const dim3 grid(static_cast<unsigned>((anchors_per_image + 256 * kDecodeUnroll - 1) / (256 * kDecodeUnroll)),
                  static_cast<unsigned>(batch));
  transform_inv_kernel<1, false><<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const float4*>(anchors), reinterpret_cast<const float4*>(deltas),
      reinterpret_cast<float4*>(out), nullptr, static_cast<unsigned>(anchors_per_image), anchors_batched, mean, stdv);
  PPB_CUDA(cudaGetLastError());
  count_launch();
  return kOk;
}

int ppb_clip_boxes_f32(const float* boxes, float* out, long long total_boxes, float width, float height,
                       void* stream) {
  PPB_REQUIRE(total_boxes >= 0, "ppb_clip_boxes_f32: bad size");
  if (total_boxes == 0) return kOk;
  PPB_REQUIRE(boxes != nullptr && out != nullptr, "ppb_clip_boxes_f32: null buffer");
  clip_boxes_kernel<<<static_cast<unsigned>((total_boxes + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const float4*>(boxes), reinterpret_cast<float4*>(out), total_boxes, width, height);
  PPB_CUDA(cudaGetLastError());
  count_launch();
  return kOk;
}

int ppb_upsample_like_f32(const float* source, float* out, int batch, int src_h, int src_w, int dst_h, int dst_w,
                          int channels, void* stream) {
  PPB_REQUIRE(batch >= 0 && src_h > 0 && src_w > 0 && dst_h > 0 && dst_w > 0 && channels > 0,
              "ppb_upsample_like_f32: bad shape");
  if (batch == 0) return kOk;
  PPB_REQUIRE(source != nullptr && out != nullptr, "ppb_upsample_like_f32: null buffer");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const bool v4 = channels % 4 == 0 && (reinterpret_cast<uintptr_t>(source) & 15) == 0 && (reinterpret_cast<uintptr_t>(out) & 15) == 0;
  const int cv = v4 ? channels / 4 : channels;
  const long long total = static_cast<long long>(batch) * dst_h * dst_w * cv;
  const unsigned blocks = static_cast<unsigned>((total + 255) / 256);
  if (v4)
    upsample_like_kernel<float4><<<blocks, 256, 0, st>>>(reinterpret_cast<const float4*>(source), reinterpret_cast<float4*>(out),
                                                         batch, src_h, src_w, dst_h, dst_w, cv);
  else
    upsample_like_kernel<float><<<blocks, 256, 0, st>>>(source, out, batch, src_h, src_w, dst_h, dst_w, cv);
  PPB_CUDA(cudaGetLastError());
  count_launch();
  return kOk;
}

}  // extern "C"
