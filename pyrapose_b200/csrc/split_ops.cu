// Elementwise kernels of the fp32 path (BASELINE.json north_star "1e-4 fp32 path"): the engine's
// activations are three bf16 parts per value (ConvParams::split), so Add / UpsampleLike / MaxPool must sum or compare
// the fp32 values and split the result again, and the 3-channel stem runs as a plain fp32 convolution. This path exists for
// parity, not for speed: one thread per output value, consecutive threads on consecutive channels.
#include "elementwise.h"

#include <stdint.h>

namespace ppb {
namespace {

__device__ __forceinline__ float ldf(const __nv_bfloat16* p) { return __bfloat162float(*p); }

constexpr int kParts = 3;  // engine.h kSplitParts: p0 = bf16(v), p1 = bf16(v - p0), p2 = bf16(v - p0 - p1)

__device__ __forceinline__ void store_split(__nv_bfloat16* row, int C, int c, float v) {
#pragma unroll
  for (int i = 0; i < kParts; ++i) {
    const __nv_bfloat16 part = __float2bfloat16_rn(v);
    row[i * C + c] = part;
    v -= __bfloat162float(part);
  }
}

__device__ __forceinline__ float load_split(const __nv_bfloat16* row, int C, int c) {
  return (ldf(row + 2 * C + c) + ldf(row + C + c)) + ldf(row + c);  // exact: the parts tile one fp32 mantissa
}

template <typename T>
__global__ void __launch_bounds__(256) stem_conv_split_kernel(const T* __restrict__ image, const float* __restrict__ w,
                                                              const float* __restrict__ bias, __nv_bfloat16* __restrict__ out,
                                                              int B, int H, int W, int Ho, int Wo, float m0, float m1,
                                                              float m2) {
  // block = 4 output pixels x 64 channels; the 7x7x3 input patch of each pixel is staged in shared memory
  __shared__ float patch[4][148];
  const int c = threadIdx.x & 63, pl = threadIdx.x >> 6;
  const long long pix0 = static_cast<long long>(blockIdx.x) * 4;
  const long long total = static_cast<long long>(B) * Ho * Wo;
  const float mean[3] = {m0, m1, m2};
  for (int i = threadIdx.x; i < 4 * 147; i += 256) {
    const int q = i / 147, k = i - q * 147;
    const long long pix = pix0 + q;
    float v = 0.0f;
    if (pix < total) {
      const int ox = static_cast<int>(pix % Wo), oy = static_cast<int>((pix / Wo) % Ho), b = static_cast<int>(pix / (static_cast<long long>(Wo) * Ho));
      const int r = k / 21, s = (k / 3) % 7, ch = k % 3;
      const int iy = 2 * oy + r - 3, ix = 2 * ox + s - 3;
      if (iy >= 0 && iy < H && ix >= 0 && ix < W)
        v = __fsub_rn(static_cast<float>(image[((static_cast<long long>(b) * H + iy) * W + ix) * 3 + ch]), mean[ch]);
    }
    patch[q][k] = v;
  }
  __syncthreads();
  const long long pix = pix0 + pl;
  if (pix >= total) return;
  float acc = 0.0f;
#pragma unroll 7
  for (int k = 0; k < 147; ++k) acc = fmaf(patch[pl][k], __ldg(w + k * 64 + c), acc);
  acc = fmaxf(acc + __ldg(bias + c), 0.0f);
  const int ox = static_cast<int>(pix % Wo), oy = static_cast<int>((pix / Wo) % Ho), b = static_cast<int>(pix / (static_cast<long long>(Wo) * Ho));
  __nv_bfloat16* row = out + ((static_cast<long long>(b) * (Ho + 3) + oy) * (Wo + 3) + ox) * (kParts * 64);
  store_split(row, 64, c, acc);
}

__global__ void __launch_bounds__(256) maxpool_split_kernel(const __nv_bfloat16* __restrict__ in, __nv_bfloat16* __restrict__ out,
                                                            int B, int Hi, int Wi, int in_hs, int in_ws, int C, int Ho,
                                                            int Wo, int pad_top, int pad_left) {
  const long long t = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const int Hp = Ho + 2, Wp = Wo + 2;
  const long long total = static_cast<long long>(B) * Hp * Wp * C;
  if (t >= total) return;
  const int c = static_cast<int>(t % C);
  const long long pix = t / C;
  const int wx = static_cast<int>(pix % Wp), hy = static_cast<int>((pix / Wp) % Hp), b = static_cast<int>(pix / (static_cast<long long>(Wp) * Hp));
  __nv_bfloat16* row = out + pix * kParts * C;
  float best = 0.0f;
  if (hy >= 1 && hy <= Ho && wx >= 1 && wx <= Wo) {
    best = -INFINITY;  // TF 'same' max-pool ignores padded positions
    const int iy0 = 2 * (hy - 1) - pad_top, ix0 = 2 * (wx - 1) - pad_left;
    for (int dy = 0; dy < 3; ++dy)
      for (int dx = 0; dx < 3; ++dx) {
        const int iy = iy0 + dy, ix = ix0 + dx;
        if (iy < 0 || iy >= Hi || ix < 0 || ix >= Wi) continue;
        const __nv_bfloat16* s = in + ((static_cast<long long>(b) * in_hs + iy) * in_ws + ix) * kParts * C;
        best = fmaxf(best, load_split(s, C, c));
      }
  }
  store_split(row, C, c, best);
}

__global__ void __launch_bounds__(256) add_split_kernel(const __nv_bfloat16* __restrict__ a, const __nv_bfloat16* __restrict__ b,
                                                        const __nv_bfloat16* __restrict__ cc, const __nv_bfloat16* __restrict__ u,
                                                        int uH, int uW, __nv_bfloat16* __restrict__ out, int B, int H,
                                                        int W, int C) {
  const long long t = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const int Hp = H + 2, Wp = W + 2;
  const long long total = static_cast<long long>(B) * Hp * Wp * C;
  if (t >= total) return;
  const int c = static_cast<int>(t % C);
  const long long pix = t / C;
  const int wx = static_cast<int>(pix % Wp), hy = static_cast<int>((pix / Wp) % Hp), bi = static_cast<int>(pix / (static_cast<long long>(Wp) * Hp));
  const long long o = pix * kParts * C;
  // Keras Add sums its inputs left to right in fp32 (retinanet.py:198-211); borders of every input are zero
  float f = load_split(a + o, C, c);
  if (b != nullptr) f += load_split(b + o, C, c);
  if (cc != nullptr) f += load_split(cc + o, C, c);
  if (u != nullptr && hy >= 1 && hy <= H && wx >= 1 && wx <= W) {
    int sy = static_cast<int>(floorf((static_cast<float>(hy - 1) + 0.5f) * (static_cast<float>(uH) / H)));
    int sx = static_cast<int>(floorf((static_cast<float>(wx - 1) + 0.5f) * (static_cast<float>(uW) / W)));
    sy = min(sy, uH - 1);
    sx = min(sx, uW - 1);
    const __nv_bfloat16* s = u + ((static_cast<long long>(bi) * (uH + 2) + sy + 1) * (uW + 2) + sx + 1) * kParts * C;
    f += load_split(s, C, c);
  }
  store_split(out + o, C, c, f);
}

inline unsigned blocks_for(long long total) { return static_cast<unsigned>((total + 255) / 256); }

}  // namespace

cudaError_t launch_stem_conv_split(const void* image, int is_u8, const float mean[3], const float* w, const float* bias,
                                   __nv_bfloat16* out, int B, int H, int W, int Ho, int Wo, cudaStream_t st) {
  const long long total = static_cast<long long>(B) * Ho * Wo;
  if (is_u8)
    stem_conv_split_kernel<unsigned char><<<blocks_for(total * 64), 256, 0, st>>>(
        static_cast<const unsigned char*>(image), w, bias, out, B, H, W, Ho, Wo, mean[0], mean[1], mean[2]);
  else
    stem_conv_split_kernel<float><<<blocks_for(total * 64), 256, 0, st>>>(static_cast<const float*>(image), w, bias, out,
                                                                          B, H, W, Ho, Wo, 0.0f, 0.0f, 0.0f);
  return cudaGetLastError();
}

cudaError_t launch_maxpool3x3s2_split(const __nv_bfloat16* in, __nv_bfloat16* out, int B, int Hi, int Wi, int in_hs,
                                      int in_ws, int C, int Ho, int Wo, cudaStream_t st) {
  const int pt = ((Ho - 1) * 2 + 3 - Hi) > 0 ? ((Ho - 1) * 2 + 3 - Hi) / 2 : 0;
  const int pl = ((Wo - 1) * 2 + 3 - Wi) > 0 ? ((Wo - 1) * 2 + 3 - Wi) / 2 : 0;
  const long long total = static_cast<long long>(B) * (Ho + 2) * (Wo + 2) * C;
  maxpool_split_kernel<<<blocks_for(total), 256, 0, st>>>(in, out, B, Hi, Wi, in_hs, in_ws, C, Ho, Wo, pt, pl);
  return cudaGetLastError();
}

cudaError_t launch_add_split(const __nv_bfloat16* a, const __nv_bfloat16* b, const __nv_bfloat16* c,
                             const __nv_bfloat16* u, int uH, int uW, __nv_bfloat16* out, int B, int H, int W, int C,
                             cudaStream_t st) {
  const long long total = static_cast<long long>(B) * (H + 2) * (W + 2) * C;
  add_split_kernel<<<blocks_for(total), 256, 0, st>>>(a, b, c, u, uH, uW, out, B, H, W, C);
  return cudaGetLastError();
}

}  // namespace ppb
