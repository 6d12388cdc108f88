// Bandwidth-bound helpers of the PFPN engine; see elementwise.h. Threads move 16-byte vectors (8 bf16 channels),
// consecutive threads walk consecutive channels -> fully coalesced 128-bit accesses; the copy-like kernels give every
// thread several vectors 256 apart and issue all their loads before the first use.
#include "elementwise.h"

#include <stdint.h>

namespace ppb {
namespace {

__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ float lo_f(uint32_t u) { return __uint_as_float(u << 16); }
__device__ __forceinline__ float hi_f(uint32_t u) { return __uint_as_float(u & 0xFFFF0000u); }

template <typename T>
__global__ void __launch_bounds__(256) stem_s2d_kernel(const T* __restrict__ image, uint4* __restrict__ out, int B,
                                                       int H, int W, int Hs, int Ws, float m0, float m1, float m2) {
  // one thread per S pixel: 2x2 input pixels x 3 channels -> 12 bf16 (+4 zero), two 16-byte stores
  const long long t = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const long long total = static_cast<long long>(B) * Hs * Ws;
  if (t >= total) return;
  const int X = static_cast<int>(t % Ws);
  const int Y = static_cast<int>((t / Ws) % Hs);
  const int b = static_cast<int>(t / (static_cast<long long>(Ws) * Hs));
  float v[12];
#pragma unroll
  for (int py = 0; py < 2; ++py) {
    const int iy = 2 * (Y - 2) + py;
#pragma unroll
    for (int px = 0; px < 2; ++px) {
      const int ix = 2 * (X - 2) + px;
      const bool ok = iy >= 0 && iy < H && ix >= 0 && ix < W;
      const T* src = image + ((static_cast<long long>(b) * H + (ok ? iy : 0)) * W + (ok ? ix : 0)) * 3;
      // uint8 input: preprocess_image(mode='caffe') (PyraPose/utils/image.py:52-60) fused in: float32(pixel) - mean[c],
      // the same float32 subtraction the host code does, so the bf16 values are identical to the float32 input path
      v[(py * 2 + px) * 3 + 0] = ok ? __fsub_rn(static_cast<float>(__ldg(src + 0)), m0) : 0.0f;
      v[(py * 2 + px) * 3 + 1] = ok ? __fsub_rn(static_cast<float>(__ldg(src + 1)), m1) : 0.0f;
      v[(py * 2 + px) * 3 + 2] = ok ? __fsub_rn(static_cast<float>(__ldg(src + 2)), m2) : 0.0f;
    }
  }
  uint4 lo, hi;
  lo.x = pack2(v[0], v[1]);
  lo.y = pack2(v[2], v[3]);
  lo.z = pack2(v[4], v[5]);
  lo.w = pack2(v[6], v[7]);
  hi.x = pack2(v[8], v[9]);
  hi.y = pack2(v[10], v[11]);
  hi.z = 0u;
  hi.w = 0u;
  out[t * 2] = lo;
  out[t * 2 + 1] = hi;
}

// max of 8 bf16 pairs as four packed HMNMX2.BF16 (the maximum of two bf16 values is one of them: no rounding, so this
// equals the fp32 max the unpacked form computed, at a seventh of the instructions)
__device__ __forceinline__ uint32_t max2(uint32_t a, uint32_t b) {
  const __nv_bfloat162 r = __hmax2(*reinterpret_cast<const __nv_bfloat162*>(&a), *reinterpret_cast<const __nv_bfloat162*>(&b));
  return *reinterpret_cast<const uint32_t*>(&r);
}
__device__ __forceinline__ uint4 max8(uint4 a, uint4 b) {
  return make_uint4(max2(a.x, b.x), max2(a.y, b.y), max2(a.z, b.z), max2(a.w, b.w));
}

// Splits a vector index of a padded [B, Hp, Wp, C8] tensor; 32-bit arithmetic (the launchers reject tensors with 2^31
// or more vectors): 64-bit divisions and modulos were most of these kernels' instructions.
struct VecPos {
  int b, hy, wx, c;
};
__device__ __forceinline__ VecPos vec_pos(unsigned t, unsigned Hp, unsigned Wp, unsigned C8) {
  VecPos v;
  const unsigned pix = t / C8;
  v.c = static_cast<int>(t - pix * C8);
  const unsigned rowi = pix / Wp;
  v.wx = static_cast<int>(pix - rowi * Wp);
  v.b = static_cast<int>(rowi / Hp);
  v.hy = static_cast<int>(rowi - v.b * Hp);
  return v;
}

constexpr int kEwVecs = 4;  // vectors per thread of the copy-like kernels, 256 apart: all loads issued before the first use

__global__ void __launch_bounds__(256) maxpool_kernel(const uint4* __restrict__ in, uint4* __restrict__ out, int B,
                                                      int Hi, int Wi, int in_hs, int in_ws, int C8, int Ho, int Wo,
                                                      int pad_top, int pad_left) {
  const unsigned t = blockIdx.x * blockDim.x + threadIdx.x;
  const unsigned Hp = Ho + 2, Wp = Wo + 2;
  const unsigned total = static_cast<unsigned>(B) * Hp * Wp * C8;
  if (t >= total) return;
  const VecPos q = vec_pos(t, Hp, Wp, C8);
  uint4 r = make_uint4(0, 0, 0, 0);
  if (q.hy >= 1 && q.hy <= Ho && q.wx >= 1 && q.wx <= Wo) {
    // TF 'same' max-pool ignores padded positions: they enter the maximum as -inf (bf16 0xFF80); the window always
    // holds at least one real pixel. The nine loads are independent and issued together.
    const uint4 ninf = make_uint4(0xFF80FF80u, 0xFF80FF80u, 0xFF80FF80u, 0xFF80FF80u);
    const int iy0 = 2 * (q.hy - 1) - pad_top, ix0 = 2 * (q.wx - 1) - pad_left;
    uint4 v[9];
#pragma unroll
    for (int dy = 0; dy < 3; ++dy) {
#pragma unroll
      for (int dx = 0; dx < 3; ++dx) {
        const int iy = iy0 + dy, ix = ix0 + dx;
        const bool ok = iy >= 0 && iy < Hi && ix >= 0 && ix < Wi;
        v[dy * 3 + dx] = ok ? __ldg(in + ((static_cast<long long>(q.b) * in_hs + iy) * in_ws + ix) * C8 + q.c) : ninf;
      }
    }
    r = max8(max8(max8(v[0], v[1]), max8(v[2], v[3])), max8(max8(v[4], v[5]), max8(max8(v[6], v[7]), v[8])));
  }
  out[t] = r;
}

__global__ void __launch_bounds__(256) parity_planes_kernel(const uint4* __restrict__ in, uint4* __restrict__ out, int B,
                                                            int H, int W, int C8, int Ho, int Wo, int n_planes) {
  const unsigned t0 = blockIdx.x * (256u * kEwVecs) + threadIdx.x;
  const unsigned Hp = Ho + 2, Wp = Wo + 2;
  const unsigned plane_elems = static_cast<unsigned>(B) * Hp * Wp * C8;
  const unsigned total = plane_elems * n_planes;
  uint4 r[kEwVecs];
#pragma unroll
  for (int k = 0; k < kEwVecs; ++k) {
    const unsigned t = t0 + k * 256u;
    r[k] = make_uint4(0, 0, 0, 0);
    if (t >= total) continue;
    const int plane = static_cast<int>(t / plane_elems);
    const VecPos q = vec_pos(t - plane * plane_elems, Hp, Wp, C8);
    if (q.hy >= 1 && q.hy <= Ho && q.wx >= 1 && q.wx <= Wo) {
      const int iy = 2 * (q.hy - 1) + (plane >> 1), ix = 2 * (q.wx - 1) + (plane & 1);
      if (iy < H && ix < W)
        r[k] = __ldg(in + ((static_cast<long long>(q.b) * (H + 2) + iy + 1) * (W + 2) + ix + 1) * C8 + q.c);
    }
  }
#pragma unroll
  for (int k = 0; k < kEwVecs; ++k) {
    const unsigned t = t0 + k * 256u;
    if (t < total) out[t] = r[k];
  }
}

__device__ __forceinline__ void acc8(float (&f)[8], uint4 v) {
  f[0] += lo_f(v.x); f[1] += hi_f(v.x); f[2] += lo_f(v.y); f[3] += hi_f(v.y);
  f[4] += lo_f(v.z); f[5] += hi_f(v.z); f[6] += lo_f(v.w); f[7] += hi_f(v.w);
}

__global__ void __launch_bounds__(256) add_kernel(const uint4* __restrict__ a, const uint4* __restrict__ b,
                                                  const uint4* __restrict__ c, const uint4* __restrict__ u, int uH,
                                                  int uW, uint4* __restrict__ out, int B, int H, int W, int C8) {
  constexpr int V = 2;
  const unsigned t0 = blockIdx.x * (256u * V) + threadIdx.x;
  const unsigned Hp = H + 2, Wp = W + 2;
  const unsigned total = static_cast<unsigned>(B) * Hp * Wp * C8;
  const uint4 zero = make_uint4(0, 0, 0, 0);
  uint4 va[V], vb[V], vc[V], vu[V];
  // every load of the thread's V vectors is issued before the first sum
#pragma unroll
  for (int k = 0; k < V; ++k) {
    const unsigned t = t0 + k * 256u;
    const bool valid = t < total;
    va[k] = valid ? __ldg(a + t) : zero;
    vb[k] = (valid && b != nullptr) ? __ldg(b + t) : zero;
    vc[k] = (valid && c != nullptr) ? __ldg(c + t) : zero;
    vu[k] = zero;
    if (valid && u != nullptr) {
      const VecPos q = vec_pos(t, Hp, Wp, C8);
      if (q.hy >= 1 && q.hy <= H && q.wx >= 1 && q.wx <= W) {
        // tf.image.resize NEAREST, half_pixel_centers: src = min(floor((dst + 0.5) * in / out), in - 1)
        int sy = static_cast<int>(floorf((static_cast<float>(q.hy - 1) + 0.5f) * (static_cast<float>(uH) / H)));
        int sx = static_cast<int>(floorf((static_cast<float>(q.wx - 1) + 0.5f) * (static_cast<float>(uW) / W)));
        sy = min(sy, uH - 1);
        sx = min(sx, uW - 1);
        vu[k] = __ldg(u + ((static_cast<long long>(q.b) * (uH + 2) + sy + 1) * (uW + 2) + sx + 1) * C8 + q.c);
      }
    }
  }
#pragma unroll
  for (int k = 0; k < V; ++k) {
    const unsigned t = t0 + k * 256u;
    if (t >= total) continue;
    // Keras Add sums its inputs left to right in fp32; the operands here are bf16 so the fp32 sums are exact up to
    // the final rounding (adding an absent operand's +0 changes nothing). Borders: all inputs carry zero borders.
    float f[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    acc8(f, va[k]);
    if (b != nullptr) acc8(f, vb[k]);
    if (c != nullptr) acc8(f, vc[k]);
    if (u != nullptr) acc8(f, vu[k]);
    uint4 o;
    o.x = pack2(f[0], f[1]);
    o.y = pack2(f[2], f[3]);
    o.z = pack2(f[4], f[5]);
    o.w = pack2(f[6], f[7]);
    out[t] = o;
  }
}

inline unsigned blocks_for(long long total) { return static_cast<unsigned>((total + 255) / 256); }
// the kernels index with 32 bits
#define PPB_EW_CHECK(total) \
  do {                      \
    if ((total) >= (1ll << 31)) return cudaErrorInvalidValue; \
  } while (0)

}  // namespace

cudaError_t launch_stem_s2d(const float* image, __nv_bfloat16* out, int B, int H, int W, int Ho, int Wo,
                            cudaStream_t st) {
  const long long total = static_cast<long long>(B) * (Ho + 3) * (Wo + 3);
  stem_s2d_kernel<float><<<blocks_for(total), 256, 0, st>>>(image, reinterpret_cast<uint4*>(out), B, H, W, Ho + 3,
                                                            Wo + 3, 0.0f, 0.0f, 0.0f);
  return cudaGetLastError();
}

cudaError_t launch_stem_s2d_u8(const unsigned char* image, const float mean[3], __nv_bfloat16* out, int B, int H, int W,
                               int Ho, int Wo, cudaStream_t st) {
  const long long total = static_cast<long long>(B) * (Ho + 3) * (Wo + 3);
  stem_s2d_kernel<unsigned char><<<blocks_for(total), 256, 0, st>>>(image, reinterpret_cast<uint4*>(out), B, H, W,
                                                                    Ho + 3, Wo + 3, mean[0], mean[1], mean[2]);
  return cudaGetLastError();
}

cudaError_t launch_maxpool3x3s2(const __nv_bfloat16* in, __nv_bfloat16* out, int B, int Hi, int Wi, int in_hs,
                                int in_ws, int C, int Ho, int Wo, cudaStream_t st) {
  const int pt = ((Ho - 1) * 2 + 3 - Hi) > 0 ? ((Ho - 1) * 2 + 3 - Hi) / 2 : 0;
  const int pl = ((Wo - 1) * 2 + 3 - Wi) > 0 ? ((Wo - 1) * 2 + 3 - Wi) / 2 : 0;
  const long long total = static_cast<long long>(B) * (Ho + 2) * (Wo + 2) * (C / 8);
  PPB_EW_CHECK(total);
  maxpool_kernel<<<blocks_for(total), 256, 0, st>>>(reinterpret_cast<const uint4*>(in), reinterpret_cast<uint4*>(out), B,
                                                    Hi, Wi, in_hs, in_ws, C / 8, Ho, Wo, pt, pl);
  return cudaGetLastError();
}

cudaError_t launch_parity_planes(const __nv_bfloat16* in, __nv_bfloat16* out, int B, int H, int W, int C, int Ho,
                                 int Wo, int n_planes, cudaStream_t st) {
  const long long total = static_cast<long long>(B) * (Ho + 2) * (Wo + 2) * (C / 8) * n_planes;
  PPB_EW_CHECK(total);
  parity_planes_kernel<<<blocks_for((total + kEwVecs - 1) / kEwVecs), 256, 0, st>>>(reinterpret_cast<const uint4*>(in),
                                                          reinterpret_cast<uint4*>(out), B, H, W, C / 8, Ho, Wo,
                                                          n_planes);
  return cudaGetLastError();
}

cudaError_t launch_add(const __nv_bfloat16* a, const __nv_bfloat16* b, const __nv_bfloat16* c,
                       const __nv_bfloat16* u, int uH, int uW, __nv_bfloat16* out, int B, int H, int W, int C,
                       cudaStream_t st) {
  const long long total = static_cast<long long>(B) * (H + 2) * (W + 2) * (C / 8);
  PPB_EW_CHECK(total);
  add_kernel<<<blocks_for((total + 1) / 2), 256, 0, st>>>(reinterpret_cast<const uint4*>(a), reinterpret_cast<const uint4*>(b),
                                                reinterpret_cast<const uint4*>(c), reinterpret_cast<const uint4*>(u),
                                                uH, uW, reinterpret_cast<uint4*>(out), B, H, W, C / 8);
  return cudaGetLastError();
}

}  // namespace ppb
