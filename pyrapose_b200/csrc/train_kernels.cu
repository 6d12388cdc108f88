// Bandwidth-bound kernels of the training step; see train_kernels.h. Coalesced 128-bit accesses where the layout
// allows, warp-shuffle + shared-memory reductions, double-precision loss accumulators.
#include "train_kernels.h"

namespace ppb {
namespace {

__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ float lo_f(uint32_t u) { return __uint_as_float(u << 16); }
__device__ __forceinline__ float hi_f(uint32_t u) { return __uint_as_float(u & 0xFFFF0000u); }

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, o);
  return v;
}

// Block-wide sum into one atomicAdd (blockDim.x multiple of 32, <= 1024).
__device__ __forceinline__ void block_accumulate(double v, double* dst) {
  __shared__ double part[32];
  v = warp_sum(v);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) part[warp] = v;
  __syncthreads();
  if (warp == 0) {
    const int nw = (blockDim.x + 31) >> 5;
    double s = lane < nw ? part[lane] : 0.0;
    s = warp_sum(s);
    if (lane == 0 && s != 0.0) atomicAdd(dst, s);
  }
  __syncthreads();
}

__global__ void __launch_bounds__(256) count_positive_kernel(const float* __restrict__ y_true, long long rows,
                                                             int width, int* __restrict__ counts, int slot) {
  int local = 0;
  for (long long r = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; r < rows;
       r += static_cast<long long>(gridDim.x) * blockDim.x)
    local += (__ldg(y_true + r * width + (width - 1)) == 1.0f) ? 1 : 0;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) local += __shfl_xor_sync(0xFFFFFFFFu, local, o);
  if ((threadIdx.x & 31) == 0 && local) atomicAdd(counts + slot, local);
}

// Row of location `loc` (index inside one image over all levels) of image b in the padded-flat matrix.
__device__ __forceinline__ long long head_row(const HeadGeom& g, int b, int loc) {
  int l = 0;
  if (g.n_levels > 1 && loc >= g.loc_base[1]) l = 1;
  if (g.n_levels > 2 && loc >= g.loc_base[2]) l = 2;
  const int rem = loc - g.loc_base[l];
  const int y = rem / g.w[l], x = rem - y * g.w[l];
  return g.row_begin[l] + (static_cast<long long>(b) * (g.h[l] + 2) + y + 1) * (g.w[l] + 2) + x + 1;
}

// One thread per (image, location, group of 8 columns): prob and grad are both contiguous over (anchor, class) inside
// a location, so a thread reads 8 consecutive probabilities and writes one 16-byte piece of the gradient row; the
// integer divisions are paid once per 8 columns and the (anchor, class) pair is advanced incrementally.
// Measured per step (two launches, 190 MB; r02, ncu): one thread per column PAIR with five integer divisions and two
// double additions per pair 0.244 ms; this form 0.141 ms; one warp per location with lanes over the columns (every
// load one coalesced 128-byte request) 0.205 ms, the same with all loads of 256 columns issued up front 0.240 ms.
__global__ void __launch_bounds__(256)
focal_loss_grad_kernel(const float* __restrict__ prob, const float* __restrict__ y_true,
                       __nv_bfloat16* __restrict__ grad, int ld, const HeadGeom g, int batch, int A, int C, float alpha,
                       float gamma, const int* __restrict__ counts, int slot, double* __restrict__ loss) {
  const int vals = A * C;
  const unsigned chunks = static_cast<unsigned>((vals + 7) >> 3);
  // 32-bit index arithmetic (the launcher rejects 2^31 or more column groups)
  const unsigned total = static_cast<unsigned>(batch) * static_cast<unsigned>(g.locs) * chunks;
  const float inv_norm = 1.0f / fmaxf(1.0f, static_cast<float>(__ldg(counts + slot)));
  const bool g2 = gamma == 2.0f;
  double local = 0.0;
  for (unsigned t = blockIdx.x * blockDim.x + threadIdx.x; t < total; t += gridDim.x * blockDim.x) {
    const unsigned bl = t / chunks;                      // (image, location)
    const int col0 = static_cast<int>(t - bl * chunks) << 3;
    const int b = static_cast<int>(bl / static_cast<unsigned>(g.locs));
    const int loc = static_cast<int>(bl - static_cast<unsigned>(b) * g.locs);
    int a = col0 / C, c = col0 - a * C;
    const long long n0 = static_cast<long long>(bl) * A;  // first anchor of the location
    const float* yt = y_true + (n0 + a) * (C + 1);
    const float* pp = prob + n0 * C + col0;               // prob[(n0 + a) * C + c] is contiguous over the columns
    float state = __ldg(yt + C);
    float gz[8];
    float part = 0.0f;
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      gz[e] = 0.0f;
      if (col0 + e >= vals) continue;
      if (state != -1.0f) {  // "ignore" anchors are filtered out (losses.py:48-50)
        const float label = __ldg(yt + c);
        const float p = __ldg(pp + e);
        const bool pos = label == 1.0f;
        const float af = pos ? alpha : 1.0f - alpha;
        const float q = pos ? 1.0f - p : p;          // focal_weight base (losses.py:55)
        const float pc = fminf(fmaxf(p, 1e-7f), 1.0f - 1e-7f);  // keras.backend.binary_crossentropy clips to [eps, 1-eps]
        const bool inside = p > 1e-7f && p < 1.0f - 1e-7f;       // gradient of the clip
        // labels are one-hot in practice: one logarithm and one division suffice; the general expression is kept for
        // soft labels (for label 1 / 0 it reduces to exactly these values)
        float bce, dbce;
        // (the kernel is instruction-bound — ncu r02: 63 % issue slots busy, 18 % DRAM — so the one-hot cases use the
        // SFU forms: __logf is accurate to ~2 ulp on [1e-7, 1), the quotient feeds a bf16 gradient)
        if (label == 1.0f) {
          bce = -__logf(pc);
          dbce = -__fdividef(1.0f, pc);
        } else if (label == 0.0f) {
          bce = -__logf(1.0f - pc);
          dbce = __fdividef(1.0f, 1.0f - pc);
        } else {
          bce = -(label * logf(pc) + (1.0f - label) * logf(1.0f - pc));
          dbce = -(label / pc - (1.0f - label) / (1.0f - pc));
        }
        if (!inside) dbce = 0.0f;
        const float fw = af * (g2 ? q * q : powf(q, gamma));
        part += fw * bce;
        // d/dp: focal_weight' * bce + focal_weight * bce'
        const float dq = pos ? -1.0f : 1.0f;
        const float dfw = af * (g2 ? 2.0f * q : gamma * powf(q, gamma - 1.0f)) * dq;
        const float dp = dfw * bce + fw * dbce;
        gz[e] = dp * p * (1.0f - p) * inv_norm;  // through the sigmoid
      }
      if (++c == C) {  // next anchor of the location
        c = 0;
        yt += C + 1;
        if (col0 + e + 1 < vals) state = __ldg(yt + C);
      }
    }
    local += static_cast<double>(part);
    const long long row = head_row(g, b, loc);
    __nv_bfloat16* dst = grad + row * ld + col0;
    if (col0 + 8 <= ld) {
      *reinterpret_cast<uint4*>(dst) = make_uint4(pack2(gz[0], gz[1]), pack2(gz[2], gz[3]), pack2(gz[4], gz[5]),
                                                  pack2(gz[6], gz[7]));
    } else {
#pragma unroll
      for (int e = 0; e < 8; ++e)
        if (col0 + e < ld) dst[e] = __float2bfloat16(gz[e]);
    }
  }
  block_accumulate(local * static_cast<double>(inv_norm), loss + slot);
}

// The 12 (x, y) edge-parallelism differences of losses.py:338-361 as index quadruples of the x coordinate:
// o = (r[a] - r[b]) - (r[c] - r[d]); the y twin uses every index + 1.
__constant__ int kOrthQuad[12][4] = {{0, 6, 2, 4},    {0, 6, 8, 14},  {0, 2, 6, 4},    {0, 2, 8, 10},
                                     {0, 8, 2, 10},   {0, 8, 6, 14},  {12, 10, 14, 8}, {12, 10, 4, 2},
                                     {12, 4, 10, 2},  {12, 4, 14, 6}, {12, 14, 4, 6},  {12, 14, 10, 8}};

// One thread per anchor.
__global__ void __launch_bounds__(256)
orth_l1_loss_grad_kernel(const float* __restrict__ pred, const float* __restrict__ y_true,
                         __nv_bfloat16* __restrict__ grad, int ld, const HeadGeom g, int batch, int A, float weight,
                         float sigma, const int* __restrict__ counts, int slot, double* __restrict__ loss) {
  const long long total = static_cast<long long>(batch) * g.locs * A;
  const float inv_norm = 1.0f / fmaxf(1.0f, static_cast<float>(__ldg(counts + slot)));
  const float s2 = sigma * sigma;
  double local = 0.0;
  for (long long n = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; n < total;
       n += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int a = static_cast<int>(n % A);
    const long long bl = n / A;
    const int loc = static_cast<int>(bl % g.locs);
    const int b = static_cast<int>(bl / g.locs);
    const float* yt = y_true + n * 17;
    float gr[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) gr[j] = 0.0f;
    if (__ldg(yt + 16) == 1.0f) {  // positives only (losses.py:334-336)
      float r[16], t[16];
      const float4* p4 = reinterpret_cast<const float4*>(pred + n * 16);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float4 v = __ldg(p4 + j);
        r[j * 4] = v.x; r[j * 4 + 1] = v.y; r[j * 4 + 2] = v.z; r[j * 4 + 3] = v.w;
      }
#pragma unroll
      for (int j = 0; j < 16; ++j) t[j] = __ldg(yt + j);
      float xy = 0.0f;
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        const float d = r[j] - t[j];
        const float ad = fabsf(d);
        const bool quad = ad < 1.0f / s2;
        xy += quad ? 0.5f * s2 * ad * ad : ad - 0.5f / s2;
        const float sgn = d > 0.0f ? 1.0f : (d < 0.0f ? -1.0f : 0.0f);
        gr[j] = 0.8f * (quad ? s2 * d : sgn);
      }
      float orth = 0.0f;
#pragma unroll
      for (int k = 0; k < 12; ++k) {
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const int ia = kOrthQuad[k][0] + e, ib = kOrthQuad[k][1] + e, ic = kOrthQuad[k][2] + e,
                    id = kOrthQuad[k][3] + e;
          const float o = ((r[ia] - r[ib]) - (r[ic] - r[id])) - ((t[ia] - t[ib]) - (t[ic] - t[id]));
          orth += fabsf(o);
          const float sg = (o > 0.0f ? 1.0f : (o < 0.0f ? -1.0f : 0.0f)) * (0.2f / 24.0f);
          gr[ia] += sg;
          gr[ib] -= sg;
          gr[ic] -= sg;
          gr[id] += sg;
        }
      }
      local += static_cast<double>(weight * (0.8f * xy + 0.2f * orth / 24.0f));
#pragma unroll
      for (int j = 0; j < 16; ++j) gr[j] *= weight * inv_norm;
    }
    const long long row = head_row(g, b, loc);
    uint4 o0, o1;
    o0.x = pack2(gr[0], gr[1]);   o0.y = pack2(gr[2], gr[3]);   o0.z = pack2(gr[4], gr[5]);   o0.w = pack2(gr[6], gr[7]);
    o1.x = pack2(gr[8], gr[9]);   o1.y = pack2(gr[10], gr[11]); o1.z = pack2(gr[12], gr[13]); o1.w = pack2(gr[14], gr[15]);
    uint4* dst = reinterpret_cast<uint4*>(grad + row * ld + a * 16);
    dst[0] = o0;
    dst[1] = o1;
  }
  block_accumulate(local * static_cast<double>(inv_norm), loss + slot);
}

__device__ __forceinline__ void acc8(float (&f)[8], uint4 v) {
  f[0] += lo_f(v.x); f[1] += hi_f(v.x); f[2] += lo_f(v.y); f[3] += hi_f(v.y);
  f[4] += lo_f(v.z); f[5] += hi_f(v.z); f[6] += lo_f(v.w); f[7] += hi_f(v.w);
}

// tf.image.resize NEAREST with half_pixel_centers as UpsampleLike runs it (elementwise.cu add_kernel).
__device__ __forceinline__ int nearest_src(int dst, int in, int out) {
  const int s = static_cast<int>(floorf((static_cast<float>(dst) + 0.5f) * (static_cast<float>(in) / out)));
  return min(s, in - 1);
}

__global__ void __launch_bounds__(256)
grad_combine_kernel(const uint4* __restrict__ a, const uint4* __restrict__ b, const uint4* __restrict__ c,
                    const uint4* __restrict__ d, int dH, int dW, const uint4* __restrict__ p, int pH, int pW,
                    int n_planes, const uint4* __restrict__ mask, uint4* __restrict__ out, int B, int H, int W, int C8) {
  // 32-bit index arithmetic (the launcher rejects 2^32 or more vectors): the five 64-bit divisions of the first
  // version cost more than the memory traffic on the small maps (1.6 TB/s over the seven launches of a step)
  const unsigned t = blockIdx.x * blockDim.x + threadIdx.x;
  const int Hp = H + 2, Wp = W + 2;
  const unsigned total = static_cast<unsigned>(B) * Hp * Wp * C8;
  if (t >= total) return;
  // all loads that do not depend on the pixel coordinates first
  const uint4 va = a != nullptr ? __ldg(a + t) : make_uint4(0u, 0u, 0u, 0u);
  const uint4 vb = b != nullptr ? __ldg(b + t) : make_uint4(0u, 0u, 0u, 0u);
  const uint4 vc = c != nullptr ? __ldg(c + t) : make_uint4(0u, 0u, 0u, 0u);
  const uint4 vm = mask != nullptr ? __ldg(mask + t) : make_uint4(0u, 0u, 0u, 0u);
  float f[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (a != nullptr) acc8(f, va);
  if (b != nullptr) acc8(f, vb);
  if (c != nullptr) acc8(f, vc);
  const unsigned pix = t / static_cast<unsigned>(C8);
  const int ch = static_cast<int>(t - pix * C8);
  const unsigned rowi = pix / static_cast<unsigned>(Wp);
  const int wx = static_cast<int>(pix - rowi * Wp);
  const int bi = static_cast<int>(rowi / static_cast<unsigned>(Hp));
  const int hy = static_cast<int>(rowi - static_cast<unsigned>(bi) * Hp);
  const bool interior = hy >= 1 && hy <= H && wx >= 1 && wx <= W;
  if (interior && d != nullptr) {
    // adjoint of `up = nearest_resize(src HxW -> dH x dW)`: sum the gradient of every fine pixel that read (y, x)
    const int y = hy - 1, x = wx - 1;
    const int y0 = max(0, (y * dH) / H - 1), x0 = max(0, (x * dW) / W - 1);
    for (int fy = y0; fy < min(dH, y0 + 5); ++fy) {
      if (nearest_src(fy, H, dH) != y) continue;
      for (int fx = x0; fx < min(dW, x0 + 5); ++fx) {
        if (nearest_src(fx, W, dW) != x) continue;
        acc8(f, __ldg(d + ((static_cast<long long>(bi) * (dH + 2) + fy + 1) * (dW + 2) + fx + 1) * C8 + ch));
      }
    }
  }
  if (interior && p != nullptr) {
    // adjoint of launch_parity_planes: plane (ph, pw)[i, j] = in[2i + ph, 2j + pw]
    const int y = hy - 1, x = wx - 1;
    const int ph = y & 1, pw = x & 1, i = y >> 1, j = x >> 1;
    if (i < pH && j < pW && (n_planes == 4 || (ph == 0 && pw == 0))) {
      const long long plane_elems = static_cast<long long>(B) * (pH + 2) * (pW + 2) * C8;
      const int plane = n_planes == 4 ? ph * 2 + pw : 0;
      acc8(f, __ldg(p + plane * plane_elems + ((static_cast<long long>(bi) * (pH + 2) + i + 1) * (pW + 2) + j + 1) * C8 + ch));
    }
  }
  if (mask != nullptr) {
    const uint4 m = vm;
    if (!(lo_f(m.x) > 0.f)) f[0] = 0.f;
    if (!(hi_f(m.x) > 0.f)) f[1] = 0.f;
    if (!(lo_f(m.y) > 0.f)) f[2] = 0.f;
    if (!(hi_f(m.y) > 0.f)) f[3] = 0.f;
    if (!(lo_f(m.z) > 0.f)) f[4] = 0.f;
    if (!(hi_f(m.z) > 0.f)) f[5] = 0.f;
    if (!(lo_f(m.w) > 0.f)) f[6] = 0.f;
    if (!(hi_f(m.w) > 0.f)) f[7] = 0.f;
  }
  uint4 o;
  o.x = pack2(f[0], f[1]);
  o.y = pack2(f[2], f[3]);
  o.z = pack2(f[4], f[5]);
  o.w = pack2(f[6], f[7]);
  out[t] = o;
}

// Column sums of a bf16 matrix: block = 8 column vectors (64 columns) x 32 row lanes; 3.9 TB/s over the 25 launches of a
// step. Measured and rejected (r02): four independent loads per thread in flight (3.84 TB/s, no change) and a block of
// 256 columns x 8 rows so that a warp reads 512 contiguous bytes of a row (3.25 TB/s, worse).
__global__ void __launch_bounds__(256) bias_grad_kernel(const uint4* __restrict__ g, long long rows, int ld8, int cout,
                                                        float* __restrict__ dbias) {
  const int cv = threadIdx.x & 7;
  const int rl = threadIdx.x >> 3;
  const int col8 = blockIdx.x * 8 + cv;
  float f[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (col8 < ld8) {
    for (long long r = static_cast<long long>(blockIdx.y) * 32 + rl; r < rows; r += static_cast<long long>(gridDim.y) * 32)
      acc8(f, __ldg(g + r * ld8 + col8));
  }
  __shared__ float red[32][65];
#pragma unroll
  for (int e = 0; e < 8; ++e) red[rl][cv * 8 + e] = f[e];
  __syncthreads();
  if (threadIdx.x < 64) {
    float s = 0.f;
#pragma unroll 8
    for (int r = 0; r < 32; ++r) s += red[r][threadIdx.x];
    const int col = blockIdx.x * 64 + threadIdx.x;
    if (col < cout && s != 0.0f) atomicAdd(dbias + col, s);
  }
}

__global__ void __launch_bounds__(256)
grad_prepare_kernel(const ParamDesc* __restrict__ descs, const int* __restrict__ chunk_desc,
                    const int* __restrict__ chunk_index, float* __restrict__ grads, const float* __restrict__ master,
                    float grad_scale, double* __restrict__ loss) {
  const ParamDesc d = descs[chunk_desc[blockIdx.x]];
  const long long begin = static_cast<long long>(chunk_index[blockIdx.x]) * kOptChunk;
  const long long end = min(d.count, begin + kOptChunk);
  const int k = d.kind == 0 ? d.k_total : 1;
  double sq = 0.0, l2 = 0.0;
  // four elements per round, loads first (the store to grads[] would otherwise order the next load behind it)
  for (long long i0 = begin + threadIdx.x; i0 < end; i0 += 4ll * blockDim.x) {
    float gq[4], wq[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const long long i = i0 + static_cast<long long>(u) * blockDim.x;
      gq[u] = i < end ? grads[d.offset + i] : 0.0f;
      wq[u] = (i < end && d.l2 > 0.0f) ? master[d.offset + i] : 0.0f;
    }
    // output channel of the round's first element by ONE 32-bit division (a tensor has far fewer than 2^31 elements),
    // advanced by subtraction for the other three (they lie blockDim.x apart); only BatchNorm-folded layers need it
    unsigned co = 0, rem = 0;
    if (d.scale != nullptr) {
      co = static_cast<unsigned>(i0) / static_cast<unsigned>(k);
      rem = static_cast<unsigned>(i0) - co * static_cast<unsigned>(k);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const long long i = i0 + static_cast<long long>(u) * blockDim.x;
      if (i >= end) continue;
      float gv = gq[u] * grad_scale;
      if (d.scale != nullptr) {
        gv *= __ldg(d.scale + co);  // chain rule through the folded BatchNorm scale
        rem += blockDim.x;
        while (rem >= static_cast<unsigned>(k)) {
          rem -= static_cast<unsigned>(k);
          ++co;
        }
      }
      if (d.l2 > 0.0f) {
        const float w = wq[u];
        gv += 2.0f * d.l2 * w;
        l2 += static_cast<double>(d.l2 * w * w);
      }
      grads[d.offset + i] = gv;
      sq += static_cast<double>(gv) * gv;
    }
  }
  block_accumulate(sq, loss + kGradSumSq);
  if (d.l2 > 0.0f) block_accumulate(l2, loss + kLossL2);
}

__global__ void __launch_bounds__(256)
adam_update_kernel(const ParamDesc* __restrict__ descs, const int* __restrict__ chunk_desc,
                   const int* __restrict__ chunk_index, const float* __restrict__ grads, float* __restrict__ master,
                   float* __restrict__ m, float* __restrict__ v, float lr_t, float beta1, float beta2, float eps,
                   float clipnorm, const double* __restrict__ loss) {
  const ParamDesc d = descs[chunk_desc[blockIdx.x]];
  // keras.optimizers.clip_norm: g * c / norm when norm >= c (global norm over all gradients, Keras 2.3.1)
  const float norm = static_cast<float>(sqrt(loss[kGradSumSq]));
  const float clip = (clipnorm > 0.0f && norm >= clipnorm) ? clipnorm / norm : 1.0f;
  auto step = [&](long long gi) {
    const float gv = grads[gi] * clip;
    const float mt = beta1 * m[gi] + (1.0f - beta1) * gv;
    const float vt = beta2 * v[gi] + (1.0f - beta2) * gv * gv;
    const float w = master[gi] - lr_t * mt / (sqrtf(vt) + eps);
    m[gi] = mt;
    v[gi] = vt;
    master[gi] = w;
    return w;
  };
  if (d.kind == 1) {
    const long long begin = static_cast<long long>(chunk_index[blockIdx.x]) * kOptChunk;
    const long long end = min(d.count, begin + kOptChunk);
    for (long long i = begin + threadIdx.x; i < end; i += blockDim.x) d.bias_fwd[i] = step(d.offset + i);
    return;
  }
  // kernel tensor: this block owns the 64 x 64 tile (co0.., kk0..) of the [cout][k_total] matrix. Thread = one column
  // kk (its tap / input channel are fixed) and every fourth row: loads and the forward bf16 store run along kk
  // (coalesced); the transposed dgrad copy [ci][tap_pos * g_ld + co] goes through a shared tile and leaves as 128-byte
  // runs along co (it used to be 2-byte stores a whole matrix row apart — most of the optimizer's time).
  __shared__ __nv_bfloat16 tile[64][66];
  const int tiles_k = (d.k_total + 63) >> 6;
  const int tile_co = chunk_index[blockIdx.x] / tiles_k;
  const int co0 = tile_co << 6, kk0 = (chunk_index[blockIdx.x] - tile_co * tiles_k) << 6;
  const int kl = threadIdx.x & 63, r0 = threadIdx.x >> 6;
  const int kk = kk0 + kl;
  if (kk < d.k_total) {
    // four rows per round with all sixteen loads (g, m, v, w of each row) issued before the first store: the one-row
    // loop waited out a full memory latency per row (stores to m / v / w may alias the next row's loads as far as the
    // compiler knows), 0.52 ms for 1.7 GB
    for (int rb = r0; rb < 64; rb += 16) {
      float gq[4], mq[4], vq[4], wq[4];
      long long iq[4];
      bool ok[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int co = co0 + rb + 4 * u;
        ok[u] = co < d.cout;
        iq[u] = static_cast<long long>(ok[u] ? co : co0) * d.k_total + kk;
        const long long gi = d.offset + iq[u];
        gq[u] = grads[gi];
        mq[u] = m[gi];
        vq[u] = v[gi];
        wq[u] = master[gi];
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        if (!ok[u]) continue;
        const int rl = rb + 4 * u, co = co0 + rl;
        const long long gi = d.offset + iq[u];
        const float gv = gq[u] * clip;
        const float mt = beta1 * mq[u] + (1.0f - beta1) * gv;
        const float vt = beta2 * vq[u] + (1.0f - beta2) * gv * gv;
        const float w = wq[u] - lr_t * mt / (sqrtf(vt) + eps);
        m[gi] = mt;
        v[gi] = vt;
        master[gi] = w;
        const float wf = d.scale != nullptr ? w * __ldg(d.scale + co) : w;
        const __nv_bfloat16 wb = __float2bfloat16(wf);
        d.w_fwd[iq[u]] = wb;
        tile[kl][rl] = wb;
      }
    }
  }
  if (d.w_dgrad == nullptr) return;
  __syncthreads();
  const int n_co = min(64, d.cout - co0);
  const int lane = threadIdx.x & 31, wp = threadIdx.x >> 5;
  for (int c = wp; c < 64; c += 8) {  // one warp per kk column of the tile = one row segment of the dgrad matrix
    const int k2 = kk0 + c;
    if (k2 >= d.k_total) break;
    const int tap = k2 / d.cin, ci = k2 - tap * d.cin;
    __nv_bfloat16* dst = d.w_dgrad + static_cast<long long>(ci) * (d.n_taps * d.g_ld) + d.tap_pos[tap] * d.g_ld + co0;
    if (n_co == 64) {
      reinterpret_cast<uint32_t*>(dst)[lane] = *reinterpret_cast<const uint32_t*>(&tile[c][2 * lane]);
    } else {
      for (int r = lane; r < n_co; r += 32) dst[r] = tile[c][r];
    }
  }
}


}  // namespace

inline unsigned blocks_for(long long total, int per = 256) { return static_cast<unsigned>((total + per - 1) / per); }

cudaError_t launch_count_positive(const float* y_true, long long rows, int width, int* counts, int slot,
                                  cudaStream_t st) {
  const unsigned blocks = static_cast<unsigned>(min(static_cast<long long>(148 * 8), (rows + 255) / 256));
  count_positive_kernel<<<blocks, 256, 0, st>>>(y_true, rows, width, counts, slot);
  return cudaGetLastError();
}

cudaError_t launch_focal_loss_grad(const float* prob, const float* y_true, __nv_bfloat16* grad, int ld, HeadGeom g,
                                   int batch, int A, int C, float alpha, float gamma, const int* counts, int slot,
                                   double* loss, cudaStream_t st) {
  const long long total = static_cast<long long>(batch) * g.locs * ((A * C + 7) / 8);
  if (total >= (1ll << 31) || ld % 8 != 0) return cudaErrorInvalidValue;
  const unsigned blocks = static_cast<unsigned>(min(static_cast<long long>(148 * 16), (total + 255) / 256));
  focal_loss_grad_kernel<<<blocks, 256, 0, st>>>(prob, y_true, grad, ld, g, batch, A, C, alpha, gamma, counts, slot,
                                                 loss);
  return cudaGetLastError();
}

cudaError_t launch_orth_l1_loss_grad(const float* pred, const float* y_true, __nv_bfloat16* grad, int ld, HeadGeom g,
                                     int batch, int A, float weight, float sigma, const int* counts, int slot,
                                     double* loss, cudaStream_t st) {
  const long long total = static_cast<long long>(batch) * g.locs * A;
  const unsigned blocks = static_cast<unsigned>(min(static_cast<long long>(148 * 16), (total + 255) / 256));
  orth_l1_loss_grad_kernel<<<blocks, 256, 0, st>>>(pred, y_true, grad, ld, g, batch, A, weight, sigma, counts, slot,
                                                   loss);
  return cudaGetLastError();
}

cudaError_t launch_grad_combine(const __nv_bfloat16* a, const __nv_bfloat16* b, const __nv_bfloat16* c,
                                const __nv_bfloat16* d, int dH, int dW, const __nv_bfloat16* p, int pH, int pW,
                                int n_planes, const __nv_bfloat16* mask, __nv_bfloat16* out, int B, int H, int W, int C,
                                cudaStream_t st) {
  const long long total = static_cast<long long>(B) * (H + 2) * (W + 2) * (C / 8);
  if (total >= (1ll << 32) - 256) return cudaErrorInvalidValue;
  grad_combine_kernel<<<blocks_for(total), 256, 0, st>>>(
      reinterpret_cast<const uint4*>(a), reinterpret_cast<const uint4*>(b), reinterpret_cast<const uint4*>(c),
      reinterpret_cast<const uint4*>(d), dH, dW, reinterpret_cast<const uint4*>(p), pH, pW, n_planes,
      reinterpret_cast<const uint4*>(mask), reinterpret_cast<uint4*>(out), B, H, W, C / 8);
  return cudaGetLastError();
}

cudaError_t launch_bias_grad(const __nv_bfloat16* g, long long rows, int ld, int cout, float* dbias, cudaStream_t st) {
  const int ld8 = ld / 8;
  dim3 grid((cout + 63) / 64, 1);
  long long want = (148 * 4 + grid.x - 1) / grid.x;
  const long long max_y = (rows + 31) / 32;
  if (want > max_y) want = max_y;
  if (want < 1) want = 1;
  grid.y = static_cast<unsigned>(want);
  bias_grad_kernel<<<grid, 256, 0, st>>>(reinterpret_cast<const uint4*>(g), rows, ld8, cout, dbias);
  return cudaGetLastError();
}

cudaError_t launch_grad_prepare(const ParamDesc* descs, int n_descs, const int* chunk_desc, const int* chunk_index,
                                int n_chunks, float* grads, const float* master, float grad_scale, double* loss,
                                cudaStream_t st) {
  (void)n_descs;
  if (n_chunks <= 0) return cudaSuccess;
  grad_prepare_kernel<<<n_chunks, 256, 0, st>>>(descs, chunk_desc, chunk_index, grads, master, grad_scale, loss);
  return cudaGetLastError();
}

cudaError_t launch_adam_update(const ParamDesc* descs, int n_descs, const int* chunk_desc, const int* chunk_index,
                               int n_chunks, const float* grads, float* master, float* m, float* v, float lr_t,
                               float beta1, float beta2, float eps, float clipnorm, const double* loss,
                               cudaStream_t st) {
  (void)n_descs;
  if (n_chunks <= 0) return cudaSuccess;
  adam_update_kernel<<<n_chunks, 256, 0, st>>>(descs, chunk_desc, chunk_index, grads, master, m, v, lr_t, beta1, beta2,
                                               eps, clipnorm, loss);
  return cudaGetLastError();
}

}  // namespace ppb
