// Score thresholding, per-class NMS, top-k and -1 padding on the GPU.
//
// Replaces (a) FilterDetections / filter_detections (PyraPose/layers/filter_detections.py:21-118, 121-234:
// tf.where -> tf.image.non_max_suppression -> concat -> tf.nn.top_k -> gather -> tf.pad(-1)) and (b) the
// evaluators' host-side vote filter (PyraPose/utils/linemod_eval.py:307-334, 427-430).
//
// Pipeline (all device-resident, no host round trip for the FilterDetections path):
//   class_max_kernel   (class_specific_filter=False only) max / argmax over classes
//   score_mask_kernel  the only bandwidth-relevant pass: reads the (B, N, C) scores once, coalesced through
//                      shared memory, and emits one bit per score (warp ballot) -> mask[B][C][ceil(N/32)]
//   compact_keys_kernel  per (image, class): popc scan over the mask words -> ascending candidate list,
//                      emitted directly as 64-bit sort keys  (~orderable(score) << 32 | anchor index)
//   sort_keys_kernel   per (image, class) bitonic sort (registers + warp shuffles + shared memory up to 8192 keys,
//                      global memory beyond)
//   nms_kernel         greedy NMS in score order, TensorFlow semantics (IoU > thr suppresses, no "+1",
//                      min/max-normalised corners, area <= 0 -> 0, stop at max_detections), float32, no FMA
//   topk_pad_kernel    per image: merge the per-class survivors, top-k by score (ties -> lower position),
//                      gather boxes / boxes3D / labels, pad with -1
// Compiled with -fmad=false so the IoU arithmetic rounds exactly like the reference's.
#include "../../include/pyrapose_b200.h"
#include "errors.h"
#include "launch_count.h"
#include "ptx.cuh"

#include <stdint.h>
#include <cstdlib>

namespace ppb {
namespace {

constexpr int kSortSmemKeys = 8192;
constexpr int kTopkSmemKeys = 16384;
constexpr int kMaxDetCap = 1024;
constexpr int kNmsChunk = 256;

// Monotone map float -> uint32 (ascending order preserved, -0 < +0 irrelevant here).
__device__ __forceinline__ uint32_t orderable(float f) {
  const uint32_t u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
// Key whose ASCENDING order is (score descending, index ascending).
__device__ __forceinline__ uint64_t make_key(float score, uint32_t idx) {
  return (static_cast<uint64_t>(~orderable(score)) << 32) | idx;
}

__host__ __device__ inline int next_pow2(int v) {
  int p = 1;
  while (p < v) p <<= 1;
  return p;
}

// Exclusive block scan of one int per thread (blockDim.x <= 1024); returns the block total in `total`.
__device__ __forceinline__ int block_exclusive_scan(int v, int* warp_sums, int& total) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int incl = v;
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    const int n = __shfl_up_sync(0xffffffffu, incl, off);
    if (lane >= off) incl += n;
  }
  if (lane == 31) warp_sums[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    const int nw = (blockDim.x + 31) >> 5;
    int s = lane < nw ? warp_sums[lane] : 0;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      const int n = __shfl_up_sync(0xffffffffu, s, off);
      if (lane >= off) s += n;
    }
    warp_sums[lane] = s;  // inclusive warp totals
  }
  __syncthreads();
  const int nw = (blockDim.x + 31) >> 5;
  total = warp_sums[nw - 1];
  const int warp_off = warp == 0 ? 0 : warp_sums[warp - 1];
  const int res = warp_off + incl - v;
  __syncthreads();  // warp_sums may be reused by the caller's next scan
  return res;
}

// ------------------------------------------------------------------ class max / argmax
// filter_detections.py:83-84 (keras.backend.max / argmax over axis 1; first maximum wins).
__global__ void __launch_bounds__(256) class_max_kernel(const float* __restrict__ cls, long long total, int C,
                                                        float* __restrict__ smax, int* __restrict__ sarg) {
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const float* p = cls + i * C;
  float best = p[0];
  int arg = 0;
  for (int c = 1; c < C; ++c) {
    const float v = p[c];
    if (v > best) {
      best = v;
      arg = c;
    }
  }
  smax[i] = best;
  sarg[i] = arg;
}

// One warp, one slab row per lane: one LDS + compare + ballot per class (row stride C words: conflict-free for odd
// C); lane c keeps the word of class c so that the C words leave in one store instruction. (Building a per-row bit
// field first and transposing it afterwards measured slower: 30.0 vs 24.3 us at B=32, N=56700, C=15.)
// cnt != nullptr: the population count of each class word is added to cnt[class] (shared or global memory).
template <int CT>
__device__ __forceinline__ void ballot_rows(const float* __restrict__ my, bool in_rows, int C, float thr,
                                            uint32_t* __restrict__ dst, int W, int lane, int* cnt) {
  if constexpr (CT > 0) {  // class count known at compile time (15 / 21 / 30: LineMOD, YCB-V, T-LESS): the loop unrolls fully and
                 // the CT shared-memory loads are all in flight before the first ballot
    float v[CT];
#pragma unroll
    for (int c = 0; c < CT; ++c) v[c] = my[c];
    uint32_t mine = 0u;
#pragma unroll
    for (int c = 0; c < CT; ++c) {
      const uint32_t word = __ballot_sync(0xffffffffu, in_rows && v[c] > thr);
      if (lane == c) mine = word;
    }
    if (lane < CT) {
      dst[static_cast<long long>(lane) * W] = mine;
      if (cnt != nullptr && mine != 0u) atomicAdd(cnt + lane, __popc(mine));
    }
  } else {
    for (int c0 = 0; c0 < C; c0 += 32) {
      uint32_t mine = 0u;
      const int cn = min(32, C - c0);
      for (int c = 0; c < cn; ++c) {
        const uint32_t word = __ballot_sync(0xffffffffu, in_rows && my[c0 + c] > thr);
        if (lane == c) mine = word;
      }
      if (lane < cn) {
        dst[static_cast<long long>(c0 + lane) * W] = mine;
        if (cnt != nullptr && mine != 0u) atomicAdd(cnt + c0 + lane, __popc(mine));
      }
    }
  }
}

// ------------------------------------------------------------------ score -> bit mask
// grid (ceil(N/TILE), B), TILE/32 warps. mask[(b*C + c)*W + n/32] bit (n%32) = scores[b, n, c] > thr.
// The [TILE][C] slab of scores is contiguous in memory: one elected thread brings it to shared memory with a single
// bulk asynchronous copy (cp.async.bulk, completion on an mbarrier) so the whole slab is in flight at once and no
// thread holds load registers; a slab whose base is not 16-byte aligned falls back to coalesced loads, all issued
// before the first shared-memory write. Then warp g owns anchors [32g, 32g+32): one LDS + compare + ballot per
// class (row stride C words: conflict-free for odd C). Lane c keeps the word of class c so that the C words leave in
// one store instruction.
template <int TILE, int CT>
__global__ void __launch_bounds__(TILE) score_mask_kernel(const float* __restrict__ scores, int N, int C, float thr,
                                                          uint32_t* __restrict__ mask, int W,
                                                          int* __restrict__ counts) {
  extern __shared__ __align__(128) float tile[];  // [TILE][C]
  __shared__ __align__(8) uint64_t bar;
  // vote path: per-class hit counts of this block, flushed with one global atomic per class (counts is zeroed by the
  // launcher); TILE = 32 blocks are a single warp and add straight to global memory
  __shared__ int scnt[TILE > 32 ? 64 : 1];
  if (TILE > 32 && counts != nullptr && threadIdx.x < 64) scnt[threadIdx.x] = 0;
  const int b = blockIdx.y;
  const int n0 = blockIdx.x * TILE;
  const int rows = min(TILE, N - n0);
  const long long base = (static_cast<long long>(b) * N + n0) * C;
  const int elems = rows * C;
  const float* src = scores + base;
  if ((reinterpret_cast<uintptr_t>(src) & 15) == 0) {
    const uint32_t bulk_bytes = (static_cast<uint32_t>(elems) * 4u) & ~15u;
    if (threadIdx.x == 0) {
      mbar_init(&bar, 1);
      fence_mbar_init();
      if (bulk_bytes != 0) {
        mbar_arrive_expect_tx(&bar, bulk_bytes);
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                         smem_u32(tile)),
                     "l"(src), "r"(bulk_bytes), "r"(smem_u32(&bar))
                     : "memory");
      } else {
        mbar_arrive(&bar);
      }
    }
    for (int i = static_cast<int>(bulk_bytes >> 2) + threadIdx.x; i < elems; i += TILE) tile[i] = __ldg(src + i);
    __syncthreads();  // the barrier is initialised (and the tail is written) before anyone polls it
    mbar_wait(&bar, 0, 0x51u);
  } else {
    constexpr int kU = 4;
    for (int i0 = threadIdx.x; i0 < elems; i0 += kU * TILE) {
      float v[kU];
#pragma unroll
      for (int u = 0; u < kU; ++u) v[u] = (i0 + u * TILE < elems) ? __ldg(src + i0 + u * TILE) : 0.0f;
#pragma unroll
      for (int u = 0; u < kU; ++u)
        if (i0 + u * TILE < elems) tile[i0 + u * TILE] = v[u];
    }
    __syncthreads();
  }
  const int lane = threadIdx.x & 31, g = threadIdx.x >> 5;
  const int r = g * 32 + lane;
  const int widx = (n0 >> 5) + g;
  int* cnt = counts == nullptr ? nullptr : (TILE > 32 ? scnt : counts + static_cast<long long>(b) * C);
  if (widx < W) {
    const bool in_rows = r < rows;
    const float* my = tile + (in_rows ? r : 0) * C;
    ballot_rows<CT>(my, in_rows, C, thr, mask + static_cast<long long>(b) * C * W + widx, W, lane, cnt);
  }
  if (TILE > 32 && counts != nullptr) {
    __syncthreads();
    if (threadIdx.x < C && scnt[threadIdx.x] != 0)
      atomicAdd(counts + static_cast<long long>(b) * C + threadIdx.x, scnt[threadIdx.x]);
  }
}

// ------------------------------------------------------------------ mask -> sorted-key candidates
// grid (C, B). keys segment of (b,c) starts at (b*C+c)*cap. scores index: (b*N + n)*score_stride + c*class_step.
__global__ void __launch_bounds__(256) compact_keys_kernel(const uint32_t* __restrict__ mask,
                                                           const float* __restrict__ scores, int N, int C, int W,
                                                           int score_stride, int cap, uint64_t* __restrict__ keys,
                                                           int* __restrict__ counts) {
  __shared__ int warp_sums[32];
  const int c = blockIdx.x, b = blockIdx.y;
  const long long bc = static_cast<long long>(b) * C + c;
  const uint32_t* words = mask + bc * W;
  uint64_t* seg = keys + bc * cap;
  int running = 0;
  for (int base = 0; base < W; base += 256) {
    const int wi = base + threadIdx.x;
    uint32_t w = wi < W ? words[wi] : 0u;
    int total;
    int pos = running + block_exclusive_scan(__popc(w), warp_sums, total);
    while (w) {
      const int bit = __ffs(w) - 1;
      w &= w - 1;
      const uint32_t n = static_cast<uint32_t>(wi) * 32u + bit;
      const float s = __ldg(scores + (static_cast<long long>(b) * N + n) * score_stride + c);
      seg[pos++] = make_key(s, n);
    }
    running += total;
  }
  if (threadIdx.x == 0) counts[bc] = running;
}

// ------------------------------------------------------------------ vote gather (evaluator path)
// grid (C, B). Output is CSR: offsets[b*C+c] .. offsets[b*C+c+1] in indices / votes.
__global__ void __launch_bounds__(256) vote_gather_kernel(const uint32_t* __restrict__ mask,
                                                          const float* __restrict__ scores,
                                                          const float* __restrict__ boxes3D, int N, int C, int W,
                                                          const long long* __restrict__ offsets,
                                                          long long* __restrict__ indices, float* __restrict__ votes,
                                                          float* __restrict__ conf) {
  __shared__ int warp_sums[32];
  __shared__ float conf_sum[8];
  const int c = blockIdx.x, b = blockIdx.y;
  const long long bc = static_cast<long long>(b) * C + c;
  const uint32_t* words = mask + bc * W;
  const long long out0 = offsets[bc];
  int running = 0;
  float csum = 0.0f;
  for (int base = 0; base < W; base += 256) {
    const int wi = base + threadIdx.x;
    uint32_t w = wi < W ? words[wi] : 0u;
    int total;
    int pos = running + block_exclusive_scan(__popc(w), warp_sums, total);
    while (w) {
      const int bit = __ffs(w) - 1;
      w &= w - 1;
      const long long n = static_cast<long long>(wi) * 32 + bit;
      const long long row = static_cast<long long>(b) * N + n;
      csum += __ldg(scores + row * C + c);
      indices[out0 + pos] = n;
      const float4* src = reinterpret_cast<const float4*>(boxes3D + row * 16);
      float4* dst = reinterpret_cast<float4*>(votes + (out0 + pos) * 16);
#pragma unroll
      for (int q = 0; q < 4; ++q) dst[q] = __ldg(src + q);
      ++pos;
    }
    running += total;
  }
  // confidence = sum of the hit scores (PyraPose_ROS_wrapper/scripts/pyrapose_node.py:638); summation order is
  // not specified by numpy either, so this value carries a tolerance, not bit-exactness.
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) csum += __shfl_xor_sync(0xffffffffu, csum, off);
  if ((threadIdx.x & 31) == 0) conf_sum[threadIdx.x >> 5] = csum;
  __syncthreads();
  if (threadIdx.x == 0 && conf != nullptr) {
    float t = 0.0f;
    for (int i = 0; i < 8; ++i) t += conf_sum[i];
    conf[bc] = t;
  }
}

// ------------------------------------------------------------------ per-segment bitonic sort (ascending)
__device__ __forceinline__ void bitonic_sort_block(uint64_t* a, int n_pow2) {
  for (int k = 2; k <= n_pow2; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = threadIdx.x; i < n_pow2; i += blockDim.x) {
        const int ixj = i ^ j;
        if (ixj > i) {
          const uint64_t x = a[i], y = a[ixj];
          const bool up = (i & k) == 0;
          if ((x > y) == up) {
            a[i] = y;
            a[ixj] = x;
          }
        }
      }
      __syncthreads();
    }
  }
}

// Bitonic sort of np2 = 1024 * E keys held in REGISTERS, E consecutive keys per thread (key index i = tid * E + e):
// compare distances below E stay inside the thread, distances up to 16 E are warp shuffles, only the rest go through
// shared memory (two buffers alternate, so one barrier per such pass). For 2048 keys that is 15 shared-memory passes out
// of 66; the all-shared-memory form was issue-bound at ~47 instructions per warp and pass (r02 capture: 73 us for 480
// lists of ~1200 keys, no dominant stall reason).
template <int E>
__device__ __forceinline__ void bitonic_sort_regs(uint64_t (&v)[E], int np2, uint64_t* buf0, uint64_t* buf1) {
  const int tid = threadIdx.x;
  int flip = 0;
  for (int k = 2; k <= np2; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      if (j < E) {
        // partner inside the thread; J is a compile-time constant in every copy so v[] stays in registers (indexing it
        // with the runtime j sent it to local memory: 8192 keys 0.29 ms slower than the shared-memory sort)
#pragma unroll
        for (int J = 1; J < E; J <<= 1) {
          if (j == J) {
#pragma unroll
            for (int e = 0; e < E; ++e) {
              if ((e & J) == 0) {
                const int i = tid * E + e;
                const bool up = (i & k) == 0;
                const uint64_t a = v[e], b = v[e | J];
                const bool sw = (a > b) == up;
                v[e] = sw ? b : a;
                v[e | J] = sw ? a : b;
              }
            }
          }
        }
      } else if (j < 32 * E) {
        const int lane_xor = j / E;
#pragma unroll
        for (int e = 0; e < E; ++e) {
          const int i = tid * E + e;
          const uint64_t other = __shfl_xor_sync(0xffffffffu, v[e], lane_xor);
          const bool lower = (i & j) == 0, up = (i & k) == 0;
          const bool take_min = lower == up;
          v[e] = (v[e] < other) == take_min ? v[e] : other;
        }
      } else {
        uint64_t* w = flip ? buf1 : buf0;
        flip ^= 1;
#pragma unroll
        for (int e = 0; e < E; ++e) w[tid * E + e] = v[e];
        __syncthreads();
#pragma unroll
        for (int e = 0; e < E; ++e) {
          const int i = tid * E + e;
          const uint64_t other = w[i ^ j];
          const bool lower = (i & j) == 0, up = (i & k) == 0;
          const bool take_min = lower == up;
          v[e] = (v[e] < other) == take_min ? v[e] : other;
        }
      }
    }
  }
}

template <int E>
__device__ __forceinline__ void sort_segment_regs(uint64_t* seg, int n, int np2, uint64_t* smem) {
  uint64_t v[E];
#pragma unroll
  for (int e = 0; e < E; ++e) {
    const int i = threadIdx.x * E + e;
    v[e] = i < n ? seg[i] : ~0ull;
  }
  bitonic_sort_regs<E>(v, np2, smem, smem + 1024 * E);
#pragma unroll
  for (int e = 0; e < E; ++e) {
    const int i = threadIdx.x * E + e;
    if (i < n) seg[i] = v[e];
  }
}

// grid (segments), 1024 threads, dynamic smem kSortSmemKeys * 8. Segment s holds counts[s] keys at keys + s*cap;
// cap is a power of two, so the global fallback can pad in place.
__global__ void __launch_bounds__(1024) sort_keys_kernel(uint64_t* __restrict__ keys, const int* __restrict__ counts,
                                                         int cap) {
  extern __shared__ uint64_t skeys[];
  const int n = counts[blockIdx.x];
  if (n <= 1) return;
  uint64_t* seg = keys + static_cast<long long>(blockIdx.x) * cap;
  const int np2 = next_pow2(n);
  if (np2 <= 1024) {
    sort_segment_regs<1>(seg, n, np2, skeys);
  } else if (np2 == 2048) {
    sort_segment_regs<2>(seg, n, np2, skeys);
  } else if (np2 == 4096) {
    sort_segment_regs<4>(seg, n, np2, skeys);
  } else if (np2 <= kSortSmemKeys) {
    // 8192 keys: all passes through shared memory (eight keys per thread in registers need the whole 64 KB for ONE
    // exchange buffer and a second barrier per pass: measured 0.59 vs 0.56 ms on the dense set)
    for (int i = threadIdx.x; i < np2; i += blockDim.x) skeys[i] = i < n ? seg[i] : ~0ull;
    __syncthreads();
    bitonic_sort_block(skeys, np2);
    for (int i = threadIdx.x; i < n; i += blockDim.x) seg[i] = skeys[i];
  } else {  // more than kSortSmemKeys = 8192 candidates in one (image, class) list: in place in global memory
    for (int i = n + threadIdx.x; i < np2; i += blockDim.x) seg[i] = ~0ull;
    __syncthreads();
    bitonic_sort_block(seg, np2);
  }
}

// ------------------------------------------------------------------ greedy NMS
struct NBox {
  float a0, a1, a2, a3, area;  // min/max-normalised corners (TF treats them as y1,x1,y2,x2; symmetric)
};

__device__ __forceinline__ NBox normalise(const float4 b) {
  NBox r;
  r.a0 = fminf(b.x, b.z);
  r.a1 = fminf(b.y, b.w);
  r.a2 = fmaxf(b.x, b.z);
  r.a3 = fmaxf(b.y, b.w);
  r.area = (r.a2 - r.a0) * (r.a3 - r.a1);
  return r;
}

// tensorflow/core/kernels/non_max_suppression_op.cc: IOU(); float32, division last.
__device__ __forceinline__ bool iou_exceeds(const NBox& i, const NBox& j, float thr) {
  if (i.area <= 0.0f || j.area <= 0.0f) return 0.0f > thr;
  const float y0 = fmaxf(i.a0, j.a0), x0 = fmaxf(i.a1, j.a1);
  const float y1 = fminf(i.a2, j.a2), x1 = fminf(i.a3, j.a3);
  const float dy = y1 - y0, dx = x1 - x0;
  // Disjoint boxes — nearly every pair — have intersection 0 and IoU 0, which exceeds no threshold >= 0: leave before
  // the union / threshold arithmetic (a third of the instructions of a test; the NMS kernel is issue-bound).
  if (thr >= 0.0f && (dy <= 0.0f || dx <= 0.0f)) return false;
  const float inter = fmaxf(dy, 0.0f) * fmaxf(dx, 0.0f);
  const float uni = i.area + j.area - inter;
  // The IEEE division costs as much as the rest of the test and nearly every pair is far from the threshold: decide by
  // multiplication when inter is outside thr * uni * (1 +- 1e-6) — a margin 16x the rounding error of the quotient and
  // of the product, so the answer is the one the division gives — and divide only inside that band.
  const float t = thr * uni;
  if (inter > t * 1.000001f) return true;
  if (inter < t * 0.999999f) return false;
  return inter / uni > thr;
}

// grid (C, B), 1024 threads = 256 candidates x 4 helpers. Reads the sorted keys of (b,c); writes up to max_det anchor
// indices in selection order to sel_idx[(b*C+c)*max_det ...] and their number to sel_cnt.
// Per 256-candidate chunk (descending score): (1) every candidate is tested against the boxes selected so far, the
// list split four ways over its helpers; (2) the candidate-vs-earlier-candidate suppression bit matrix is built, two
// 32-bit words per helper; (3) ONE warp resolves the greedy order 32 candidates at a time with ballots/shuffles.
// (Building the bit matrix one chunk AHEAD, on the other 31 warps while warp 0 resolves, was measured in r02 and lost:
// the matrix can then no longer skip the candidates phase 1 has killed, and that pruning is worth more than the overlap
// — 0.267 -> 0.296 ms per step on the sparse set, 0.564 -> 0.879 on the dense one.)
constexpr int kNmsWords = kNmsChunk / 32;

// NT threads = NT / 256 helper threads per candidate of the 256-wide chunk (PPB_NMS_THREADS=256|512|1024; smaller
// blocks let several share an SM but measured slightly slower, so 1024 stays the default).
template <int NT>
__global__ void __launch_bounds__(NT) nms_kernel(const uint64_t* __restrict__ keys,
                                                          const int* __restrict__ counts, int cap,
                                                          const float* __restrict__ boxes, int N, int C, int do_nms,
                                                          float thr, int max_det, int* __restrict__ sel_idx,
                                                          int* __restrict__ sel_cnt) {
  __shared__ NBox sel_box[kMaxDetCap];
  __shared__ NBox chunk_box[kNmsChunk];
  __shared__ uint32_t chunk_idx[kNmsChunk];
  __shared__ uint32_t supp[kNmsChunk][kNmsWords + 1];  // +1: odd row stride, conflict-free row-parallel access
  __shared__ unsigned char alive_s[kNmsChunk];
  __shared__ int n_sel_s;
  const int c = blockIdx.x, b = blockIdx.y;
  const long long bc = static_cast<long long>(b) * C + c;
  const uint64_t* seg = keys + bc * cap;
  int* out = sel_idx + bc * max_det;
  const int n = counts[bc];
  const int tid = threadIdx.x;
  if (!do_nms) {  // filter_detections.py:56: without NMS every thresholded candidate goes to top-k
    const int m = min(n, max_det);
    for (int r = tid; r < m; r += blockDim.x) out[r] = static_cast<int>(seg[r] & 0xffffffffu);
    if (tid == 0) sel_cnt[bc] = m;
    return;
  }
  const int cand = tid & (kNmsChunk - 1);
  constexpr int kNmsSubs = NT / kNmsChunk;
  static_assert(kNmsSubs >= 1 && kNmsWords % kNmsSubs == 0, "helpers must divide the 8 suppression words");
  const int sub = tid >> 8;  // warp-uniform helper index 0..kNmsSubs-1
  if (tid == 0) n_sel_s = 0;
  __syncthreads();
  for (int cs = 0; cs < n; cs += kNmsChunk) {
    const int n_sel = n_sel_s;
    if (n_sel >= max_det) break;
    const int i = cs + cand;
    const bool valid = i < n;
    if (sub == 0) {
      NBox me = {0.f, 0.f, 0.f, 0.f, 0.f};
      uint32_t idx = 0;
      if (valid) {
        idx = static_cast<uint32_t>(seg[i] & 0xffffffffu);
        me = normalise(__ldg(reinterpret_cast<const float4*>(boxes) + static_cast<long long>(b) * N + idx));
      }
      chunk_box[cand] = me;
      chunk_idx[cand] = idx;
      alive_s[cand] = valid;
    }
    __syncthreads();
    const NBox me = chunk_box[cand];
    // phase 1: against boxes selected in earlier chunks (helper `sub` takes every fourth one)
    if (valid) {
      bool dead = false;
      for (int s = sub; s < n_sel && !dead; s += kNmsSubs) dead = iou_exceeds(me, sel_box[s], thr);
      if (dead) alive_s[cand] = 0;  // benign race: every writer stores 0
    }
    __syncthreads();
    // phase 2: which earlier (higher-scored) alive members of this chunk would suppress me
    const bool alive = alive_s[cand] != 0;
#pragma unroll
    for (int wi = 0; wi < kNmsWords / kNmsSubs; ++wi) {
      const int w = sub + wi * kNmsSubs;
      uint32_t bits = 0u;
      if (alive && w * 32 < cand) {
        const int jend = min(32, cand - w * 32);
        for (int jj = 0; jj < jend; ++jj) {
          const int j = w * 32 + jj;
          if (alive_s[j] && iou_exceeds(me, chunk_box[j], thr)) bits |= 1u << jj;
        }
      }
      supp[cand][w] = bits;
    }
    __syncthreads();
    // phase 3: greedy resolution by one warp, 32 candidates per step
    if (tid < 32) {
      const int lane = tid;
      uint32_t kept_w[kNmsWords];
#pragma unroll
      for (int w = 0; w < kNmsWords; ++w) kept_w[w] = 0u;
      int ns = n_sel;
#pragma unroll
      for (int sb = 0; sb < kNmsWords; ++sb) {
        if (ns < max_det) {  // warp-uniform
          const int cd = sb * 32 + lane;
          bool a = alive_s[cd] != 0;
#pragma unroll
          for (int w = 0; w < kNmsWords; ++w)
            if (w < sb) a = a && ((supp[cd][w] & kept_w[w]) == 0u);
          const uint32_t intra = supp[cd][sb];
          uint32_t rem = __ballot_sync(0xffffffffu, a);
          uint32_t kept = 0u;
          if (!__any_sync(0xffffffffu, a && (intra & rem) != 0u)) {
            kept = rem;  // no alive member of this group overlaps an alive earlier one: all of them are kept
          } else {
            // Greedy order without walking the lanes one by one: a lane is dead as soon as a kept lane suppresses
            // it, and kept once none of its (lower-lane) suppressors is kept or still undecided. The lowest undecided
            // lane is always decided, so this ends in at most 32 rounds — in practice the depth of the longest
            // suppression chain, 2-4 — and gives exactly the serial result (r02 capture: 47 % of this kernel's warp
            // samples sat at the barrier behind the one-lane-per-step loop this replaces).
            uint32_t und = rem;
            while (und) {  // warp-uniform
              const bool u = (und >> lane) & 1u;
              const bool dead = u && (intra & kept) != 0u;
              const bool keep = u && !dead && (intra & und) == 0u;
              const uint32_t kb = __ballot_sync(0xffffffffu, keep);
              const uint32_t db = __ballot_sync(0xffffffffu, dead);
              kept |= kb;
              und &= ~(kb | db);
            }
          }
          kept_w[sb] = kept;
          if ((kept >> lane) & 1u) {
            const int pos = ns + __popc(kept & ((1u << lane) - 1u));
            if (pos < max_det) {
              sel_box[pos] = chunk_box[cd];
              out[pos] = static_cast<int>(chunk_idx[cd]);
            }
          }
          ns = min(ns + __popc(kept), max_det);
        }
      }
      if (lane == 0) n_sel_s = ns;
    }
    __syncthreads();
  }
  if (tid == 0) sel_cnt[bc] = n_sel_s;
}

// ------------------------------------------------------------------ top-k over classes, gather, pad
// grid (B), 1024 threads, dynamic smem np2*8 (or global scratch when np2 > kTopkSmemKeys).
// Position p = c*max_det + rank preserves the order of the concatenated per-class lists
// (filter_detections.py:81), so "ties -> lower position" equals tf.nn.top_k's lower-index rule.
__global__ void __launch_bounds__(1024) topk_pad_kernel(const int* __restrict__ sel_idx, const int* __restrict__ sel_cnt,
                                                        int Ceff, int max_det, const float* __restrict__ cls,
                                                        const int* __restrict__ sarg, int N, int C,
                                                        const float* __restrict__ boxes,
                                                        const float* __restrict__ boxes3D, float* __restrict__ out_boxes,
                                                        float* __restrict__ out_boxes3D, float* __restrict__ out_scores,
                                                        int* __restrict__ out_labels, int* __restrict__ out_indices,
                                                        uint64_t* __restrict__ scratch, int np2) {
  extern __shared__ uint64_t skeys[];
  __shared__ int total_s;
  __shared__ int order_s[kMaxDetCap];  // slot index of the r-th detection
  const int b = blockIdx.x;
  uint64_t* a = np2 <= kTopkSmemKeys ? skeys : scratch + static_cast<long long>(b) * np2;
  const int slots = Ceff * max_det;
  if (threadIdx.x == 0) {
    int t = 0;
    for (int c = 0; c < Ceff; ++c) t += sel_cnt[b * Ceff + c];
    total_s = t;
  }
  for (int p = threadIdx.x; p < np2; p += blockDim.x) {
    uint64_t key = ~0ull;
    if (p < slots) {
      const int c = p / max_det;
      const int r = p - c * max_det;
      if (r < sel_cnt[b * Ceff + c]) {
        const int idx = sel_idx[(static_cast<long long>(b) * Ceff + c) * max_det + r];
        const long long row = static_cast<long long>(b) * N + idx;
        const int label = sarg != nullptr ? sarg[row] : c;
        key = make_key(__ldg(cls + row * C + label), static_cast<uint32_t>(p));  // filter_detections.py:88
      }
    }
    a[p] = key;
  }
  __syncthreads();
  // top_k over the concatenation (filter_detections.py:81-90) without sorting it: every per-class list is already in
  // ascending key order (NMS selects in descending score, ties by position), so the rank of an element is its position
  // in its own list plus, for every other class, the number of that list's keys below it (binary search). Keys are
  // unique (the slot index is their low word). ~9 shared-memory probes per (element, class) instead of a 91-stage
  // bitonic sort of 8192 keys.
  const int k = min(max_det, total_s);  // filter_detections.py:90
  for (int p = threadIdx.x; p < slots; p += blockDim.x) {
    const uint64_t key = a[p];
    if (key == ~0ull) continue;
    const int c = p / max_det;
    int rank = p - c * max_det;
    for (int c2 = 0; c2 < Ceff && rank < k; ++c2) {
      if (c2 == c) continue;
      const uint64_t* lst = a + c2 * max_det;
      int lo = 0, hi = __ldg(sel_cnt + b * Ceff + c2);
      while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (lst[mid] < key) lo = mid + 1;
        else hi = mid;
      }
      rank += lo;
    }
    if (rank < k) order_s[rank] = p;
  }
  __syncthreads();
  for (int t = threadIdx.x; t < max_det * 16; t += blockDim.x) {
    const int r = t >> 4, j = t & 15;
    float v3 = -1.0f;
    if (r < k) {
      const int p = order_s[r];
      const int c = p / max_det;
      const int idx = sel_idx[(static_cast<long long>(b) * Ceff + c) * max_det + (p - c * max_det)];
      const long long row = static_cast<long long>(b) * N + idx;
      v3 = __ldg(boxes3D + row * 16 + j);
      if (j < 4) out_boxes[(static_cast<long long>(b) * max_det + r) * 4 + j] = __ldg(boxes + row * 4 + j);
      if (j == 4) {
        const int label = sarg != nullptr ? sarg[row] : c;
        out_scores[static_cast<long long>(b) * max_det + r] = __ldg(cls + row * C + label);
        out_labels[static_cast<long long>(b) * max_det + r] = label;
        if (out_indices != nullptr) out_indices[static_cast<long long>(b) * max_det + r] = idx;
      }
    } else {  // filter_detections.py:100-107: pad with -1
      if (j < 4) out_boxes[(static_cast<long long>(b) * max_det + r) * 4 + j] = -1.0f;
      if (j == 4) {
        out_scores[static_cast<long long>(b) * max_det + r] = -1.0f;
        out_labels[static_cast<long long>(b) * max_det + r] = -1;
        if (out_indices != nullptr) out_indices[static_cast<long long>(b) * max_det + r] = -1;
      }
    }
    out_boxes3D[(static_cast<long long>(b) * max_det + r) * 16 + j] = v3;
  }
}

// ------------------------------------------------------------------ generic row gather (FilterDetections `other`)
__global__ void __launch_bounds__(256) gather_rows_kernel(const float* __restrict__ src, const int* __restrict__ indices,
                                                          float* __restrict__ dst, int B, int N, int D, int row_elems) {
  const long long t = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const long long total = static_cast<long long>(B) * D * row_elems;
  if (t >= total) return;
  const int e = static_cast<int>(t % row_elems);
  const long long bd = t / row_elems;
  const int b = static_cast<int>(bd / D);
  const int idx = indices[bd];
  dst[t] = idx >= 0 ? __ldg(src + (static_cast<long long>(b) * N + idx) * row_elems + e) : -1.0f;
}

struct FilterWorkspace {
  uint32_t* mask;
  int* counts;
  uint64_t* keys;
  int* sel_idx;
  int* sel_cnt;
  float* smax;
  int* sarg;
  uint64_t* topk_scratch;
  size_t bytes;
};

size_t align256(size_t v) { return (v + 255) & ~static_cast<size_t>(255); }

FilterWorkspace carve(void* base, int B, int N, int C, int class_specific, int max_det) {
  FilterWorkspace w;
  const int Ceff = class_specific ? C : 1;
  const int W = (N + 31) / 32;
  const size_t cap = next_pow2(N > 0 ? N : 1);
  size_t off = 0;
  char* p = static_cast<char*>(base);
  auto take = [&](size_t bytes) {
    char* r = p != nullptr ? p + off : nullptr;
    off += align256(bytes);
    return r;
  };
  w.mask = reinterpret_cast<uint32_t*>(take(static_cast<size_t>(B) * Ceff * W * 4));
  w.counts = reinterpret_cast<int*>(take(static_cast<size_t>(B) * Ceff * 4));
  w.keys = reinterpret_cast<uint64_t*>(take(static_cast<size_t>(B) * Ceff * cap * 8));
  w.sel_idx = reinterpret_cast<int*>(take(static_cast<size_t>(B) * Ceff * max_det * 4));
  w.sel_cnt = reinterpret_cast<int*>(take(static_cast<size_t>(B) * Ceff * 4));
  w.smax = reinterpret_cast<float*>(take(class_specific ? 0 : static_cast<size_t>(B) * N * 4));
  w.sarg = reinterpret_cast<int*>(take(class_specific ? 0 : static_cast<size_t>(B) * N * 4));
  const int np2 = next_pow2(Ceff * max_det);
  w.topk_scratch = reinterpret_cast<uint64_t*>(take(np2 > kTopkSmemKeys ? static_cast<size_t>(B) * np2 * 8 : 0));
  w.bytes = off;
  return w;
}

// counts != nullptr (vote path): int [B][C] hit counts, produced by the same pass (zeroed here first).
int launch_score_mask(const float* scores, int B, int N, int C, float thr, uint32_t* mask, int* counts,
                      cudaStream_t st) {
  const int W = (N + 31) / 32;
  if (counts != nullptr) PPB_CUDA(cudaMemsetAsync(counts, 0, static_cast<size_t>(B) * C * sizeof(int), st));
  if (C <= 44) {
    const dim3 grid((N + 255) / 256, B);
    const size_t smem = 256 * static_cast<size_t>(C) * 4;
    if (C == 15) score_mask_kernel<256, 15><<<grid, 256, smem, st>>>(scores, N, C, thr, mask, W, counts);
    else if (C == 21) score_mask_kernel<256, 21><<<grid, 256, smem, st>>>(scores, N, C, thr, mask, W, counts);
    else if (C == 30) score_mask_kernel<256, 30><<<grid, 256, smem, st>>>(scores, N, C, thr, mask, W, counts);
    else score_mask_kernel<256, 0><<<grid, 256, smem, st>>>(scores, N, C, thr, mask, W, counts);
  } else {
    PPB_REQUIRE(C <= 352, "score mask: at most 352 classes are supported, got %d", C);
    score_mask_kernel<32, 0><<<dim3((N + 31) / 32, B), 32, 32 * C * 4, st>>>(scores, N, C, thr, mask, W, counts);
  }
  PPB_CUDA(cudaGetLastError());
  count_launch();
  return kOk;
}

// Built-in scratch of ppb_filter_detections_static (see there). 16 MiB: one 480x640 image (56 700 anchors) with up to
// 31 classes class-specific, or two with 15 classes; class-agnostic up to batch 17 at that size.
constexpr size_t kFilterStaticBytes = 16u << 20;
constexpr int kFilterStaticMaxDevices = 64;
static __device__ __align__(256) unsigned char g_filter_static_ws[kFilterStaticBytes];
std::atomic<void*> g_filter_static_ptr[kFilterStaticMaxDevices];  // device address of the array, per device ordinal

}  // namespace
}  // namespace ppb

using namespace ppb;

extern "C" {

size_t ppb_filter_workspace_bytes(int batch, int n, int num_classes, int class_specific, int max_detections) {
  if (batch <= 0 || n <= 0 || num_classes <= 0 || max_detections <= 0) return 256;
  return carve(nullptr, batch, n, num_classes, class_specific, max_detections).bytes + 256;
}

int ppb_filter_detections(const float* boxes, const float* boxes3D, const float* classification, int batch, int n,
                          int num_classes, int class_specific, int nms, float score_threshold, float nms_threshold,
                          int max_detections, float* out_boxes, float* out_boxes3D, float* out_scores,
                          int* out_labels, int* out_indices, void* workspace, size_t workspace_bytes, void* stream) {
  PPB_REQUIRE(batch >= 0 && n > 0 && num_classes > 0, "ppb_filter_detections: bad shape batch=%d n=%d classes=%d",
              batch, n, num_classes);
  PPB_REQUIRE(max_detections > 0 && max_detections <= kMaxDetCap,
              "ppb_filter_detections: max_detections must be in [1, %d], got %d", kMaxDetCap, max_detections);
  if (batch == 0) return kOk;
  PPB_REQUIRE(boxes != nullptr && boxes3D != nullptr && classification != nullptr, "ppb_filter_detections: null input");
  PPB_REQUIRE(out_boxes != nullptr && out_boxes3D != nullptr && out_scores != nullptr && out_labels != nullptr,
              "ppb_filter_detections: null output");
  PPB_REQUIRE((reinterpret_cast<uintptr_t>(boxes) & 15) == 0 && (reinterpret_cast<uintptr_t>(boxes3D) & 15) == 0,
              "ppb_filter_detections: boxes / boxes3D must be 16-byte aligned");
  const size_t need = ppb_filter_workspace_bytes(batch, n, num_classes, class_specific, max_detections);
  PPB_REQUIRE(workspace != nullptr && workspace_bytes >= need,
              "ppb_filter_detections: workspace too small (%zu < %zu bytes)", workspace_bytes, need);
  void* ws_base = reinterpret_cast<void*>((reinterpret_cast<uintptr_t>(workspace) + 255) & ~uintptr_t(255));
  const FilterWorkspace w = carve(ws_base, batch, n, num_classes, class_specific, max_detections);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int Ceff = class_specific ? num_classes : 1;
  const int W = (n + 31) / 32;
  const int cap = next_pow2(n);
  const float* score_src = classification;
  int score_stride = num_classes;
  if (!class_specific) {
    const long long total = static_cast<long long>(batch) * n;
    class_max_kernel<<<static_cast<unsigned>((total + 255) / 256), 256, 0, st>>>(classification, total, num_classes,
                                                                                  w.smax, w.sarg);
    PPB_CUDA(cudaGetLastError());
    count_launch();
    score_src = w.smax;
    score_stride = 1;
  }
  int rc = launch_score_mask(score_src, batch, n, Ceff, score_threshold, w.mask, nullptr, st);
  if (rc != kOk) return rc;
  compact_keys_kernel<<<dim3(Ceff, batch), 256, 0, st>>>(w.mask, score_src, n, Ceff, W, score_stride, cap, w.keys,
                                                         w.counts);
  PPB_CUDA(cudaGetLastError());
  static PerDeviceOnce filter_attr_once;  // per device, see launch_count.h
  if (filter_attr_once.first_on_current_device()) {
    PPB_CUDA(cudaFuncSetAttribute(sort_keys_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSortSmemKeys * 8));
    PPB_CUDA(cudaFuncSetAttribute(topk_pad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kTopkSmemKeys * 8));
  }
  sort_keys_kernel<<<Ceff * batch, 1024, kSortSmemKeys * 8, st>>>(w.keys, w.counts, cap);
  PPB_CUDA(cudaGetLastError());
  static const int nms_threads = [] {
    const char* e = getenv("PPB_NMS_THREADS");
    const int v = e != nullptr ? atoi(e) : 1024;  // measured at B=32, 15 classes: 0.468 ms (1024), 0.503 (512), 0.506 (256)
    return (v == 256 || v == 512) ? v : 1024;
  }();
  if (nms_threads == 256) nms_kernel<256><<<dim3(Ceff, batch), 256, 0, st>>>(w.keys, w.counts, cap, boxes, n, Ceff, nms, nms_threshold,
                                                      max_detections, w.sel_idx, w.sel_cnt);
  else if (nms_threads == 512) nms_kernel<512><<<dim3(Ceff, batch), 512, 0, st>>>(w.keys, w.counts, cap, boxes, n, Ceff, nms, nms_threshold,
                                                      max_detections, w.sel_idx, w.sel_cnt);
  else nms_kernel<1024><<<dim3(Ceff, batch), 1024, 0, st>>>(w.keys, w.counts, cap, boxes, n, Ceff, nms, nms_threshold,
                                                      max_detections, w.sel_idx, w.sel_cnt);
  PPB_CUDA(cudaGetLastError());
  const int np2 = next_pow2(Ceff * max_detections);
  const size_t topk_smem = np2 <= kTopkSmemKeys ? static_cast<size_t>(np2) * 8 : 0;
  topk_pad_kernel<<<batch, 1024, topk_smem, st>>>(w.sel_idx, w.sel_cnt, Ceff, max_detections, classification,
                                                  class_specific ? nullptr : w.sarg, n, num_classes, boxes, boxes3D,
                                                  out_boxes, out_boxes3D, out_scores, out_labels, out_indices,
                                                  w.topk_scratch, np2);
  PPB_CUDA(cudaGetLastError());
  count_launch(4);
  return kOk;
}

int ppb_gather_rows_f32(const float* src, const int* indices, float* dst, int batch, int n, int max_detections,
                        int row_elems, void* stream) {
  PPB_REQUIRE(batch >= 0 && n > 0 && max_detections > 0 && row_elems > 0, "ppb_gather_rows_f32: bad shape");
  const long long total = static_cast<long long>(batch) * max_detections * row_elems;
  if (total == 0) return kOk;
  PPB_REQUIRE(src != nullptr && indices != nullptr && dst != nullptr, "ppb_gather_rows_f32: null buffer");
  gather_rows_kernel<<<static_cast<unsigned>((total + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      src, indices, dst, batch, n, max_detections, row_elems);
  PPB_CUDA(cudaGetLastError());
  count_launch();
  return kOk;
}

/* Vote filter, step 1: mask + per-(image, class) hit counts. mask: uint32 [batch, classes, ceil(n/32)]. */
int ppb_score_mask(const float* scores, int batch, int n, int num_classes, float threshold, uint32_t* mask,
                   int* counts, void* stream) {
  PPB_REQUIRE(batch >= 0 && n > 0 && num_classes > 0, "ppb_score_mask: bad shape");
  if (batch == 0) return kOk;
  PPB_REQUIRE(scores != nullptr && mask != nullptr, "ppb_score_mask: null buffer");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  return launch_score_mask(scores, batch, n, num_classes, threshold, mask, counts, st);
}

/* Vote filter, step 2: ascending hit indices and their 16-vectors in CSR order given by `offsets`
 * (int64 [batch*classes + 1], exclusive prefix sum of the counts from step 1). */
int ppb_vote_gather(const uint32_t* mask, const float* scores, const float* boxes3D, int batch, int n,
                    int num_classes, const long long* offsets, long long* indices, float* votes, float* confidence,
                    void* stream) {
  PPB_REQUIRE(batch >= 0 && n > 0 && num_classes > 0, "ppb_vote_gather: bad shape");
  if (batch == 0) return kOk;
  PPB_REQUIRE(mask != nullptr && scores != nullptr && boxes3D != nullptr && offsets != nullptr,
              "ppb_vote_gather: null input");
  vote_gather_kernel<<<dim3(num_classes, batch), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      mask, scores, boxes3D, n, num_classes, (n + 31) / 32, offsets, indices, votes, confidence);
  PPB_CUDA(cudaGetLastError());
  count_launch();
  return kOk;
}

/* FilterDetections without any allocation by caller or library: the workspace of ppb_filter_detections is the
 * file-scope device array g_filter_static_ws (kFilterStaticBytes; resident on every GPU that loads this library until
 * the process exits). No counterpart in the reference: added for callers that
 * cannot call cudaMalloc (graph capture, tight serving loops).
 *
 * In HOST code `g_filter_static_ws` is only the host-side shadow of the symbol, not a device address: the device
 * address is looked up with cudaGetSymbolAddress once per device (with that device current) and cached.
 * Nothing in the workspace has to start at zero: every buffer carved from it (mask, counts, keys, selections, class
 * maxima, top-k scratch) is fully written by the kernel that produces it before the next one reads it, so the
 * previous call's contents are harmless. All calls on a device share the one array: calls that overlap on the GPU
 * (different streams, no ordering between them) race; callers serialise them. */
int ppb_filter_detections_static(const float* boxes, const float* boxes3D, const float* classification, int batch,
                                 int n, int num_classes, int class_specific, int nms, float score_threshold,
                                 float nms_threshold, int max_detections, float* out_boxes, float* out_boxes3D,
                                 float* out_scores, int* out_labels, int* out_indices, void* stream) {
  PPB_REQUIRE(batch >= 0 && n > 0 && num_classes > 0,
              "ppb_filter_detections_static: bad shape batch=%d n=%d classes=%d", batch, n, num_classes);
  PPB_REQUIRE(max_detections > 0 && max_detections <= kMaxDetCap,
              "ppb_filter_detections_static: max_detections must be in [1, %d], got %d", kMaxDetCap, max_detections);
  if (batch == 0) return kOk;
  const size_t need = ppb_filter_workspace_bytes(batch, n, num_classes, class_specific, max_detections);
  PPB_REQUIRE(need <= kFilterStaticBytes,
              "ppb_filter_detections_static: batch=%d n=%d classes=%d class_specific=%d max_detections=%d needs %zu "
              "bytes of scratch, the built-in array holds %zu (use ppb_filter_detections with a caller workspace)",
              batch, n, num_classes, class_specific, max_detections, need, kFilterStaticBytes);
  int dev = -1;
  PPB_CUDA(cudaGetDevice(&dev));
  PPB_REQUIRE(dev >= 0 && dev < kFilterStaticMaxDevices, "ppb_filter_detections_static: device ordinal %d not supported",
              dev);
  void* ws = g_filter_static_ptr[dev].load(std::memory_order_acquire);
  if (ws == nullptr) {  // set-up, once per device, with that device current
    void* p = nullptr;
    PPB_CUDA(cudaGetSymbolAddress(&p, g_filter_static_ws));
    PPB_REQUIRE(p != nullptr, "ppb_filter_detections_static: cudaGetSymbolAddress returned a null address");
    g_filter_static_ptr[dev].store(p, std::memory_order_release);
    ws = p;
  }
  return ppb_filter_detections(boxes, boxes3D, classification, batch, n, num_classes, class_specific, nms,
                               score_threshold, nms_threshold, max_detections, out_boxes, out_boxes3D, out_scores,
                               out_labels, out_indices, ws, kFilterStaticBytes, stream);
}

size_t ppb_filter_static_capacity_bytes(void) { return kFilterStaticBytes; }

}  // extern "C"
