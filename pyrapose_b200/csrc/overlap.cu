// Anchor <-> ground-truth IoU on the GPU, float64 and bit-exact with the reference's Cython
// routine (PyraPose/utils/compute_overlap.pyx:13-53) and the numpy lines that consume it
// (PyraPose/utils/anchors.py:310-316). Compiled with -fmad=false: every multiply/add is rounded
// separately, exactly as the x86-64 build of the reference does.
//
//  * overlap_matrix_kernel  — materialises the (N, K) matrix (drop-in compute_overlap API).
//    Store-bound: N*K*8 bytes; consecutive threads write consecutive k (16-byte stores when K is even).
//  * gt_assign_kernel       — fused compute_gt_annotations: argmax / max / positive / ignore per anchor
//    without ever writing the matrix; query boxes staged in shared memory.
#include "../../include/pyrapose_b200.h"
#include "errors.h"
#include "launch_count.h"

namespace ppb {
namespace {

__device__ __forceinline__ double iou_plus1(const double b0, const double b1, const double b2, const double b3,
                                            const double q0, const double q1, const double q2, const double q3,
                                            const double q_area) {
  const double iw = (b2 < q2 ? b2 : q2) - (b0 > q0 ? b0 : q0) + 1;  // pyx:37-40
  if (iw > 0) {
    const double ih = (b3 < q3 ? b3 : q3) - (b1 > q1 ? b1 : q1) + 1;  // pyx:42-45
    if (ih > 0) {
      const double ua = (b2 - b0 + 1) * (b3 - b1 + 1) + q_area - iw * ih;  // pyx:47-51
      return iw * ih / ua;                                                // pyx:52
    }
  }
  return 0.0;
}

constexpr int kQChunk = 512;  // query boxes staged per pass (512 * 5 doubles = 20 KiB)

// grid: (ceil(N / kRowsPerBlock), batch), 256 threads. A block owns kRowsPerBlock consecutive anchors = one contiguous
// slab of the output; the image's query boxes and their areas (pyx:32-35, computed once instead of once per pair) sit
// in shared memory. Threads walk the slab with 16-byte stores; all index arithmetic is 32-bit and block-local.
constexpr int kRowsPerBlock = 32;
constexpr int kQSmemMax = 1024;  // query boxes held in shared memory (5 doubles each = 40 KiB)

template <int VEC>
__global__ void __launch_bounds__(256) overlap_matrix_kernel(const double* __restrict__ boxes,
                                                             const double* __restrict__ query,
                                                             double* __restrict__ out, long long n, int k,
                                                             long long boxes_bstride, long long query_bstride,
                                                             long long out_bstride) {
  extern __shared__ double sq[];  // [min(k, kQSmemMax)][5]
  __shared__ double sb[kRowsPerBlock][4];
  const double* bx = boxes + blockIdx.y * boxes_bstride;
  const double* qx = query + blockIdx.y * query_bstride;
  const long long row0 = static_cast<long long>(blockIdx.x) * kRowsPerBlock;
  const int rows = static_cast<int>(min(static_cast<long long>(kRowsPerBlock), n - row0));
  double* o = out + blockIdx.y * out_bstride + row0 * k;
  const int ks = min(k, kQSmemMax);
  for (int t = threadIdx.x; t < ks; t += 256) {
    const double2 q01 = __ldg(reinterpret_cast<const double2*>(qx + t * 4));
    const double2 q23 = __ldg(reinterpret_cast<const double2*>(qx + t * 4 + 2));
    sq[t * 5 + 0] = q01.x; sq[t * 5 + 1] = q01.y; sq[t * 5 + 2] = q23.x; sq[t * 5 + 3] = q23.y;
    sq[t * 5 + 4] = (q23.x - q01.x + 1) * (q23.y - q01.y + 1);  // pyx:32-35
  }
  if (threadIdx.x < rows * 4) (&sb[0][0])[threadIdx.x] = bx[row0 * 4 + threadIdx.x];
  __syncthreads();
  const unsigned total = static_cast<unsigned>(rows) * static_cast<unsigned>(k);
  for (unsigned e0 = threadIdx.x * VEC; e0 < total; e0 += 256 * VEC) {
    const unsigned row = e0 / static_cast<unsigned>(k);
    const unsigned col = e0 - row * static_cast<unsigned>(k);
    const double b0 = sb[row][0], b1 = sb[row][1], b2 = sb[row][2], b3 = sb[row][3];
    double r[VEC];
#pragma unroll
    for (int v = 0; v < VEC; ++v) {
      const unsigned c = col + v;
      if (c < static_cast<unsigned>(ks)) {
        r[v] = iou_plus1(b0, b1, b2, b3, sq[c * 5], sq[c * 5 + 1], sq[c * 5 + 2], sq[c * 5 + 3], sq[c * 5 + 4]);
      } else {  // more query boxes than the shared-memory window holds: read them from global memory
        const double q0 = qx[c * 4], q1 = qx[c * 4 + 1], q2 = qx[c * 4 + 2], q3 = qx[c * 4 + 3];
        r[v] = iou_plus1(b0, b1, b2, b3, q0, q1, q2, q3, (q2 - q0 + 1) * (q3 - q1 + 1));
      }
    }
    if (VEC == 2) {
      *reinterpret_cast<double2*>(o + e0) = make_double2(r[0], r[VEC - 1]);
    } else {
      o[e0] = r[0];
    }
  }
}

// grid: (ceil(N/256), batch). One anchor per thread, all K queries of the image looped from smem.
__global__ void __launch_bounds__(256) gt_assign_kernel(const double* __restrict__ boxes,
                                                        const double* __restrict__ query,
                                                        const int* __restrict__ query_count, int n, int k_max,
                                                        long long boxes_bstride, double neg_thr, double pos_thr,
                                                        long long* __restrict__ argmax,
                                                        unsigned char* __restrict__ positive,
                                                        unsigned char* __restrict__ ignore,
                                                        double* __restrict__ max_overlap) {
  __shared__ double sq[kQChunk * 5];
  const int b = blockIdx.y;
  const int k = query_count != nullptr ? query_count[b] : k_max;
  const double* qx = query + static_cast<long long>(b) * k_max * 4;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  double b0 = 0, b1 = 0, b2 = 0, b3 = 0;
  if (i < n) {
    const double* bp = boxes + b * boxes_bstride + static_cast<long long>(i) * 4;
    const double2 b01 = *reinterpret_cast<const double2*>(bp);
    const double2 b23 = *reinterpret_cast<const double2*>(bp + 2);
    b0 = b01.x; b1 = b01.y; b2 = b23.x; b3 = b23.y;
  }
  double best = -1.0;  // numpy.argmax: first maximum wins; an all-zero row gives index 0
  long long best_k = 0;
  for (int base = 0; base < k; base += kQChunk) {
    const int chunk = min(kQChunk, k - base);
    __syncthreads();
    for (int t = threadIdx.x; t < chunk; t += blockDim.x) {
      const double q0 = qx[(base + t) * 4 + 0], q1 = qx[(base + t) * 4 + 1];
      const double q2 = qx[(base + t) * 4 + 2], q3 = qx[(base + t) * 4 + 3];
      sq[t * 5 + 0] = q0; sq[t * 5 + 1] = q1; sq[t * 5 + 2] = q2; sq[t * 5 + 3] = q3;
      sq[t * 5 + 4] = (q2 - q0 + 1) * (q3 - q1 + 1);
    }
    __syncthreads();
    if (i < n) {
      for (int t = 0; t < chunk; ++t) {
        const double ov = iou_plus1(b0, b1, b2, b3, sq[t * 5], sq[t * 5 + 1], sq[t * 5 + 2], sq[t * 5 + 3], sq[t * 5 + 4]);
        if (ov > best) {
          best = ov;
          best_k = base + t;
        }
      }
    }
  }
  if (i < n) {
    const long long o = static_cast<long long>(b) * n + i;
    if (k == 0) best = 0.0;
    argmax[o] = best_k;
    const bool pos = best >= pos_thr;                 // anchors.py:315
    positive[o] = pos;
    ignore[o] = (best > neg_thr) && !pos;             // anchors.py:316
    if (max_overlap != nullptr) max_overlap[o] = best;
  }
}

}  // namespace
}  // namespace ppb

using namespace ppb;

extern "C" {

int ppb_compute_overlap_f64(const double* boxes, const double* query, double* overlaps, long long n, int k,
                            int batch, int boxes_shared, void* stream) {
  PPB_REQUIRE(n >= 0 && k >= 0 && batch >= 1, "ppb_compute_overlap_f64: bad sizes n=%lld k=%d batch=%d", n, k, batch);
  if (n == 0 || k == 0) return kOk;
  PPB_REQUIRE(boxes != nullptr && query != nullptr && overlaps != nullptr, "ppb_compute_overlap_f64: null buffer");
  const long long total = n * k;
  const long long bs = boxes_shared ? 0 : n * 4;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  PPB_REQUIRE(k <= (1 << 24) && batch <= 65535, "ppb_compute_overlap_f64: k or batch out of range");
  const unsigned blocks = static_cast<unsigned>((n + kRowsPerBlock - 1) / kRowsPerBlock);
  const size_t smem = static_cast<size_t>(k < kQSmemMax ? k : kQSmemMax) * 5 * sizeof(double);
  // even k keeps every row's first element 16-byte aligned (the output base comes from the caller's allocator)
  if ((k & 1) == 0 && (reinterpret_cast<uintptr_t>(overlaps) & 15) == 0) {
    overlap_matrix_kernel<2><<<dim3(blocks, batch), 256, smem, st>>>(boxes, query, overlaps, n, k, bs,
                                                                     static_cast<long long>(k) * 4, total);
  } else {
    overlap_matrix_kernel<1><<<dim3(blocks, batch), 256, smem, st>>>(boxes, query, overlaps, n, k, bs,
                                                                     static_cast<long long>(k) * 4, total);
  }
  PPB_CUDA(cudaGetLastError());
  count_launch();
  return kOk;
}

int ppb_gt_assign_f64(const double* boxes, const double* query, const int* query_count, int n, int k_max, int batch,
                      int boxes_shared, double negative_overlap, double positive_overlap, long long* argmax,
                      unsigned char* positive, unsigned char* ignore, double* max_overlap, void* stream) {
  PPB_REQUIRE(n >= 0 && k_max >= 0 && batch >= 1, "ppb_gt_assign_f64: bad sizes n=%d k=%d batch=%d", n, k_max, batch);
  if (n == 0) return kOk;
  PPB_REQUIRE(boxes != nullptr && argmax != nullptr && positive != nullptr && ignore != nullptr,
              "ppb_gt_assign_f64: null buffer");
  PPB_REQUIRE(k_max == 0 || query != nullptr, "ppb_gt_assign_f64: null query");
  const long long bs = boxes_shared ? 0 : static_cast<long long>(n) * 4;
  gt_assign_kernel<<<dim3((n + 255) / 256, batch), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      boxes, query, query_count, n, k_max, bs, negative_overlap, positive_overlap, argmax, positive, ignore,
      max_overlap);
  PPB_CUDA(cudaGetLastError());
  count_launch();
  return kOk;
}

}  // extern "C"
