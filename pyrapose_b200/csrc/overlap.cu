// Anchor <-> ground-truth IoU on the GPU, float64 and bit-exact with the reference's Cython
// routine (PyraPose/utils/compute_overlap.pyx:13-53) and the numpy lines that consume it
// (PyraPose/utils/anchors.py:310-316). Compiled with -fmad=false: every multiply/add is rounded
// separately, exactly as the x86-64 build of the reference does.
//
//  * overlap_matrix_kernel  — materialises the (N, K) matrix (drop-in compute_overlap API).
//    Store-bound: N*K*8 bytes; anchor-per-lane compute (warp-uniform branches), shared-memory transpose, 16-byte stores.
//  * gt_assign_kernel       — fused compute_gt_annotations: argmax / max / positive / ignore per anchor
//    without ever writing the matrix; query boxes staged in shared memory.
#include "../../include/pyrapose_b200.h"
#include "errors.h"
#include "launch_count.h"

namespace ppb {
namespace {

constexpr int kQChunk = 512;  // query boxes staged per pass of gt_assign_kernel (512 * 40 B = 20 KiB)

// grid: (ceil(N / 32), batch), 256 threads = 8 warps. A block owns 32 consecutive anchors = one contiguous slab of the
// output. Compute phase: lane = anchor (its box and area live in registers), warp w takes query boxes w, w+8, ... of the
// staged chunk, so all 32 lanes of a warp test neighbouring anchors against the SAME query box: the iw > 0 / ih > 0
// branches (pyx:41, 46) are almost always warp-uniform and the IEEE fp64 division only runs where boxes really
// overlap. Results go to a [32][chunk | 1] shared tile (odd row stride: conflict-free column writes); write phase:
// one warp per row, 16-byte stores of consecutive k -> every store instruction covers 512 contiguous bytes.
constexpr int kOvRows = 32;
constexpr int kOvQChunk = 128;  // query boxes per pass: 5 KiB of queries + 33 KiB of results

template <int VEC>
__global__ void __launch_bounds__(256) overlap_matrix_kernel(const double* __restrict__ boxes,
                                                             const double* __restrict__ query,
                                                             double* __restrict__ out, long long n, int k,
                                                             long long boxes_bstride, long long query_bstride,
                                                             long long out_bstride) {
  extern __shared__ double sm[];
  const int qc = min(k, kOvQChunk);
  const int ts = qc | 1;
  double* sq = sm;             // [qc][5]: x1, y1, x2, y2, area (pyx:32-35, computed once instead of once per pair)
  double* tile = sm + qc * 5;  // [32][ts]
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const double* qx = query + blockIdx.y * query_bstride;
  const long long row0 = static_cast<long long>(blockIdx.x) * kOvRows;
  const int rows = static_cast<int>(min(static_cast<long long>(kOvRows), n - row0));
  double* o = out + blockIdx.y * out_bstride + row0 * k;
  double b0 = 0, b1 = 0, b2 = 0, b3 = 0;
  if (lane < rows) {
    const double* bp = boxes + blockIdx.y * boxes_bstride + (row0 + lane) * 4;
    const double2 b01 = __ldg(reinterpret_cast<const double2*>(bp));
    const double2 b23 = __ldg(reinterpret_cast<const double2*>(bp + 2));
    b0 = b01.x; b1 = b01.y; b2 = b23.x; b3 = b23.y;
  }
  const double b_area = (b2 - b0 + 1) * (b3 - b1 + 1);  // first product of pyx:47-49, same operands for every k
  double* trow = tile + lane * ts;
  for (int base = 0; base < k; base += kOvQChunk) {
    const int chunk = min(kOvQChunk, k - base);
    __syncthreads();  // the previous pass has finished reading sq / tile
    for (int t = threadIdx.x; t < chunk; t += 256) {
      const double2 q01 = __ldg(reinterpret_cast<const double2*>(qx + (base + t) * 4));
      const double2 q23 = __ldg(reinterpret_cast<const double2*>(qx + (base + t) * 4 + 2));
      sq[t * 5 + 0] = q01.x; sq[t * 5 + 1] = q01.y; sq[t * 5 + 2] = q23.x; sq[t * 5 + 3] = q23.y;
      sq[t * 5 + 4] = (q23.x - q01.x + 1) * (q23.y - q01.y + 1);
    }
    __syncthreads();
    for (int q = warp; q < chunk; q += 8) {
      const double q0 = sq[q * 5], q1 = sq[q * 5 + 1], q2 = sq[q * 5 + 2], q3 = sq[q * 5 + 3];
      double r = 0.0;
      const double iw = (b2 < q2 ? b2 : q2) - (b0 > q0 ? b0 : q0) + 1;  // pyx:37-40
      if (iw > 0) {
        const double ih = (b3 < q3 ? b3 : q3) - (b1 > q1 ? b1 : q1) + 1;  // pyx:42-45
        if (ih > 0) {
          const double ua = b_area + sq[q * 5 + 4] - iw * ih;  // pyx:47-51
          r = iw * ih / ua;                                    // pyx:52
        }
      }
      trow[q] = r;
    }
    __syncthreads();
    for (int r = warp; r < rows; r += 8) {
      const double* src = tile + r * ts;
      double* dst = o + static_cast<long long>(r) * k + base;
      if (VEC == 2) {  // k and base even, output base 16-byte aligned: every row segment starts 16-byte aligned
        for (int c = lane * 2; c < chunk; c += 64)
          *reinterpret_cast<double2*>(dst + c) = make_double2(src[c], src[c + 1]);
      } else {
        for (int c = lane; c < chunk; c += 32) dst[c] = src[c];
      }
    }
  }
}

// grid: (ceil(N / (256 * kGtPerThread)), batch). Every thread owns kGtPerThread anchors 256 apart (coalesced box
// loads) and walks all K queries of the image from shared memory (two 16-byte and one 8-byte broadcast load per query).
// The kernel is bound by the fp64 pipe (compare / select / add chains per pair, the IEEE division where boxes overlap),
// not by shared memory: four anchors per thread measured SLOWER (0.456 vs 0.400 ms at B=64, N=56700, K=100), so the
// blocking factor stays 1.
constexpr int kGtPerThread = 1;

__global__ void __launch_bounds__(256) gt_assign_kernel(const double* __restrict__ boxes,
                                                        const double* __restrict__ query,
                                                        const int* __restrict__ query_count, int n, int k_max,
                                                        long long boxes_bstride, double neg_thr, double pos_thr,
                                                        long long* __restrict__ argmax,
                                                        unsigned char* __restrict__ positive,
                                                        unsigned char* __restrict__ ignore,
                                                        double* __restrict__ max_overlap) {
  __shared__ double2 sq01[kQChunk];
  __shared__ double2 sq23[kQChunk];
  __shared__ double sqa[kQChunk];
  const int b = blockIdx.y;
  const int k = query_count != nullptr ? query_count[b] : k_max;
  const double* qx = query + static_cast<long long>(b) * k_max * 4;
  const int i0 = blockIdx.x * (256 * kGtPerThread) + threadIdx.x;
  double b0[kGtPerThread], b1[kGtPerThread], b2[kGtPerThread], b3[kGtPerThread], barea[kGtPerThread];
  double best[kGtPerThread];
  int best_k[kGtPerThread];
#pragma unroll
  for (int r = 0; r < kGtPerThread; ++r) {
    const int i = i0 + r * 256;
    b0[r] = b1[r] = b2[r] = b3[r] = 0.0;
    if (i < n) {
      const double* bp = boxes + b * boxes_bstride + static_cast<long long>(i) * 4;
      const double2 b01 = __ldg(reinterpret_cast<const double2*>(bp));
      const double2 b23 = __ldg(reinterpret_cast<const double2*>(bp + 2));
      b0[r] = b01.x; b1[r] = b01.y; b2[r] = b23.x; b3[r] = b23.y;
    }
    barea[r] = (b2[r] - b0[r] + 1) * (b3[r] - b1[r] + 1);  // first product of pyx:47-49, same operands for every k
    best[r] = -1.0;  // numpy.argmax: first maximum wins; an all-zero row gives index 0
    best_k[r] = 0;
  }
  for (int base = 0; base < k; base += kQChunk) {
    const int chunk = min(kQChunk, k - base);
    __syncthreads();
    for (int t = threadIdx.x; t < chunk; t += blockDim.x) {
      const double2 q01 = __ldg(reinterpret_cast<const double2*>(qx + (base + t) * 4));
      const double2 q23 = __ldg(reinterpret_cast<const double2*>(qx + (base + t) * 4 + 2));
      sq01[t] = q01;
      sq23[t] = q23;
      sqa[t] = (q23.x - q01.x + 1) * (q23.y - q01.y + 1);  // pyx:32-35
    }
    __syncthreads();
    for (int t = 0; t < chunk; ++t) {
      const double2 q01 = sq01[t];
      const double2 q23 = sq23[t];
      const double qa = sqa[t];
#pragma unroll
      for (int r = 0; r < kGtPerThread; ++r) {
        double ov = 0.0;
        const double iw = (b2[r] < q23.x ? b2[r] : q23.x) - (b0[r] > q01.x ? b0[r] : q01.x) + 1;  // pyx:37-40
        if (iw > 0) {
          const double ih = (b3[r] < q23.y ? b3[r] : q23.y) - (b1[r] > q01.y ? b1[r] : q01.y) + 1;  // pyx:42-45
          if (ih > 0) {
            const double ua = barea[r] + qa - iw * ih;  // pyx:47-51
            ov = iw * ih / ua;                          // pyx:52
          }
        }
        if (ov > best[r]) {
          best[r] = ov;
          best_k[r] = base + t;
        }
      }
    }
  }
#pragma unroll
  for (int r = 0; r < kGtPerThread; ++r) {
    const int i = i0 + r * 256;
    if (i >= n) continue;
    const long long o = static_cast<long long>(b) * n + i;
    double bv = best[r];
    if (k == 0) bv = 0.0;
    argmax[o] = best_k[r];
    const bool pos = bv >= pos_thr;                 // anchors.py:315
    positive[o] = pos;
    ignore[o] = (bv > neg_thr) && !pos;             // anchors.py:316
    if (max_overlap != nullptr) max_overlap[o] = bv;
  }
}

// ------------------------------------------------------------------ anchor_targets_bbox, per-anchor part
// (PyraPose/utils/anchors.py:136-148, 267, 279-285). One thread per (image, anchor): anchor state from the fused
// assignment, one-hot label of the matched box, the 16 corner regression targets
//   ((gt16[argmax][j] - anchor[j % 4]) / (w | h) - 0) / 0.2     (box3D_transform, anchors.py:515-559; float64, then
// stored as float32 exactly like the assignment into the float32 `regression_3D` array), and the "centre outside the
// image" override. boxes16 are the projected 3D-box corners the host computes per object (anchors.py:207-215).
__global__ void __launch_bounds__(256)
anchor_targets_kernel(const double* __restrict__ anchors, const long long* __restrict__ argmax,
                      const unsigned char* __restrict__ positive, const unsigned char* __restrict__ ignore,
                      const int* __restrict__ labels, const double* __restrict__ boxes16,
                      const int* __restrict__ counts, const int* __restrict__ image_wh, int n, int k_max,
                      int num_classes, float* __restrict__ out_reg, float* __restrict__ out_cls) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int b = blockIdx.y;
  if (i >= n) return;
  const double a0 = anchors[i * 4 + 0], a1 = anchors[i * 4 + 1], a2 = anchors[i * 4 + 2], a3 = anchors[i * 4 + 3];
  const long long o = static_cast<long long>(b) * n + i;
  float* reg = out_reg + o * 17;
  float* cls = out_cls + o * (num_classes + 1);
  const int cnt = counts[b];
  float state = 0.0f;
  int label = -1;
  if (cnt > 0) {
    const bool pos = positive[o] != 0;
    if (ignore[o]) state = -1.0f;
    if (pos) state = 1.0f;
    const long long k = argmax[o];
    if (pos) label = labels[static_cast<long long>(b) * k_max + k];
    const double* g = boxes16 + (static_cast<long long>(b) * k_max + k) * 16;
    const double w = a2 - a0, h = a3 - a1;
    const double base[4] = {a0, a1, a2, a3};
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      const double t = (g[j] - base[j & 3]) / ((j & 1) ? h : w);
      reg[j] = static_cast<float>((t - 0.0) / 0.2);
    }
  } else {
#pragma unroll
    for (int j = 0; j < 16; ++j) reg[j] = 0.0f;
  }
  // anchors whose centre lies outside the image are ignored (anchors.py:279-285); width < 0 = image without a shape
  const int iw = image_wh[b * 2], ih = image_wh[b * 2 + 1];
  if (iw >= 0 && ((a0 + a2) / 2 >= iw || (a1 + a3) / 2 >= ih)) state = -1.0f;
  reg[16] = state;
  for (int c = 0; c < num_classes; ++c) cls[c] = (c == label) ? 1.0f : 0.0f;
  cls[num_classes] = state;
}

// ------------------------------------------------------------------ anchor_targets_bbox, mask part (anchors.py:152-164)
// The instance-id image is resized to the P3 grid with PIL's NEAREST rule (src = floor((dst + 0.5) * in / out)); every
// object whose id covers MORE THAN ONE cell marks those cells with its class and state 1.
__device__ __forceinline__ int nearest_index(int dst, int in, int out) {
  int s = static_cast<int>(floor((static_cast<double>(dst) + 0.5) * static_cast<double>(in) / static_cast<double>(out)));
  return s < in - 1 ? s : in - 1;
}

__global__ void __launch_bounds__(256)
mask_count_cells_kernel(const unsigned char* __restrict__ masks, int H, int W, int mh, int mw,
                        const int* __restrict__ mask_ids, const int* __restrict__ counts, int k_max,
                        int* __restrict__ cells) {
  const int cell = blockIdx.x * blockDim.x + threadIdx.x;
  const int b = blockIdx.y;
  if (cell >= mh * mw) return;
  const int y = cell / mw, x = cell - y * mw;
  const int id = masks[(static_cast<long long>(b) * H + nearest_index(y, H, mh)) * W + nearest_index(x, W, mw)];
  for (int k = 0; k < counts[b]; ++k)
    if (mask_ids[b * k_max + k] == id) atomicAdd(cells + b * k_max + k, 1);
}

__global__ void __launch_bounds__(256)
mask_targets_kernel(const unsigned char* __restrict__ masks, int H, int W, int mh, int mw,
                    const int* __restrict__ mask_ids, const int* __restrict__ labels, const int* __restrict__ counts,
                    const int* __restrict__ cells, int k_max, int num_classes, float* __restrict__ out) {
  const int cell = blockIdx.x * blockDim.x + threadIdx.x;
  const int b = blockIdx.y;
  if (cell >= mh * mw) return;
  const int y = cell / mw, x = cell - y * mw;
  const int id = masks[(static_cast<long long>(b) * H + nearest_index(y, H, mh)) * W + nearest_index(x, W, mw)];
  float* row = out + (static_cast<long long>(b) * mh * mw + cell) * (num_classes + 1);
  for (int c = 0; c <= num_classes; ++c) row[c] = 0.0f;
  for (int k = 0; k < counts[b]; ++k)
    if (mask_ids[b * k_max + k] == id && cells[b * k_max + k] > 1) {
      row[labels[b * k_max + k]] = 1.0f;
      row[num_classes] = 1.0f;
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
  const unsigned blocks = static_cast<unsigned>((n + kOvRows - 1) / kOvRows);
  const int qc = k < kOvQChunk ? k : kOvQChunk;
  const size_t smem = (static_cast<size_t>(qc) * 5 + static_cast<size_t>(kOvRows) * (qc | 1)) * sizeof(double);
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
  gt_assign_kernel<<<dim3((n + 256 * kGtPerThread - 1) / (256 * kGtPerThread), batch), 256, 0,
                     static_cast<cudaStream_t>(stream)>>>(
      boxes, query, query_count, n, k_max, bs, negative_overlap, positive_overlap, argmax, positive, ignore,
      max_overlap);
  PPB_CUDA(cudaGetLastError());
  count_launch();
  return kOk;
}

int ppb_anchor_targets(const double* anchors, const long long* argmax, const unsigned char* positive,
                       const unsigned char* ignore, const int* labels, const double* boxes16, const int* counts,
                       const int* image_wh, int n, int k_max, int batch, int num_classes, float* out_regression,
                       float* out_labels, void* stream) {
  PPB_REQUIRE(n >= 0 && k_max >= 0 && batch >= 1 && num_classes >= 1, "ppb_anchor_targets: bad sizes n=%d k=%d batch=%d",
              n, k_max, batch);
  if (n == 0) return kOk;
  PPB_REQUIRE(anchors && argmax && positive && ignore && counts && image_wh && out_regression && out_labels,
              "ppb_anchor_targets: null buffer");
  PPB_REQUIRE(k_max == 0 || (labels != nullptr && boxes16 != nullptr), "ppb_anchor_targets: null annotations");
  PPB_REQUIRE(batch <= 65535, "ppb_anchor_targets: batch out of range");
  anchor_targets_kernel<<<dim3((n + 255) / 256, batch), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      anchors, argmax, positive, ignore, labels, boxes16, counts, image_wh, n, k_max, num_classes, out_regression,
      out_labels);
  PPB_CUDA(cudaGetLastError());
  count_launch();
  return kOk;
}

int ppb_mask_targets(const unsigned char* masks, int height, int width, const int* mask_ids, const int* labels,
                     const int* counts, int k_max, int batch, int mask_h, int mask_w, int num_classes, int* cells,
                     float* out_mask, void* stream) {
  PPB_REQUIRE(height > 0 && width > 0 && mask_h > 0 && mask_w > 0 && batch >= 1 && batch <= 65535 && num_classes >= 1,
              "ppb_mask_targets: bad sizes");
  PPB_REQUIRE(masks && counts && out_mask && (k_max == 0 || (mask_ids && labels && cells)),
              "ppb_mask_targets: null buffer");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const dim3 grid((mask_h * mask_w + 255) / 256, batch);
  if (k_max > 0) {
    PPB_CUDA(cudaMemsetAsync(cells, 0, static_cast<size_t>(batch) * k_max * sizeof(int), st));
    mask_count_cells_kernel<<<grid, 256, 0, st>>>(masks, height, width, mask_h, mask_w, mask_ids, counts, k_max, cells);
    PPB_CUDA(cudaGetLastError());
    count_launch();
  }
  mask_targets_kernel<<<grid, 256, 0, st>>>(masks, height, width, mask_h, mask_w, mask_ids, labels, counts, cells, k_max,
                                            num_classes, out_mask);
  PPB_CUDA(cudaGetLastError());
  count_launch();
  return kOk;
}

}  // extern "C"
