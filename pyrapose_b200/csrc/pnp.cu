// Pose from corner votes on the GPU: the step right after the hot path, which every evaluator of
// the reference runs on the host per (image, class):
//   cv2.solvePnPRansac(obj_points (k*8,1,3), est_points (k*8,1,2), K, iterationsCount=300, reprojectionError=5.0,
//                      confidence=0.99, flags=SOLVEPNP_ITERATIVE)
//   (PyraPose/utils/linemod_eval.py:479-485, PyraPose_ROS_wrapper/scripts/pyrapose_node.py:652-659)
// where every vote (anchor with score > threshold) contributes the 8 projected corners of the object's 3D box.
//
// B200-native formulation, batched over all (image, class) groups of a batch at once:
//   1. hypotheses   one per sampled vote (<= `iterations`): a vote already holds all 8 non-coplanar box corners, so the
//                   minimal solver is POSIT (DeMenthon & Davis) on that vote — a 3x7 pseudo-inverse per group, a few
//                   fixed-point iterations per thread, no polynomial roots.
//   2. scoring      every hypothesis against every correspondence of its group; inlier <=> reprojection error <= thr
//                   (OpenCV's rule: squared error <= thr^2). One block per (group, hypothesis).
//   3. refinement   the best hypothesis (most inliers, lowest index on ties) is refined on its inliers by damped
//                   Gauss-Newton on the 6-DoF reprojection error (what OpenCV's final SOLVEPNP_ITERATIVE pass
//                   minimises): normal equations reduced over the block in double precision, 6x6 Cholesky.
// Parity with OpenCV is statistical in the RANSAC part (different sampler) and exact in the objective: with the same
// inlier set both converge to the same minimiser (tests/test_pnp_gpu.py).
#include "../../include/pyrapose_b200.h"
#include "errors.h"
#include "launch_count.h"

#include <math.h>

namespace ppb {
namespace {

struct Pose {
  float r[9];  // row-major rotation
  float t[3];
};

__device__ __forceinline__ void cross3(const float* a, const float* b, float* c) {
  c[0] = a[1] * b[2] - a[2] * b[1];
  c[1] = a[2] * b[0] - a[0] * b[2];
  c[2] = a[0] * b[1] - a[1] * b[0];
}

__device__ __forceinline__ float norm3(const float* a) { return sqrtf(a[0] * a[0] + a[1] * a[1] + a[2] * a[2]); }

// pinv of the 7x3 matrix of corner offsets (P_i - P_0): B = (A^T A)^-1 A^T, 3x7. Returns false if singular.
__device__ bool posit_pinv(const float* obj /*[8][3]*/, float* B /*[3][7]*/) {
  float a[7][3];
  float m[6] = {0, 0, 0, 0, 0, 0};  // xx xy xz yy yz zz
  for (int i = 0; i < 7; ++i) {
    for (int c = 0; c < 3; ++c) a[i][c] = obj[(i + 1) * 3 + c] - obj[c];
    m[0] += a[i][0] * a[i][0]; m[1] += a[i][0] * a[i][1]; m[2] += a[i][0] * a[i][2];
    m[3] += a[i][1] * a[i][1]; m[4] += a[i][1] * a[i][2]; m[5] += a[i][2] * a[i][2];
  }
  const float c00 = m[3] * m[5] - m[4] * m[4], c01 = m[2] * m[4] - m[1] * m[5], c02 = m[1] * m[4] - m[2] * m[3];
  const float det = m[0] * c00 + m[1] * c01 + m[2] * c02;
  if (!(fabsf(det) > 1e-12f * (m[0] + m[3] + m[5]) * (m[0] + m[3] + m[5]) * (m[0] + m[3] + m[5]))) return false;
  const float id = 1.0f / det;
  const float inv[9] = {c00 * id, c01 * id, c02 * id,
                        c01 * id, (m[0] * m[5] - m[2] * m[2]) * id, (m[1] * m[2] - m[0] * m[4]) * id,
                        c02 * id, (m[1] * m[2] - m[0] * m[4]) * id, (m[0] * m[3] - m[1] * m[1]) * id};
  for (int r = 0; r < 3; ++r)
    for (int i = 0; i < 7; ++i) B[r * 7 + i] = inv[r * 3] * a[i][0] + inv[r * 3 + 1] * a[i][1] + inv[r * 3 + 2] * a[i][2];
  return true;
}

// POSIT on one vote: x, y are the 8 corners in NORMALISED image coordinates ((u - cx) / fx, (v - cy) / fy).
__device__ bool posit(const float* obj, const float* B, const float* x, const float* y, Pose& out) {
  float eps[7] = {0, 0, 0, 0, 0, 0, 0};
  float iv[3], jv[3], kv[3], z0 = 0.0f;
  for (int it = 0; it < 12; ++it) {
    float I[3] = {0, 0, 0}, J[3] = {0, 0, 0};
    for (int i = 0; i < 7; ++i) {
      const float xp = x[i + 1] * (1.0f + eps[i]) - x[0];
      const float yp = y[i + 1] * (1.0f + eps[i]) - y[0];
      for (int r = 0; r < 3; ++r) {
        I[r] += B[r * 7 + i] * xp;
        J[r] += B[r * 7 + i] * yp;
      }
    }
    const float s1 = norm3(I), s2 = norm3(J);
    if (!(s1 > 1e-12f) || !(s2 > 1e-12f)) return false;
    for (int r = 0; r < 3; ++r) {
      iv[r] = I[r] / s1;
      jv[r] = J[r] / s2;
    }
    cross3(iv, jv, kv);
    const float nk = norm3(kv);
    if (!(nk > 1e-6f)) return false;
    for (int r = 0; r < 3; ++r) kv[r] /= nk;
    z0 = 2.0f / (s1 + s2);
    for (int i = 0; i < 7; ++i) {
      const float a0 = obj[(i + 1) * 3] - obj[0], a1 = obj[(i + 1) * 3 + 1] - obj[1], a2 = obj[(i + 1) * 3 + 2] - obj[2];
      eps[i] = (a0 * kv[0] + a1 * kv[1] + a2 * kv[2]) / z0;
    }
  }
  // orthonormal frame: keep i, k; j = k x i
  cross3(kv, iv, jv);
  for (int r = 0; r < 3; ++r) {
    out.r[r] = iv[r];
    out.r[3 + r] = jv[r];
    out.r[6 + r] = kv[r];
  }
  // translation of the object ORIGIN: camera position of P_0 minus R * P_0
  const float c0[3] = {x[0] * z0, y[0] * z0, z0};
  for (int r = 0; r < 3; ++r)
    out.t[r] = c0[r] - (out.r[r * 3] * obj[0] + out.r[r * 3 + 1] * obj[1] + out.r[r * 3 + 2] * obj[2]);
  return z0 > 0.0f && isfinite(z0);
}

// Vote sampled by hypothesis h of a group with k votes: every vote when k <= n_hyp, else an even stride.
__device__ __forceinline__ long long sample_vote(int h, long long k, int n_hyp) {
  return k <= n_hyp ? h : (static_cast<long long>(h) * k) / n_hyp;
}

// grid (ceil(n_hyp / 128), groups): one hypothesis per thread.
__global__ void __launch_bounds__(128)
pnp_hypotheses_kernel(const float* __restrict__ votes, const long long* __restrict__ offsets,
                      const float* __restrict__ obj, const float* __restrict__ cam, int n_hyp,
                      Pose* __restrict__ hyp, unsigned char* __restrict__ valid) {
  const int g = blockIdx.y;
  const int h = blockIdx.x * blockDim.x + threadIdx.x;
  __shared__ float sobj[24];
  __shared__ float sB[21];
  __shared__ int ok;
  if (threadIdx.x < 24) sobj[threadIdx.x] = obj[g * 24 + threadIdx.x];
  __syncthreads();
  if (threadIdx.x == 0) ok = posit_pinv(sobj, sB) ? 1 : 0;
  __syncthreads();
  if (h >= n_hyp) return;
  const long long begin = offsets[g], k = offsets[g + 1] - begin;
  const long long o = static_cast<long long>(g) * n_hyp + h;
  if (!ok || h >= k) {
    valid[o] = 0;
    return;
  }
  const float fx = cam[g * 4], fy = cam[g * 4 + 1], cx = cam[g * 4 + 2], cy = cam[g * 4 + 3];
  const float4* v4 = reinterpret_cast<const float4*>(votes + (begin + sample_vote(h, k, n_hyp)) * 16);
  float x[8], y[8];
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const float4 v = __ldg(v4 + q);
    x[q * 2] = (v.x - cx) / fx;
    y[q * 2] = (v.y - cy) / fy;
    x[q * 2 + 1] = (v.z - cx) / fx;
    y[q * 2 + 1] = (v.w - cy) / fy;
  }
  Pose p;
  const bool good = posit(sobj, sB, x, y, p);
  hyp[o] = p;
  valid[o] = good ? 1 : 0;
}

// Number of correspondences of vote `v` (8 corners) within `thr2` (squared pixels) of their projection under `p`;
// `mask` (optional) receives one bit per corner.
__device__ __forceinline__ int vote_inliers(const Pose& p, const float* sobj, const float4* v4, float fx, float fy,
                                            float cx, float cy, float thr2, unsigned* mask) {
  int n = 0;
  unsigned m = 0;
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const float4 v = __ldg(v4 + q);
    const float u[2] = {v.x, v.z}, w[2] = {v.y, v.w};
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      const float* X = sobj + (q * 2 + e) * 3;
      const float Y0 = p.r[0] * X[0] + p.r[1] * X[1] + p.r[2] * X[2] + p.t[0];
      const float Y1 = p.r[3] * X[0] + p.r[4] * X[1] + p.r[5] * X[2] + p.t[1];
      const float Y2 = p.r[6] * X[0] + p.r[7] * X[1] + p.r[8] * X[2] + p.t[2];
      const float du = fx * Y0 / Y2 + cx - u[e], dv = fy * Y1 / Y2 + cy - w[e];
      const bool in = Y2 > 0.0f && du * du + dv * dv <= thr2;
      n += in ? 1 : 0;
      m |= in ? (1u << (q * 2 + e)) : 0u;
    }
  }
  if (mask) *mask = m;
  return n;
}

// grid (n_hyp, groups), 128 threads: inlier count of one hypothesis over all votes of its group.
__global__ void __launch_bounds__(128)
pnp_score_kernel(const float* __restrict__ votes, const long long* __restrict__ offsets, const float* __restrict__ obj,
                 const float* __restrict__ cam, int n_hyp, float thr2, const Pose* __restrict__ hyp,
                 const unsigned char* __restrict__ valid, int* __restrict__ score) {
  const int g = blockIdx.y, h = blockIdx.x;
  const long long o = static_cast<long long>(g) * n_hyp + h;
  if (!valid[o]) {
    if (threadIdx.x == 0) score[o] = -1;
    return;
  }
  __shared__ float sobj[24];
  __shared__ Pose sp;
  __shared__ int part[4];
  if (threadIdx.x < 24) sobj[threadIdx.x] = obj[g * 24 + threadIdx.x];
  if (threadIdx.x == 0) sp = hyp[o];
  __syncthreads();
  const Pose p = sp;
  const float fx = cam[g * 4], fy = cam[g * 4 + 1], cx = cam[g * 4 + 2], cy = cam[g * 4 + 3];
  const long long begin = offsets[g], k = offsets[g + 1] - begin;
  int n = 0;
  for (long long v = threadIdx.x; v < k; v += blockDim.x)
    n += vote_inliers(p, sobj, reinterpret_cast<const float4*>(votes + (begin + v) * 16), fx, fy, cx, cy, thr2, nullptr);
#pragma unroll
  for (int s = 16; s > 0; s >>= 1) n += __shfl_xor_sync(0xFFFFFFFFu, n, s);
  if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = n;
  __syncthreads();
  if (threadIdx.x == 0) score[o] = part[0] + part[1] + part[2] + part[3];
}

// ---- refinement helpers (thread 0 of the block)
__device__ void rodrigues_exp(const double* w, double* R) {  // R = exp([w]x)
  const double th = sqrt(w[0] * w[0] + w[1] * w[1] + w[2] * w[2]);
  double a, b;  // R = I + a [w]x + b [w]x^2
  if (th < 1e-8) {
    a = 1.0;
    b = 0.5;
  } else {
    a = sin(th) / th;
    b = (1.0 - cos(th)) / (th * th);
  }
  const double x = w[0], y = w[1], z = w[2];
  R[0] = 1.0 - b * (y * y + z * z); R[1] = -a * z + b * x * y;       R[2] = a * y + b * x * z;
  R[3] = a * z + b * x * y;         R[4] = 1.0 - b * (x * x + z * z); R[5] = -a * x + b * y * z;
  R[6] = -a * y + b * x * z;        R[7] = a * x + b * y * z;         R[8] = 1.0 - b * (x * x + y * y);
}

__device__ void rodrigues_log(const float* R, float* w) {  // rotation vector of a rotation matrix (cv2.Rodrigues)
  const double tr = static_cast<double>(R[0]) + R[4] + R[8];
  double c = (tr - 1.0) * 0.5;
  c = c > 1.0 ? 1.0 : (c < -1.0 ? -1.0 : c);
  const double th = acos(c);
  double v[3] = {static_cast<double>(R[7]) - R[5], static_cast<double>(R[2]) - R[6], static_cast<double>(R[3]) - R[1]};
  const double s = sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]) * 0.5;  // sin(theta)
  if (s > 1e-6) {
    const double f = th / (2.0 * s);
    for (int i = 0; i < 3; ++i) w[i] = static_cast<float>(v[i] * f);
  } else if (c > 0.0) {
    for (int i = 0; i < 3; ++i) w[i] = static_cast<float>(v[i] * 0.5);
  } else {  // theta ~ pi: axis from the diagonal of (R + I) / 2
    double ax[3] = {sqrt(fmax((R[0] + 1.0) * 0.5, 0.0)), sqrt(fmax((R[4] + 1.0) * 0.5, 0.0)),
                    sqrt(fmax((R[8] + 1.0) * 0.5, 0.0))};
    if (R[1] + R[3] < 0.0) ax[1] = -ax[1];
    if (R[2] + R[6] < 0.0) ax[2] = -ax[2];
    for (int i = 0; i < 3; ++i) w[i] = static_cast<float>(ax[i] * th);
  }
}

// Solves (H + lambda * diag(H)) d = -b for the symmetric 6x6 H given as its upper triangle (21 values, row-major).
__device__ bool solve6(const double* Hu, const double* b, double lambda, double* d) {
  double A[6][6];
  int q = 0;
  for (int i = 0; i < 6; ++i)
    for (int j = i; j < 6; ++j) {
      A[i][j] = A[j][i] = Hu[q++];
    }
  for (int i = 0; i < 6; ++i) A[i][i] += lambda * A[i][i] + 1e-12;
  double L[6][6] = {};
  for (int i = 0; i < 6; ++i)
    for (int j = 0; j <= i; ++j) {
      double s = A[i][j];
      for (int k = 0; k < j; ++k) s -= L[i][k] * L[j][k];
      if (i == j) {
        if (!(s > 0.0)) return false;
        L[i][i] = sqrt(s);
      } else {
        L[i][j] = s / L[j][j];
      }
    }
  double y[6];
  for (int i = 0; i < 6; ++i) {
    double s = -b[i];
    for (int k = 0; k < i; ++k) s -= L[i][k] * y[k];
    y[i] = s / L[i][i];
  }
  for (int i = 5; i >= 0; --i) {
    double s = y[i];
    for (int k = i + 1; k < 6; ++k) s -= L[k][i] * d[k];
    d[i] = s / L[i][i];
  }
  return true;
}

constexpr int kRefThreads = 256;
constexpr int kNormal = 28;  // 21 (J^T J upper) + 6 (J^T r) + 1 (sum r^2)

// Block-wide sum of kNormal doubles per thread into `out` (shared, kNormal entries).
__device__ void block_sum_normal(double* acc, double* red /*[8][kNormal]*/, double* out) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int i = 0; i < kNormal; ++i) {
    double v = acc[i];
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, s);
    if (lane == 0) red[warp * kNormal + i] = v;
  }
  __syncthreads();
  if (threadIdx.x < kNormal) {
    double s = 0.0;
    for (int w = 0; w < kRefThreads / 32; ++w) s += red[w * kNormal + threadIdx.x];
    out[threadIdx.x] = s;
  }
  __syncthreads();
}

// Accumulates the normal equations of the reprojection error at pose (R, t) over the inlier correspondences
// (bit q of inl[v] = corner q of vote v). Perturbation: R <- exp([dw]x) R, t <- t + dt.
__device__ void accumulate_normal(const double* R, const double* t, const float* sobj, const float* votes,
                                  const unsigned char* inl, long long begin, long long k, float fx, float fy, float cx,
                                  float cy, double* acc) {
  for (int i = 0; i < kNormal; ++i) acc[i] = 0.0;
  for (long long v = threadIdx.x; v < k; v += kRefThreads) {
    const unsigned m = inl[begin + v];
    if (!m) continue;
    const float* uv = votes + (begin + v) * 16;
    for (int q = 0; q < 8; ++q) {
      if (!(m & (1u << q))) continue;
      const double X0 = sobj[q * 3], X1 = sobj[q * 3 + 1], X2 = sobj[q * 3 + 2];
      const double a0 = R[0] * X0 + R[1] * X1 + R[2] * X2, a1 = R[3] * X0 + R[4] * X1 + R[5] * X2,
                   a2 = R[6] * X0 + R[7] * X1 + R[8] * X2;  // R X
      const double Y0 = a0 + t[0], Y1 = a1 + t[1], Y2 = a2 + t[2];
      const double iz = 1.0 / Y2;
      const double ru = fx * Y0 * iz + cx - uv[q * 2], rv = fy * Y1 * iz + cy - uv[q * 2 + 1];
      // d(u)/dY = [fx/Z, 0, -fx Y0/Z^2], d(v)/dY = [0, fy/Z, -fy Y1/Z^2]; dY = -[RX]x dw + dt
      const double pu[3] = {fx * iz, 0.0, -fx * Y0 * iz * iz}, pv[3] = {0.0, fy * iz, -fy * Y1 * iz * iz};
      double ju[6], jv[6];
      // (p^T * -[a]x) = (a x p)^T
      ju[0] = a1 * pu[2] - a2 * pu[1]; ju[1] = a2 * pu[0] - a0 * pu[2]; ju[2] = a0 * pu[1] - a1 * pu[0];
      jv[0] = a1 * pv[2] - a2 * pv[1]; jv[1] = a2 * pv[0] - a0 * pv[2]; jv[2] = a0 * pv[1] - a1 * pv[0];
      for (int c = 0; c < 3; ++c) {
        ju[3 + c] = pu[c];
        jv[3 + c] = pv[c];
      }
      int e = 0;
      for (int i = 0; i < 6; ++i)
        for (int j = i; j < 6; ++j) acc[e++] += ju[i] * ju[j] + jv[i] * jv[j];
      for (int i = 0; i < 6; ++i) acc[21 + i] += ju[i] * ru + jv[i] * rv;
      acc[27] += ru * ru + rv * rv;
    }
  }
}

// grid (groups), 256 threads.
__global__ void __launch_bounds__(kRefThreads)
pnp_refine_kernel(const float* __restrict__ votes, const long long* __restrict__ offsets, const float* __restrict__ obj,
                  const float* __restrict__ cam, int n_hyp, float thr2, int min_inliers, int max_evals,
                  const Pose* __restrict__ hyp, const int* __restrict__ score, unsigned char* __restrict__ inl,
                  float* __restrict__ out_rt, int* __restrict__ out_inliers, float* __restrict__ out_rmse) {
  const int g = blockIdx.x;
  __shared__ float sobj[24];
  __shared__ int best_s[kRefThreads / 32], best_h[kRefThreads / 32];
  __shared__ double red[(kRefThreads / 32) * kNormal];
  __shared__ double sums[kNormal], cur[kNormal];
  __shared__ double R[9], t[3], Rn[9], tn[3];
  __shared__ int state;  // 0 continue, 1 stop
  __shared__ int n_in;
  if (threadIdx.x < 24) sobj[threadIdx.x] = obj[g * 24 + threadIdx.x];
  // ---- best hypothesis: most inliers, lowest index on ties
  int bs = -1, bh = 0x7fffffff;
  for (int h = threadIdx.x; h < n_hyp; h += kRefThreads) {
    const int s = score[static_cast<long long>(g) * n_hyp + h];
    if (s > bs) {
      bs = s;
      bh = h;
    }
  }
#pragma unroll
  for (int s = 16; s > 0; s >>= 1) {
    const int os = __shfl_xor_sync(0xFFFFFFFFu, bs, s), oh = __shfl_xor_sync(0xFFFFFFFFu, bh, s);
    if (os > bs || (os == bs && oh < bh)) {
      bs = os;
      bh = oh;
    }
  }
  if ((threadIdx.x & 31) == 0) {
    best_s[threadIdx.x >> 5] = bs;
    best_h[threadIdx.x >> 5] = bh;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < kRefThreads / 32; ++w)
      if (best_s[w] > best_s[0] || (best_s[w] == best_s[0] && best_h[w] < best_h[0])) {
        best_s[0] = best_s[w];
        best_h[0] = best_h[w];
      }
    n_in = best_s[0];
    if (n_in >= min_inliers) {
      const Pose p = hyp[static_cast<long long>(g) * n_hyp + best_h[0]];
      for (int i = 0; i < 9; ++i) R[i] = p.r[i];
      for (int i = 0; i < 3; ++i) t[i] = p.t[i];
    }
  }
  __syncthreads();
  if (n_in < min_inliers) {  // cv2.solvePnPRansac returns retval = False: no pose
    if (threadIdx.x < 6) out_rt[g * 6 + threadIdx.x] = 0.0f;
    if (threadIdx.x == 0) {
      out_inliers[g] = 0;
      if (out_rmse) out_rmse[g] = 0.0f;
    }
    return;
  }
  const float fx = cam[g * 4], fy = cam[g * 4 + 1], cx = cam[g * 4 + 2], cy = cam[g * 4 + 3];
  const long long begin = offsets[g], k = offsets[g + 1] - begin;
  {  // inlier set of the best model (fixed during the refinement, as OpenCV's final solvePnP on the inliers)
    Pose p;
    for (int i = 0; i < 9; ++i) p.r[i] = static_cast<float>(R[i]);
    for (int i = 0; i < 3; ++i) p.t[i] = static_cast<float>(t[i]);
    for (long long v = threadIdx.x; v < k; v += kRefThreads) {
      unsigned m;
      vote_inliers(p, sobj, reinterpret_cast<const float4*>(votes + (begin + v) * 16), fx, fy, cx, cy, thr2, &m);
      inl[begin + v] = static_cast<unsigned char>(m);
    }
  }
  __syncthreads();
  double acc[kNormal];
  accumulate_normal(R, t, sobj, votes, inl, begin, k, fx, fy, cx, cy, acc);
  block_sum_normal(acc, red, cur);
  __shared__ double lambda;
  if (threadIdx.x == 0) {
    lambda = 1e-4;
    state = 0;
  }
  __syncthreads();
  for (int ev = 0; ev < max_evals; ++ev) {
    if (threadIdx.x == 0) {
      double d[6];
      if (!solve6(cur, cur + 21, lambda, d)) {
        lambda *= 10.0;
        for (int i = 0; i < 9; ++i) Rn[i] = R[i];
        for (int i = 0; i < 3; ++i) tn[i] = t[i];
      } else {
        double E[9];
        rodrigues_exp(d, E);
        for (int r = 0; r < 3; ++r)
          for (int c = 0; c < 3; ++c) Rn[r * 3 + c] = E[r * 3] * R[c] + E[r * 3 + 1] * R[3 + c] + E[r * 3 + 2] * R[6 + c];
        for (int i = 0; i < 3; ++i) tn[i] = t[i] + d[3 + i];
        const double step = fabs(d[0]) + fabs(d[1]) + fabs(d[2]) +
                            (fabs(d[3]) + fabs(d[4]) + fabs(d[5])) / (fabs(t[2]) + 1e-9);
        if (step < 1e-9) state = 1;
      }
    }
    __syncthreads();
    if (state) break;
    accumulate_normal(Rn, tn, sobj, votes, inl, begin, k, fx, fy, cx, cy, acc);
    block_sum_normal(acc, red, sums);
    if (threadIdx.x == 0) {
      if (sums[27] < cur[27] && tn[2] + 0.0 == tn[2]) {  // accept
        const double gain = (cur[27] - sums[27]) / (cur[27] + 1e-30);
        for (int i = 0; i < 9; ++i) R[i] = Rn[i];
        for (int i = 0; i < 3; ++i) t[i] = tn[i];
        for (int i = 0; i < kNormal; ++i) cur[i] = sums[i];
        lambda = fmax(lambda * 0.1, 1e-12);
        if (gain < 1e-12) state = 1;
      } else {
        lambda *= 10.0;
        if (lambda > 1e8) state = 1;
      }
    }
    __syncthreads();
    if (state) break;
  }
  if (threadIdx.x == 0) {
    float Rf[9], w[3];
    for (int i = 0; i < 9; ++i) Rf[i] = static_cast<float>(R[i]);
    rodrigues_log(Rf, w);
    for (int i = 0; i < 3; ++i) {
      out_rt[g * 6 + i] = w[i];
      out_rt[g * 6 + 3 + i] = static_cast<float>(t[i]);
    }
    out_inliers[g] = n_in;
    if (out_rmse) out_rmse[g] = static_cast<float>(sqrt(cur[27] / (n_in > 0 ? n_in : 1)));
  }
}

}  // namespace
}  // namespace ppb

using namespace ppb;

namespace {
size_t align256(size_t v) { return (v + 255) & ~static_cast<size_t>(255); }
}  // namespace

extern "C" {

size_t ppb_pnp_workspace_bytes(int groups, long long total_votes, int iterations) {
  if (groups <= 0 || iterations <= 0 || total_votes < 0) return 0;
  const size_t gh = static_cast<size_t>(groups) * iterations;
  return align256(gh * sizeof(Pose)) + align256(gh) + align256(gh * sizeof(int)) +
         align256(static_cast<size_t>(total_votes) + 1);
}

int ppb_pnp_ransac_votes(const float* votes, const long long* offsets, const float* obj_points, const float* cam,
                         int groups, long long total_votes, int iterations, float reprojection_error, int min_inliers,
                         float* out_rt, int* out_inliers, float* out_rmse, void* workspace, size_t workspace_bytes,
                         void* stream) {
  PPB_REQUIRE(groups >= 0 && total_votes >= 0, "ppb_pnp_ransac_votes: bad sizes groups=%d votes=%lld", groups, total_votes);
  if (groups == 0) return kOk;
  PPB_REQUIRE(iterations >= 1 && iterations <= 65535, "ppb_pnp_ransac_votes: iterations must be in [1, 65535], got %d",
              iterations);
  PPB_REQUIRE(groups <= 65535, "ppb_pnp_ransac_votes: at most 65535 groups per call, got %d", groups);
  PPB_REQUIRE(reprojection_error > 0.0f, "ppb_pnp_ransac_votes: reprojection_error must be positive");
  PPB_REQUIRE(offsets && obj_points && cam && out_rt && out_inliers && (total_votes == 0 || votes),
              "ppb_pnp_ransac_votes: null buffer");
  PPB_REQUIRE((reinterpret_cast<uintptr_t>(votes) & 15) == 0, "ppb_pnp_ransac_votes: votes must be 16-byte aligned");
  PPB_REQUIRE(workspace != nullptr && workspace_bytes >= ppb_pnp_workspace_bytes(groups, total_votes, iterations),
              "ppb_pnp_ransac_votes: workspace too small");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const size_t gh = static_cast<size_t>(groups) * iterations;
  unsigned char* base = static_cast<unsigned char*>(workspace);
  Pose* hyp = reinterpret_cast<Pose*>(base);
  base += align256(gh * sizeof(Pose));
  unsigned char* valid = base;
  base += align256(gh);
  int* score = reinterpret_cast<int*>(base);
  base += align256(gh * sizeof(int));
  unsigned char* inl = base;
  const float thr2 = reprojection_error * reprojection_error;
  pnp_hypotheses_kernel<<<dim3((iterations + 127) / 128, groups), 128, 0, st>>>(votes, offsets, obj_points, cam,
                                                                                iterations, hyp, valid);
  PPB_CUDA(cudaGetLastError());
  pnp_score_kernel<<<dim3(iterations, groups), 128, 0, st>>>(votes, offsets, obj_points, cam, iterations, thr2, hyp,
                                                             valid, score);
  PPB_CUDA(cudaGetLastError());
  pnp_refine_kernel<<<groups, kRefThreads, 0, st>>>(votes, offsets, obj_points, cam, iterations, thr2,
                                                    min_inliers < 1 ? 1 : min_inliers, 40, hyp, score, inl, out_rt,
                                                    out_inliers, out_rmse);
  PPB_CUDA(cudaGetLastError());
  count_launch(3);
  return kOk;
}

}  // extern "C"
