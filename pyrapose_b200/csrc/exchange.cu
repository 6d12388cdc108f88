// Peer-to-peer gradient exchange for the data-parallel training step (the reference's
// keras.utils.multi_gpu_model / fit_generator step, PyraPose/bin/train.py:82-102, 381-390).
//
// An alternative to the overlapped NCCL all-reduce that needs no SM while the backward pass runs (NCCL's channels are
// CTAs, and every kernel of the step is a persistent grid of one CTA per SM): as soon as a gradient bucket is complete,
// every replica pushes it into a staging slot in EACH peer's memory with plain asynchronous copies (copy engines,
// NVLink / NVSwitch, peer mappings from CUDA IPC handles), and after the backward pass one bandwidth-bound kernel sums
// the N slices in rank order — the same order on every replica, so the replicas stay bit-identical.
// Measured r02 (config 4): the same step time as NCCL at 2 GPUs (13.30-13.42 vs 13.31 ms), slower at 8 (15.08 vs 13.63
// ms) — an all-gather by copies moves N-1 times the gradient (1.19 GB out and in per replica and step at N = 8), which
// the copy engines do not finish inside the backward pass. It is therefore an option (compile(exchange='p2p')), not the
// default.
#include "../../include/pyrapose_b200.h"
#include "errors.h"
#include "launch_count.h"

#include <cuda_runtime.h>
#include <string.h>

namespace ppb {
namespace {

constexpr int kMaxReplicas = 16;

struct SumSrcs {
  const float* p[kMaxReplicas];
};

// out[i] = ((s0[i] + s1[i]) + s2[i]) + ... in rank order; out may alias one of the sources (every element is read
// completely before it is written by the same thread).
__global__ void __launch_bounds__(256) sum_replicas_kernel(float* __restrict__ out, SumSrcs srcs, int n, long long count4,
                                                           long long count) {
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < count4; i += stride) {
    float4 v[kMaxReplicas];
#pragma unroll
    for (int r = 0; r < kMaxReplicas; ++r)
      if (r < n) v[r] = reinterpret_cast<const float4*>(srcs.p[r])[i];
    float4 a = v[0];
#pragma unroll
    for (int r = 1; r < kMaxReplicas; ++r)
      if (r < n) {
        a.x += v[r].x;
        a.y += v[r].y;
        a.z += v[r].z;
        a.w += v[r].w;
      }
    reinterpret_cast<float4*>(out)[i] = a;
  }
  // tail (count not a multiple of 4)
  for (long long i = count4 * 4 + static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < count; i += stride) {
    float a = srcs.p[0][i];
    for (int r = 1; r < n; ++r) a += srcs.p[r][i];
    out[i] = a;
  }
}

}  // namespace
}  // namespace ppb

using namespace ppb;

extern "C" {

int ppb_p2p_alloc(size_t bytes, void** dev_ptr) {
  PPB_REQUIRE(dev_ptr != nullptr && bytes > 0, "ppb_p2p_alloc: bad argument");
  PPB_CUDA(cudaMalloc(dev_ptr, bytes));
  PPB_CUDA(cudaMemset(*dev_ptr, 0, bytes));
  return kOk;
}

int ppb_p2p_free(void* dev_ptr) {
  if (dev_ptr != nullptr) PPB_CUDA(cudaFree(dev_ptr));
  return kOk;
}

int ppb_p2p_export(const void* dev_ptr, unsigned char* handle64) {
  PPB_REQUIRE(dev_ptr != nullptr && handle64 != nullptr, "ppb_p2p_export: null argument");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "CUDA IPC handles are 64 bytes");
  cudaIpcMemHandle_t h;
  PPB_CUDA(cudaIpcGetMemHandle(&h, const_cast<void*>(dev_ptr)));
  memcpy(handle64, &h, 64);
  return kOk;
}

int ppb_p2p_import(const unsigned char* handle64, void** peer_ptr) {
  PPB_REQUIRE(handle64 != nullptr && peer_ptr != nullptr, "ppb_p2p_import: null argument");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  PPB_CUDA(cudaIpcOpenMemHandle(peer_ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return kOk;
}

int ppb_p2p_release(void* peer_ptr) {
  if (peer_ptr != nullptr) PPB_CUDA(cudaIpcCloseMemHandle(peer_ptr));
  return kOk;
}

int ppb_p2p_copy(void* dst, const void* src, size_t bytes, void* stream) {
  PPB_REQUIRE(dst != nullptr && src != nullptr, "ppb_p2p_copy: null buffer");
  if (bytes == 0) return kOk;
  PPB_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDefault, static_cast<cudaStream_t>(stream)));
  return kOk;
}

int ppb_p2p_sum_f32(float* out, const float* const* srcs_host, int n_srcs, long long count, void* stream) {
  PPB_REQUIRE(out != nullptr && srcs_host != nullptr && n_srcs >= 1 && n_srcs <= kMaxReplicas && count >= 0,
              "ppb_p2p_sum_f32: bad argument (1..%d sources)", kMaxReplicas);
  if (count == 0) return kOk;
  SumSrcs s;
  for (int r = 0; r < kMaxReplicas; ++r) s.p[r] = r < n_srcs ? srcs_host[r] : nullptr;
  for (int r = 0; r < n_srcs; ++r)
    PPB_REQUIRE(s.p[r] != nullptr && (reinterpret_cast<uintptr_t>(s.p[r]) & 15) == 0,
                "ppb_p2p_sum_f32: source %d must be a 16-byte aligned device pointer", r);
  PPB_REQUIRE((reinterpret_cast<uintptr_t>(out) & 15) == 0, "ppb_p2p_sum_f32: out must be 16-byte aligned");
  const long long count4 = count / 4;
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  long long blocks = (count4 + 255) / 256;
  if (blocks > 8ll * sms) blocks = 8ll * sms;  // grid-stride: a multiple of the SM count
  if (blocks < 1) blocks = 1;
  sum_replicas_kernel<<<static_cast<unsigned>(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(out, s, n_srcs, count4,
                                                                                                  count);
  PPB_CUDA(cudaGetLastError());
  count_launch();
  return kOk;
}

}  // extern "C"
