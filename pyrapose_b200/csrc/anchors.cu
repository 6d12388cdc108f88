// Anchor grids on the GPU: out[(y*fw + x)*A + a][j] = base[a][j] + shift, shift = (x|y + 0.5) * stride.
// Replaces PyraPose/utils/anchors.py:415-444 (`shift`, float64) + :402-412 (level concatenation) and the
// float32 graph version PyraPose/backend/common.py:93-116 used by layers/_misc.py:24-93 (`Anchors`).
// The nine base anchors per level are computed by the Python host exactly as generate_anchors does
// (anchors.py:447-478); the add here is a single correctly-rounded operation, hence bit-exact.
#include "../../include/pyrapose_b200.h"
#include "errors.h"
#include "launch_count.h"

namespace ppb {
namespace {

template <typename T>
__global__ void __launch_bounds__(256) anchors_shift_kernel(const T* __restrict__ base, int num_base, int fh, int fw,
                                                            T stride, T* __restrict__ out) {
  // one thread per (location, base anchor); 4 coordinates written as one vector
  const long long total = static_cast<long long>(fh) * fw * num_base;
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const int a = static_cast<int>(i % num_base);
  const long long loc = i / num_base;
  const int x = static_cast<int>(loc % fw);
  const int y = static_cast<int>(loc / fw);
  const T sx = (static_cast<T>(x) + static_cast<T>(0.5)) * stride;
  const T sy = (static_cast<T>(y) + static_cast<T>(0.5)) * stride;
  T* o = out + i * 4;
  o[0] = base[a * 4 + 0] + sx;
  o[1] = base[a * 4 + 1] + sy;
  o[2] = base[a * 4 + 2] + sx;
  o[3] = base[a * 4 + 3] + sy;
}

template <typename T>
int launch_level(const T* base, int num_base, int fh, int fw, double stride, T* out, void* stream) {
  PPB_REQUIRE(num_base > 0 && fh >= 0 && fw >= 0, "ppb_anchors_shift: bad sizes A=%d fh=%d fw=%d", num_base, fh, fw);
  const long long total = static_cast<long long>(fh) * fw * num_base;
  if (total == 0) return kOk;
  PPB_REQUIRE(base != nullptr && out != nullptr, "ppb_anchors_shift: null buffer");
  anchors_shift_kernel<T><<<static_cast<unsigned>((total + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      base, num_base, fh, fw, static_cast<T>(stride), out);
  PPB_CUDA(cudaGetLastError());
  count_launch();
  return kOk;
}

}  // namespace
}  // namespace ppb

extern "C" {

int ppb_anchors_shift_f64(const double* base, int num_base, int fh, int fw, double stride, double* out, void* stream) {
  return ppb::launch_level<double>(base, num_base, fh, fw, stride, out, stream);
}

int ppb_anchors_shift_f32(const float* base, int num_base, int fh, int fw, float stride, float* out, void* stream) {
  return ppb::launch_level<float>(base, num_base, fh, fw, stride, out, stream);
}

}  // extern "C"
