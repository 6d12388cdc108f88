/* pyrapose_b200 — C-ABI of the B200 (sm_100a) implementation of PyraPose's data-parallel hot path.
 *
 * Conventions
 *  - Every function returns 0 on success or a negative ppb_status; ppb_last_error() then holds a
 *    message (thread-local). The Python host layer raises ValueError for PPB_ERR_INVALID and
 *    RuntimeError otherwise.
 *  - All buffers are caller-owned DEVICE pointers unless a parameter name ends in _host.
 *  - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream). Calls are
 *    asynchronous with respect to the host unless stated otherwise.
 *  - No entry point has a CPU fallback: without a CUDA device they fail with PPB_ERR_CUDA.
 *
 * Each declaration cites the reference interface (file:line under sThalham/PyraPose) it replaces.
 */
#ifndef PYRAPOSE_B200_H_
#define PYRAPOSE_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PPB_API __attribute__((visibility("default")))

typedef enum ppb_status {
  PPB_OK = 0,
  PPB_ERR_INVALID = -1,
  PPB_ERR_CUDA = -2,
  PPB_ERR_FAULT = -3,
  PPB_ERR_STATE = -4
} ppb_status;

/* ------------------------------------------------------------------ library */
PPB_API int ppb_version(void);
PPB_API const char* ppb_last_error(void);
/* Number of kernels this library has launched in the calling process (bench.py's gpu_launches). */
PPB_API long long ppb_launch_count(void);

/* ------------------------------------------------------------------ convolution (tcgen05 implicit GEMM)
 * Replaces keras.layers.Conv2D(+BatchNormalization(freeze)+Activation('relu')+Add) as used by
 * PyraPose/models/retinanet.py:29-46,73-88,118-126,192-212 and keras-resnet's bottleneck_2d
 * (PyraPose/models/resnet.py:87).
 *
 * x        : bf16 [batch, h+2, w+2, cin]   "padded-flat": one-pixel zero border around every image
 * weight   : bf16 [cout_pad, ksize*ksize*cin], K index = (r*ksize + s)*cin + c  (Keras HWIO -> O,HWI);
 *            cout_pad = cout rounded up to a multiple of 16, padding rows zero
 * bias     : f32 [cout] or NULL (frozen BatchNorm is folded into weight and bias by the caller)
 * residual : bf16 [batch, h+2, w+2, cout] or NULL, added before the ReLU
 * y        : bf16 [batch, h+2, w+2, cout], border written as zeros
 * ksize    : 1 or 3; stride 1; 'same' zero padding. cin % 64 == 0, cout % 16 == 0.
 */
PPB_API int ppb_conv2d_bf16(const void* x, const void* weight, const float* bias, const void* residual, void* y,
                            int batch, int h, int w, int cin, int cout, int ksize, int relu, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* PYRAPOSE_B200_H_ */
