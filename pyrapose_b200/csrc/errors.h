// Error convention of the C-ABI: every entry point returns 0 or a negative ppb_status and leaves a
// human-readable message in a thread-local buffer (ppb_last_error). The Python layer maps
// PPB_ERR_INVALID to ValueError and everything else to RuntimeError.
#pragma once
#include <cuda_runtime.h>
#include <cstdarg>
#include <cstdio>

namespace ppb {

enum Status : int {
  kOk = 0,
  kErrInvalid = -1,  // bad shape / dtype / argument
  kErrCuda = -2,     // a CUDA runtime or driver call failed
  kErrFault = -3,    // a kernel trapped on its own watchdog
  kErrState = -4,    // handle used in the wrong state (e.g. forward before weights are loaded)
};

char* last_error_buffer();
constexpr int kLastErrorSize = 512;

inline int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(last_error_buffer(), kLastErrorSize, fmt, ap);
  va_end(ap);
  return code;
}

#define PPB_CUDA(expr)                                                                             \
  do {                                                                                             \
    cudaError_t err__ = (expr);                                                                    \
    if (err__ != cudaSuccess)                                                                      \
      return ::ppb::fail(::ppb::kErrCuda, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(err__), \
                         __FILE__, __LINE__);                                                      \
  } while (0)

#define PPB_REQUIRE(cond, ...)                                         \
  do {                                                                 \
    if (!(cond)) return ::ppb::fail(::ppb::kErrInvalid, __VA_ARGS__);  \
  } while (0)

}  // namespace ppb
