// Counts kernel launches made by this library (reported as bench.py's gpu_launches).
#pragma once
#include <atomic>
#include <cuda_runtime.h>
namespace ppb {
extern std::atomic<long long> g_launches;
inline void count_launch(int n = 1) { g_launches.fetch_add(n, std::memory_order_relaxed); }

// Per-device one-time setup (cudaFuncSetAttribute is a per-device attribute: a process that drives several GPUs
// through several engine handles must opt every device in). Returns true exactly once per (flag set, current device);
// the caller performs the setup when it does. Thread-safe.
struct PerDeviceOnce {
  std::atomic<unsigned long long> done{0};
  bool first_on_current_device() {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev > 63) return true;  // set it again: harmless
    const unsigned long long bit = 1ull << dev;
    return (done.fetch_or(bit, std::memory_order_acq_rel) & bit) == 0;
  }
};
}  // namespace ppb
