// Counts kernel launches made by this library (reported as bench.py's gpu_launches).
#pragma once
#include <atomic>
namespace ppb {
extern std::atomic<long long> g_launches;
inline void count_launch(int n = 1) { g_launches.fetch_add(n, std::memory_order_relaxed); }
}  // namespace ppb
