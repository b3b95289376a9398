// Host-side plumbing shared by every translation unit: error text, launch
// counter, TMA descriptor cache.
#pragma once

#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <cstdarg>
#include <cstdio>

namespace dtb {

void set_error(const char* fmt, ...);
extern std::atomic<long long> g_launch_count;

inline void count_launch(int n = 1) { g_launch_count.fetch_add(n, std::memory_order_relaxed); }

inline int sm_count() {
  static int cached[16] = {0};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 16) dev = 0;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

// Builds (or fetches from a small cache) a rank-4 tiled tensor map with 128-byte
// swizzle.  dims / box are innermost-first; strides_bytes has 3 entries (dims 1..3).
int get_tensor_map_bf16(CUtensorMap* out, const void* base, const uint64_t dims[4], const uint64_t strides_bytes[3],
                        const uint32_t box[4]);

}  // namespace dtb

#define DTB_CUDA_OK(expr)                                                                         \
  do {                                                                                            \
    cudaError_t _e = (expr);                                                                      \
    if (_e != cudaSuccess) {                                                                      \
      dtb::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
      return 1;                                                                                   \
    }                                                                                             \
  } while (0)

#define DTB_LAUNCH_OK()                                                                                 \
  do {                                                                                                  \
    cudaError_t _e = cudaGetLastError();                                                                \
    if (_e != cudaSuccess) {                                                                            \
      dtb::set_error("kernel launch failed: %s (%s:%d)", cudaGetErrorString(_e), __FILE__, __LINE__);   \
      return 1;                                                                                         \
    }                                                                                                   \
    dtb::count_launch();                                                                                \
  } while (0)

#define DTB_REQUIRE(cond, ...)      \
  do {                              \
    if (!(cond)) {                  \
      dtb::set_error(__VA_ARGS__);  \
      return 2;                     \
    }                               \
  } while (0)
