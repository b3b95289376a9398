// Library-wide host state: last-error text, launch counter, TMA descriptor cache.
#include "host.h"

#include <cstring>
#include <mutex>
#include <unordered_map>
#include <vector>

#include "../../include/dtb200.h"

namespace dtb {

static thread_local char g_error[1024] = "";
std::atomic<long long> g_launch_count{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_error, sizeof(g_error), fmt, ap);
  va_end(ap);
}

// cuTensorMapEncodeTiled is fetched through the runtime so the library does not
// link against libcuda (it has to load on the GPU-less build box too).
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess) {
      fn = reinterpret_cast<EncodeTiledFn>(p);
    }
  });
  return fn;
}

struct MapKey {
  uint64_t v[12];
  bool operator==(const MapKey& o) const { return memcmp(v, o.v, sizeof(v)) == 0; }
};
struct MapKeyHash {
  size_t operator()(const MapKey& k) const {
    uint64_t h = 1469598103934665603ull;
    for (uint64_t x : k.v) {
      h ^= x;
      h *= 1099511628211ull;
    }
    return static_cast<size_t>(h);
  }
};

int get_tensor_map_bf16(CUtensorMap* out, const void* base, const uint64_t dims[4], const uint64_t strides_bytes[3],
                        const uint32_t box[4]) {
  static std::mutex mu;
  static std::unordered_map<MapKey, CUtensorMap, MapKeyHash> cache;
  MapKey key;
  key.v[0] = reinterpret_cast<uint64_t>(base);
  for (int i = 0; i < 4; ++i) key.v[1 + i] = dims[i];
  for (int i = 0; i < 3; ++i) key.v[5 + i] = strides_bytes[i];
  for (int i = 0; i < 4; ++i) key.v[8 + i] = box[i];
  {
    std::lock_guard<std::mutex> lock(mu);
    auto it = cache.find(key);
    if (it != cache.end()) {
      *out = it->second;
      return 0;
    }
  }
  EncodeTiledFn fn = encode_fn();
  if (!fn) {
    set_error("cuTensorMapEncodeTiled is not available from this driver");
    return 1;
  }
  cuuint64_t gdims[4] = {dims[0], dims[1], dims[2], dims[3]};
  cuuint64_t gstr[3] = {strides_bytes[0], strides_bytes[1], strides_bytes[2]};
  cuuint32_t gbox[4] = {box[0], box[1], box[2], box[3]};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), gdims, gstr, gbox, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error(
        "cuTensorMapEncodeTiled failed (%d): base=%p dims=[%llu,%llu,%llu,%llu] strides=[%llu,%llu,%llu] "
        "box=[%u,%u,%u,%u]",
        static_cast<int>(r), base, (unsigned long long)dims[0], (unsigned long long)dims[1],
        (unsigned long long)dims[2], (unsigned long long)dims[3], (unsigned long long)strides_bytes[0],
        (unsigned long long)strides_bytes[1], (unsigned long long)strides_bytes[2], box[0], box[1], box[2], box[3]);
    return 1;
  }
  std::lock_guard<std::mutex> lock(mu);
  if (cache.size() > 4096) cache.clear();
  cache.emplace(key, *out);
  return 0;
}

}  // namespace dtb

extern "C" {
const char* dtb_last_error(void) { return dtb::g_error; }
int dtb_version(void) { return 100; }
long long dtb_launch_count(void) { return dtb::g_launch_count.load(); }
void dtb_reset_launch_count(void) { dtb::g_launch_count.store(0); }
}
