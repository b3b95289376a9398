// fp32 SIMT Linear for the reference's own toy shapes (Linear(20,1), Linear(10,1),
// batch 32): far below one tensor-core tile, so a warp-per-output dot product is the
// right tool.  fwd: y = x W^T + b.  bwd: dx = dy W, dW = dy^T x, db = sum dy.
// Reference call sites: /root/reference/src/distributed_trainer.py:162,164,199.
#include <cuda_runtime.h>

#include "../../include/dtb200.h"
#include "../common/host.h"

namespace dtb {
namespace small {

__global__ void __launch_bounds__(128)
fwd_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ b, float* __restrict__ y,
           int rows, int in_f, int out_f) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= rows * out_f) return;
  const int r = warp / out_f, o = warp % out_f;
  float acc = 0.f;
  for (int k = lane; k < in_f; k += 32) acc += x[static_cast<long long>(r) * in_f + k] * w[static_cast<long long>(o) * in_f + k];
#pragma unroll
  for (int s = 16; s > 0; s >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, s);
  if (lane == 0) y[static_cast<long long>(r) * out_f + o] = acc + (b ? b[o] : 0.f);
}

__global__ void __launch_bounds__(128)
bwd_dx_kernel(const float* __restrict__ w, const float* __restrict__ dy, float* __restrict__ dx, int rows, int in_f,
              int out_f) {
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= static_cast<long long>(rows) * in_f) return;
  const int r = static_cast<int>(i / in_f), k = static_cast<int>(i % in_f);
  float acc = 0.f;
  for (int o = 0; o < out_f; ++o) acc += dy[static_cast<long long>(r) * out_f + o] * w[static_cast<long long>(o) * in_f + k];
  dx[i] = acc;
}

// one warp per (o, k) weight element (and one per bias element): reduce over rows
__global__ void __launch_bounds__(128)
bwd_dw_kernel(const float* __restrict__ x, const float* __restrict__ dy, float* __restrict__ dw, float* __restrict__ db,
              int rows, int in_f, int out_f) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  const int n_w = out_f * in_f;
  if (warp >= n_w + out_f) return;
  float acc = 0.f;
  if (warp < n_w) {
    const int o = warp / in_f, k = warp % in_f;
    for (int r = lane; r < rows; r += 32) acc += dy[static_cast<long long>(r) * out_f + o] * x[static_cast<long long>(r) * in_f + k];
  } else {
    const int o = warp - n_w;
    for (int r = lane; r < rows; r += 32) acc += dy[static_cast<long long>(r) * out_f + o];
  }
#pragma unroll
  for (int s = 16; s > 0; s >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, s);
  if (lane == 0) {
    if (warp < n_w) dw[warp] = acc;
    else if (db) db[warp - n_w] = acc;
  }
}

}  // namespace small
}  // namespace dtb

using namespace dtb;

extern "C" int dtb_linear_small_fwd(const float* x, const float* w, const float* b, float* y, int rows, int in_f,
                                    int out_f, void* stream) {
  DTB_REQUIRE(rows > 0 && in_f > 0 && out_f > 0, "dtb_linear_small_fwd: empty problem");
  const long long warps = static_cast<long long>(rows) * out_f;
  small::fwd_kernel<<<static_cast<int>((warps + 3) / 4), 128, 0, static_cast<cudaStream_t>(stream)>>>(x, w, b, y, rows,
                                                                                                     in_f, out_f);
  DTB_LAUNCH_OK();
  return 0;
}

extern "C" int dtb_linear_small_bwd(const float* x, const float* w, const float* dy, float* dx, float* dw, float* db,
                                    int rows, int in_f, int out_f, void* stream) {
  DTB_REQUIRE(rows > 0 && in_f > 0 && out_f > 0, "dtb_linear_small_bwd: empty problem");
  auto s = static_cast<cudaStream_t>(stream);
  if (dx != nullptr) {
    const long long n = static_cast<long long>(rows) * in_f;
    small::bwd_dx_kernel<<<static_cast<int>((n + 127) / 128), 128, 0, s>>>(w, dy, dx, rows, in_f, out_f);
    DTB_LAUNCH_OK();
  }
  if (dw != nullptr) {
    const long long warps = static_cast<long long>(out_f) * in_f + out_f;
    small::bwd_dw_kernel<<<static_cast<int>((warps + 3) / 4), 128, 0, s>>>(x, dy, dw, db, rows, in_f, out_f);
    DTB_LAUNCH_OK();
  }
  return 0;
}
