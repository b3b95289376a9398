// Fused losses.
//
//  * cross_entropy (index targets): one CTA per row; the bf16 logits row is read
//    once from HBM into shared memory while an online (max, sum-exp) pair is
//    built, then dlogits = (softmax - onehot) * scale is written in place from
//    the shared copy.  Vocabulary rows up to ~100 KB (50304 bf16) fit.
//  * soft-target cross entropy (float targets of the logits' shape): the form the
//    reference trainer actually calls, F.cross_entropy(output, targets) with
//    float targets (/root/reference/src/distributed_trainer.py:163).
//  * MSE (nn.MSELoss, /root/reference/src/playground/ddp_script.py:135).
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <float.h>

#include "../../include/dtb200.h"
#include "../common/host.h"

namespace dtb {
namespace loss {

constexpr int CE_THREADS = 512;

__device__ __forceinline__ void online_merge(float& m, float& s, float m2, float s2) {
  const float nm = fmaxf(m, m2);
  s = s * __expf(m - nm) + s2 * __expf(m2 - nm);
  m = nm;
}

__global__ void __launch_bounds__(CE_THREADS)
ce_kernel(__nv_bfloat16* __restrict__ logits, const long long* __restrict__ targets, float* __restrict__ row_loss,
          int vocab, long long ld, float grad_scale, int write_grad) {
  extern __shared__ __align__(16) uint8_t ce_smem[];
  __nv_bfloat16* srow = reinterpret_cast<__nv_bfloat16*>(ce_smem);
  __shared__ float red_m[CE_THREADS / 32], red_s[CE_THREADS / 32];
  __shared__ float bc_m, bc_s;
  const int row = blockIdx.x;
  __nv_bfloat16* grow = logits + static_cast<long long>(row) * ld;
  const int nvec = vocab >> 3;
  float m = -FLT_MAX, s = 0.f;
  for (int v = threadIdx.x; v < nvec; v += CE_THREADS) {
    uint4 raw = *reinterpret_cast<const uint4*>(grow + v * 8);
    *reinterpret_cast<uint4*>(srow + v * 8) = raw;
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&raw);
    float f[8];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float2 t = __bfloat1622float2(h[i]);
      f[2 * i] = t.x;
      f[2 * i + 1] = t.y;
    }
    float lm = f[0];
#pragma unroll
    for (int j = 1; j < 8; ++j) lm = fmaxf(lm, f[j]);
    float ls = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) ls += __expf(f[j] - lm);
    online_merge(m, s, lm, ls);
  }
  for (int c = nvec * 8 + threadIdx.x; c < vocab; c += CE_THREADS) {   // ragged tail (vocab 50257)
    const __nv_bfloat16 x = grow[c];
    srow[c] = x;
    online_merge(m, s, __bfloat162float(x), 1.f);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float m2 = __shfl_xor_sync(0xffffffffu, m, o);
    const float s2 = __shfl_xor_sync(0xffffffffu, s, o);
    online_merge(m, s, m2, s2);
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) {
    red_m[warp] = m;
    red_s[warp] = s;
  }
  __syncthreads();
  if (warp == 0) {
    m = lane < CE_THREADS / 32 ? red_m[lane] : -FLT_MAX;
    s = lane < CE_THREADS / 32 ? red_s[lane] : 0.f;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float m2 = __shfl_xor_sync(0xffffffffu, m, o);
      const float s2 = __shfl_xor_sync(0xffffffffu, s, o);
      online_merge(m, s, m2, s2);
    }
    if (lane == 0) {
      bc_m = m;
      bc_s = s;
    }
  }
  __syncthreads();
  m = bc_m;
  s = bc_s;
  const int tgt = static_cast<int>(targets[row]);
  if (threadIdx.x == 0) {
    const float lse = m + __logf(s);
    const float xt = (tgt >= 0 && tgt < vocab) ? __bfloat162float(srow[tgt]) : lse;   // ignore_index < 0
    row_loss[row] = lse - xt;
  }
  if (!write_grad) return;
  const float inv = grad_scale / s;
  const bool ignore = (tgt < 0 || tgt >= vocab);
  const int padded = static_cast<int>(ld);   // padding columns [vocab, ld) get zero gradient
  for (int v = threadIdx.x; v < (padded >> 3); v += CE_THREADS) {
    uint4 raw = *reinterpret_cast<const uint4*>(srow + v * 8);
    __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&raw);
    float f[8];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float2 t = __bfloat1622float2(h[i]);
      f[2 * i] = t.x;
      f[2 * i + 1] = t.y;
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int c = v * 8 + j;
      float g = (c < vocab && !ignore) ? __expf(f[j] - m) * inv : 0.f;
      if (c == tgt) g -= grad_scale;
      f[j] = g;
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
    *reinterpret_cast<uint4*>(grow + v * 8) = raw;
  }
}

// Soft-label CE over small class counts: one warp per row, fp32.
__global__ void __launch_bounds__(128)
soft_ce_kernel(const float* __restrict__ logits, const float* __restrict__ targets, float* __restrict__ row_loss,
               float* __restrict__ dlogits, int rows, int cols, float grad_scale) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= rows) return;
  const float* x = logits + static_cast<long long>(warp) * cols;
  const float* t = targets + static_cast<long long>(warp) * cols;
  float m = -FLT_MAX;
  for (int c = lane; c < cols; c += 32) m = fmaxf(m, x[c]);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
  float s = 0.f, tsum = 0.f;
  for (int c = lane; c < cols; c += 32) {
    s += __expf(x[c] - m);
    tsum += t[c];
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    s += __shfl_xor_sync(0xffffffffu, s, o);
    tsum += __shfl_xor_sync(0xffffffffu, tsum, o);
  }
  const float lse = m + __logf(s);
  // -sum_c t_c * log_softmax(x)_c, term by term: with a single class every term is exactly 0
  float loss = 0.f;
  for (int c = lane; c < cols; c += 32) loss += t[c] * (lse - x[c]);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) loss += __shfl_xor_sync(0xffffffffu, loss, o);
  if (lane == 0) row_loss[warp] = loss;
  if (dlogits != nullptr) {
    float* d = dlogits + static_cast<long long>(warp) * cols;
    for (int c = lane; c < cols; c += 32) d[c] = (__expf(x[c] - lse) * tsum - t[c]) * grad_scale;
  }
}

__global__ void __launch_bounds__(256)
mse_kernel(const float* __restrict__ pred, const float* __restrict__ target, float* __restrict__ loss_sum,
           float* __restrict__ dpred, long long n, float grad_scale) {
  __shared__ float red[8];
  float acc = 0.f;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
    const float d = pred[i] - target[i];
    acc += d * d;
    if (dpred != nullptr) dpred[i] = 2.f * d * grad_scale;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int w = 0; w < 8; ++w) t += red[w];
    atomicAdd(loss_sum, t);
  }
}

}  // namespace loss
}  // namespace dtb

using namespace dtb;
using namespace dtb::loss;

extern "C" int dtb_cross_entropy_fwd_bwd(void* logits, const long long* targets, float* row_loss, int rows, int vocab,
                                         long long ld, float grad_scale, int write_grad, void* stream) {
  DTB_REQUIRE(rows > 0 && vocab > 0 && ld >= vocab, "dtb_cross_entropy: bad shape rows=%d vocab=%d ld=%lld", rows,
              vocab, ld);
  DTB_REQUIRE(ld % 8 == 0 && (reinterpret_cast<uintptr_t>(logits) & 15) == 0,
              "dtb_cross_entropy: logits must be 16-byte aligned with ld %% 8 == 0");
  const size_t smem = static_cast<size_t>(ld) * 2;
  DTB_REQUIRE(smem <= 200 * 1024, "dtb_cross_entropy: vocab %d exceeds the shared-memory row cache", vocab);
  static size_t configured = 0;
  if (smem > 48 * 1024 && smem > configured) {
    DTB_CUDA_OK(cudaFuncSetAttribute(ce_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    configured = 200 * 1024;
  }
  ce_kernel<<<rows, CE_THREADS, smem, static_cast<cudaStream_t>(stream)>>>(
      static_cast<__nv_bfloat16*>(logits), targets, row_loss, vocab, ld, grad_scale, write_grad);
  DTB_LAUNCH_OK();
  return 0;
}

extern "C" int dtb_soft_cross_entropy_fwd_bwd(const float* logits, const float* targets, float* row_loss,
                                              float* dlogits, int rows, int cols, float grad_scale, void* stream) {
  DTB_REQUIRE(rows > 0 && cols > 0, "dtb_soft_cross_entropy: empty input");
  soft_ce_kernel<<<(rows + 3) / 4, 128, 0, static_cast<cudaStream_t>(stream)>>>(logits, targets, row_loss, dlogits, rows,
                                                                                cols, grad_scale);
  DTB_LAUNCH_OK();
  return 0;
}

extern "C" int dtb_mse_fwd_bwd(const float* pred, const float* target, float* loss_sum, float* dpred, long long n,
                               float grad_scale, void* stream) {
  DTB_REQUIRE(n > 0, "dtb_mse: empty input");
  long long blocks = (n + 255) / 256;
  if (blocks > sm_count() * 4) blocks = sm_count() * 4;
  DTB_CUDA_OK(cudaMemsetAsync(loss_sum, 0, sizeof(float), static_cast<cudaStream_t>(stream)));
  mse_kernel<<<static_cast<int>(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(pred, target, loss_sum, dpred, n,
                                                                                      grad_scale);
  DTB_LAUNCH_OK();
  return 0;
}
