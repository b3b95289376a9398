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

#include <algorithm>
#include <cstdlib>

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

// ---- streaming CE: persistent CTAs, rows double-buffered in shared memory by bulk async copies --------------------
// The ncu capture of ce_kernel (profiles/kernel_zoo_ncu_metrics.csv) shows it issue-bound, not bandwidth-bound:
// 31 instructions and 2.25 MUFU.EX2 per logit (online max/sum-exp in the load loop, a second exponential in the
// gradient loop, non-ftz ex2 sequences), 0.48 of the HBM copy rate.  This kernel does one exponential and ~10
// instructions per logit:
//   copy   cp.async.bulk global -> shared of row i+1 (one elected thread, mbarrier complete_tx) under the math of row i
//   pass 1 packed bf16x2 max over the shared row (padding columns [vocab, ld) are set to -inf here)
//   pass 2 e = ex2(x*log2e - max*log2e): summed in fp32, written back over the row as bf16
//   pass 3 dlogits = e * grad_scale/sum (fp32 multiply, one rounding), 16-byte stores straight to the row in HBM; the
//          target column is recomputed from its fp32 logit so that p_t - 1 keeps full precision when p_t -> 1.
namespace {

__device__ __forceinline__ float ex2_ftz(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

__device__ __forceinline__ void bulk_load_row(void* sdst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   static_cast<uint32_t>(__cvta_generic_to_shared(sdst))),
               "l"(gsrc), "r"(bytes), "r"(static_cast<uint32_t>(__cvta_generic_to_shared(bar)))
               : "memory");
}
__device__ __forceinline__ void bar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(static_cast<uint32_t>(__cvta_generic_to_shared(bar))),
               "r"(count)
               : "memory");
}
__device__ __forceinline__ void bar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(
                   static_cast<uint32_t>(__cvta_generic_to_shared(bar))),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bar_wait(uint64_t* bar, uint32_t parity) {
  uint32_t done;
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}\n"
        : "=r"(done)
        : "r"(static_cast<uint32_t>(__cvta_generic_to_shared(bar))), "r"(parity)
        : "memory");
  } while (!done);
}

template <int T>
__device__ __forceinline__ float block_max(float v, float* red) {   // one __syncthreads; every thread gets the result
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  float r = (threadIdx.x & 31) < T / 32 ? red[threadIdx.x & 31] : -FLT_MAX;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) r = fmaxf(r, __shfl_xor_sync(0xffffffffu, r, o));
  return r;
}
template <int T>
__device__ __forceinline__ float block_sum(float v, float* red) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  float r = (threadIdx.x & 31) < T / 32 ? red[threadIdx.x & 31] : 0.f;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) r += __shfl_xor_sync(0xffffffffu, r, o);
  return r;
}

}  // namespace

template <int T>
__global__ void __launch_bounds__(T, T == 256 ? 4 : 1)
ce_stream_kernel(__nv_bfloat16* __restrict__ logits, const long long* __restrict__ targets,
                 float* __restrict__ row_loss, int rows, int vocab, long long ld, float grad_scale, int write_grad) {
  extern __shared__ __align__(128) uint8_t ce_smem[];
  __shared__ __align__(8) uint64_t full[2];
  __shared__ float red_m[2][32], red_s[2][32];
  constexpr float LOG2E = 1.4426950408889634f;
  const uint32_t row_bytes = static_cast<uint32_t>(ld) * 2;
  const int nvec = static_cast<int>(ld >> 3);
  const int first_pad_vec = vocab >> 3;             // vectors from here on contain padding columns
  const int tid = threadIdx.x;
  if (tid == 0) {
    bar_init(&full[0], 1);
    bar_init(&full[1], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  int row = blockIdx.x;
  if (row >= rows) return;
  if (tid == 0) {
    bar_expect_tx(&full[0], row_bytes);
    bulk_load_row(ce_smem, logits + static_cast<long long>(row) * ld, row_bytes, &full[0]);
  }
  long long tgt_next = targets[row];
  for (int it = 0; row < rows; ++it, row += gridDim.x) {
    const int s = it & 1;
    uint8_t* srow = ce_smem + static_cast<size_t>(s) * row_bytes;
    const int next = row + gridDim.x;
    const int tgt = static_cast<int>(tgt_next);
    if (next < rows) tgt_next = targets[next];       // in flight under this row's passes
    if (tid == 0 && next < rows) {
      // every thread finished reading / writing the other buffer before the __syncthreads that ended the previous
      // iteration; order those generic-proxy accesses before the async-proxy write
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      bar_expect_tx(&full[s ^ 1], row_bytes);
      bulk_load_row(ce_smem + static_cast<size_t>(s ^ 1) * row_bytes, logits + static_cast<long long>(next) * ld,
                    row_bytes, &full[s ^ 1]);
    }
    const bool ignore = tgt < 0 || tgt >= vocab;      // ignore_index < 0
    bar_wait(&full[s], (it >> 1) & 1);
    const float xt = ignore ? 0.f : __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(srow)[tgt]);
    // ---- pass 1: max ----
    __nv_bfloat162 mx2 = __float2bfloat162_rn(-FLT_MAX);
    for (int v = tid; v < nvec; v += T) {
      uint4 raw = *reinterpret_cast<const uint4*>(srow + v * 16);
      if (v >= first_pad_vec) {
        __nv_bfloat16* e = reinterpret_cast<__nv_bfloat16*>(&raw);
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (v * 8 + j >= vocab) e[j] = __ushort_as_bfloat16(0xff80);   // -inf: no weight in max, sum or gradient
        *reinterpret_cast<uint4*>(srow + v * 16) = raw;
      }
      const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&raw);
      mx2 = __hmax2(mx2, __hmax2(__hmax2(h[0], h[1]), __hmax2(h[2], h[3])));
    }
    const float m = block_max<T>(fmaxf(__low2float(mx2), __high2float(mx2)), red_m[s]);
    const float ml2 = m * LOG2E;
    // ---- pass 2: exponentials, sum ----
    float sum = 0.f;
    for (int v = tid; v < nvec; v += T) {
      uint4 raw = *reinterpret_cast<const uint4*>(srow + v * 16);
      uint32_t u[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float e0 = ex2_ftz(fmaf(__uint_as_float(u[j] << 16), LOG2E, -ml2));
        const float e1 = ex2_ftz(fmaf(__uint_as_float(u[j] & 0xffff0000u), LOG2E, -ml2));
        sum += e0 + e1;
        const __nv_bfloat162 pk = __floats2bfloat162_rn(e0, e1);
        u[j] = *reinterpret_cast<const uint32_t*>(&pk);
      }
      if (write_grad) *reinterpret_cast<uint4*>(srow + v * 16) = make_uint4(u[0], u[1], u[2], u[3]);
    }
    sum = block_sum<T>(sum, red_s[s]);
    if (tid == 0) row_loss[row] = ignore ? 0.f : (m + __logf(sum)) - xt;
    // ---- pass 3: gradient, in place over the logits row ----
    if (write_grad) {
      const float inv = ignore ? 0.f : grad_scale / sum;
      uint8_t* grow = reinterpret_cast<uint8_t*>(logits + static_cast<long long>(row) * ld);
      const int tvec = ignore ? -1 : (tgt >> 3);
      for (int v = tid; v < nvec; v += T) {
        const uint4 raw = *reinterpret_cast<const uint4*>(srow + v * 16);
        const uint32_t u[4] = {raw.x, raw.y, raw.z, raw.w};
        uint4 out;
        uint32_t* o = reinterpret_cast<uint32_t*>(&out);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const __nv_bfloat162 pk = __floats2bfloat162_rn(__uint_as_float(u[j] << 16) * inv,
                                                          __uint_as_float(u[j] & 0xffff0000u) * inv);
          o[j] = *reinterpret_cast<const uint32_t*>(&pk);
        }
        if (v == tvec) {
          const float g = ex2_ftz(fmaf(xt, LOG2E, -ml2)) * inv - grad_scale;
          const uint32_t gb = __bfloat16_as_ushort(__float2bfloat16(g));
          const int word = (tgt & 7) >> 1;
          const uint32_t keep = (tgt & 1) ? 0x0000ffffu : 0xffff0000u, put = (tgt & 1) ? gb << 16 : gb;
#pragma unroll
          for (int j = 0; j < 4; ++j)
            if (j == word) o[j] = (o[j] & keep) | put;
        }
        *reinterpret_cast<uint4*>(grow + v * 16) = out;
      }
    }
    __syncthreads();      // this buffer may be refilled (two iterations from now it is loaded again)
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
  const size_t row_bytes = static_cast<size_t>(ld) * 2;
  DTB_REQUIRE(row_bytes <= 200 * 1024, "dtb_cross_entropy: vocab %d exceeds the shared-memory row cache", vocab);
  auto s = static_cast<cudaStream_t>(stream);
  auto* lg = static_cast<__nv_bfloat16*>(logits);
  static const bool legacy = [] {
    const char* e = getenv("DTB_CE_LEGACY");
    return e && e[0] == '1';
  }();
  // two row buffers must fit (227 KB per CTA): GPT-2's 50304-wide rows take 2 x 100.6 KB, one 1024-thread CTA per SM
  if (!legacy && 2 * row_bytes <= 225 * 1024 && row_bytes >= 16 * 1024) {
    const size_t smem = 2 * row_bytes;
    const bool big = row_bytes >= 48 * 1024;
    static size_t configured_big = 0, configured_small = 0;
    if (big && smem > configured_big) {
      DTB_CUDA_OK(cudaFuncSetAttribute(ce_stream_kernel<1024>, cudaFuncAttributeMaxDynamicSharedMemorySize, 225 * 1024));
      configured_big = 225 * 1024;
    }
    if (!big && smem > 48 * 1024 && smem > configured_small) {
      DTB_CUDA_OK(cudaFuncSetAttribute(ce_stream_kernel<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
      configured_small = 96 * 1024;
    }
    if (big) {
      const int grid = rows < sm_count() ? rows : sm_count();
      ce_stream_kernel<1024><<<grid, 1024, smem, s>>>(lg, targets, row_loss, rows, vocab, ld, grad_scale, write_grad);
    } else {
      // small rows: several 256-thread CTAs per SM (shared memory and the 2048-thread limit allow up to 6-8)
      const int per_sm = static_cast<int>(std::min<size_t>(8, (200 * 1024) / (smem + 1024)));
      const int cap = sm_count() * (per_sm < 1 ? 1 : per_sm);
      const int grid = rows < cap ? rows : cap;
      ce_stream_kernel<256><<<grid, 256, smem, s>>>(lg, targets, row_loss, rows, vocab, ld, grad_scale, write_grad);
    }
    DTB_LAUNCH_OK();
    return 0;
  }
  const size_t smem = row_bytes;
  static size_t configured = 0;
  if (smem > 48 * 1024 && smem > configured) {
    DTB_CUDA_OK(cudaFuncSetAttribute(ce_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    configured = 200 * 1024;
  }
  ce_kernel<<<rows, CE_THREADS, smem, s>>>(lg, targets, row_loss, vocab, ld, grad_scale, write_grad);
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
