// LayerNorm forward / backward for transformer rows (768, 1600, ... <= 4096 columns).
// One warp owns one row: 16-byte loads, row held in registers, fp32 statistics via
// shuffles.  Forward optionally fuses the residual add that precedes every LN in a
// pre-LN GPT-2 block (y = LN(x + r), xres = x + r).  Backward is ONE pass over dy and x:
// the kernel that produces dx (plus the incoming residual-stream gradient) also accumulates
// each lane's columns of dgamma/dbeta over the rows its warp walks (registers up to 1024
// columns, per-warp shared memory above), with the next row's loads issued before the
// current row's arithmetic; CTA partials go to a caller-provided scratch buffer and a small
// second kernel adds them in a fixed order (deterministic).  The allocation-free entry
// point keeps the older three-kernel schedule, whose scratch need does not depend on the
// grid size.
//
// Replaces ATen native_layer_norm(+backward) for the GPT-2 configs.
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include "../../include/dtb200.h"
#include "../common/host.h"

namespace dtb {
namespace ln {

constexpr int WARPS = 4;
constexpr int MAXV = 16;                 // 16 vectors x 8 elements x 32 lanes = 4096 columns

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__device__ __forceinline__ void unpack8(const uint4& raw, float* f) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&raw);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float2 t = __bfloat1622float2(h[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}
__device__ __forceinline__ uint4 pack8(const float* f) {
  uint4 raw;
  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&raw);
#pragma unroll
  for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
  return raw;
}

template <int NV>
__global__ void __launch_bounds__(WARPS * 32)
fwd_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ res,
           const __nv_bfloat16* __restrict__ gamma, const __nv_bfloat16* __restrict__ beta,
           __nv_bfloat16* __restrict__ y, __nv_bfloat16* __restrict__ xres, float* __restrict__ mean,
           float* __restrict__ rstd, int rows, int cols, float eps) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nvec = cols >> 3;
  const bool has_res = res != nullptr;
  const int stride = gridDim.x * WARPS;
  int row = blockIdx.x * WARPS + warp;
  // the next row's loads are issued before the current row's reductions: one warp keeps two rows in flight
  uint4 nx[NV], nr[NV];
  auto issue = [&](int r) {
    const long long base = static_cast<long long>(r) * cols;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int vi = lane + 32 * i;
      if (vi < nvec) {
        nx[i] = *reinterpret_cast<const uint4*>(x + base + vi * 8);
        if (has_res) nr[i] = *reinterpret_cast<const uint4*>(res + base + vi * 8);
      }
    }
  };
  if (row < rows) issue(row);
  while (row < rows) {
    const long long base = static_cast<long long>(row) * cols;
    float v[NV][8];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int vi = lane + 32 * i;
      if (vi < nvec) {
        unpack8(nx[i], v[i]);
        if (has_res) {
          float r[8];
          unpack8(nr[i], r);
#pragma unroll
          for (int j = 0; j < 8; ++j) v[i][j] += r[j];
          uint4 packed = pack8(v[i]);
          *reinterpret_cast<uint4*>(xres + base + vi * 8) = packed;
          unpack8(packed, v[i]);   // normalise exactly what backward will re-read
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) s += v[i][j];
      }
    }
    const int cur = row;
    row += stride;
    if (row < rows) issue(row);
    const float mu = warp_sum(s) / cols;
    float sq = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      if (lane + 32 * i < nvec) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          float d = v[i][j] - mu;
          sq += d * d;
        }
      }
    }
    const float rs = rsqrtf(warp_sum(sq) / cols + eps);
    if (lane == 0) {
      mean[cur] = mu;
      rstd[cur] = rs;
    }
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int vi = lane + 32 * i;
      if (vi < nvec) {
        float g[8], b[8], o[8];
        unpack8(*reinterpret_cast<const uint4*>(gamma + vi * 8), g);
        unpack8(*reinterpret_cast<const uint4*>(beta + vi * 8), b);
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = (v[i][j] - mu) * rs * g[j] + b[j];
        *reinterpret_cast<uint4*>(y + base + vi * 8) = pack8(o);
      }
    }
  }
}

template <int NV>
__global__ void __launch_bounds__(WARPS * 32)
bwd_dx_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
              const __nv_bfloat16* __restrict__ gamma, const float* __restrict__ mean, const float* __restrict__ rstd,
              const __nv_bfloat16* __restrict__ dres, __nv_bfloat16* __restrict__ dx, int rows, int cols) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nvec = cols >> 3;
  for (int row = blockIdx.x * WARPS + warp; row < rows; row += gridDim.x * WARPS) {
    const long long base = static_cast<long long>(row) * cols;
    const float mu = mean[row], rs = rstd[row];
    float xh[NV][8], wdy[NV][8];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int vi = lane + 32 * i;
      if (vi < nvec) {
        float g[8];
        unpack8(*reinterpret_cast<const uint4*>(gamma + vi * 8), g);
        unpack8(*reinterpret_cast<const uint4*>(dy + base + vi * 8), wdy[i]);
        unpack8(*reinterpret_cast<const uint4*>(x + base + vi * 8), xh[i]);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          xh[i][j] = (xh[i][j] - mu) * rs;
          wdy[i][j] *= g[j];
          s1 += wdy[i][j];
          s2 += wdy[i][j] * xh[i][j];
        }
      }
    }
    s1 = warp_sum(s1) / cols;
    s2 = warp_sum(s2) / cols;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int vi = lane + 32 * i;
      if (vi < nvec) {
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = rs * (wdy[i][j] - s1 - xh[i][j] * s2);
        if (dres != nullptr) {
          float r[8];
          unpack8(*reinterpret_cast<const uint4*>(dres + base + vi * 8), r);
#pragma unroll
          for (int j = 0; j < 8; ++j) o[j] += r[j];
        }
        *reinterpret_cast<uint4*>(dx + base + vi * 8) = pack8(o);
      }
    }
  }
}

// dgamma/dbeta partial sums: CTA (bx, by) owns 64 columns and every PARAM_CHUNKS-th
// block of rows; partial[by][0|1][cols].
constexpr int PARAM_CHUNKS = 32;
__global__ void __launch_bounds__(256)
bwd_param_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                 const float* __restrict__ mean, const float* __restrict__ rstd, float* __restrict__ partial,
                 int rows, int cols) {
  __shared__ float pg[32][65];
  __shared__ float pb[32][65];
  const int cx = threadIdx.x & 7, ry = threadIdx.x >> 3;
  const int col = blockIdx.x * 64 + cx * 8;
  float ag[8] = {0, 0, 0, 0, 0, 0, 0, 0}, ab[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  if (col < cols) {
    for (int r = blockIdx.y * 32 + ry; r < rows; r += gridDim.y * 32) {
      float d[8], xv[8];
      unpack8(*reinterpret_cast<const uint4*>(dy + static_cast<long long>(r) * cols + col), d);
      unpack8(*reinterpret_cast<const uint4*>(x + static_cast<long long>(r) * cols + col), xv);
      const float mu = mean[r], rs = rstd[r];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        ag[j] += d[j] * (xv[j] - mu) * rs;
        ab[j] += d[j];
      }
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    pg[ry][cx * 8 + j] = ag[j];
    pb[ry][cx * 8 + j] = ab[j];
  }
  __syncthreads();
  if (threadIdx.x < 128) {
    const int c = threadIdx.x & 63;
    const bool is_beta = threadIdx.x >= 64;
    float s = 0.f;
#pragma unroll 8
    for (int r = 0; r < 32; ++r) s += is_beta ? pb[r][c] : pg[r][c];
    const int gc = blockIdx.x * 64 + c;
    if (gc < cols) partial[(static_cast<long long>(blockIdx.y) * 2 + (is_beta ? 1 : 0)) * cols + gc] = s;
  }
}

template <typename T>
__global__ void __launch_bounds__(256)
bwd_finish_kernel(const float* __restrict__ partial, int nparts, int cols, T* __restrict__ dgamma,
                  T* __restrict__ dbeta, int accumulate) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= 2 * cols) return;
  float s = 0.f;
  for (int p = 0; p < nparts; ++p) s += partial[static_cast<long long>(p) * 2 * cols + c];
  T* dst = c < cols ? dgamma + c : dbeta + (c - cols);
  if (accumulate) s += static_cast<float>(*dst);
  *dst = static_cast<T>(s);
}


// ---- single-pass backward ----------------------------------------------------------------------------------------
// Partial-sum rows the fused kernel may write (one per CTA); dtb_layernorm_bwd_scratch_floats sizes the scratch for it.
constexpr int MAX_PARTS = 512;

// NV = 16-byte vectors per lane.  SMEM_ACC = false (<= 1024 columns): dgamma/dbeta accumulators, gamma and the
// prefetched next row all live in registers.  SMEM_ACC = true (wider rows): accumulators in the warp's own
// shared-memory slab (each lane owns its columns, so no atomics), gamma re-read through L1, no prefetch -- the lower
// register count buys the occupancy that hides the loads instead.
template <int NV, bool SMEM_ACC>
__global__ void __launch_bounds__(WARPS * 32)
bwd_fused_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                 const __nv_bfloat16* __restrict__ gamma, const float* __restrict__ mean, const float* __restrict__ rstd,
                 const __nv_bfloat16* __restrict__ dres, __nv_bfloat16* __restrict__ dx, float* __restrict__ partial,
                 int rows, int cols) {
  extern __shared__ float acc_smem[];                 // [WARPS][2][cols]
  constexpr bool PREFETCH = !SMEM_ACC;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nvec = cols >> 3;
  const bool has_res = dres != nullptr;
  float* my_g = acc_smem + static_cast<size_t>(warp) * 2 * cols;
  float* my_b = my_g + cols;
  float ag[SMEM_ACC ? 1 : NV][8], ab[SMEM_ACC ? 1 : NV][8];
  uint4 graw[SMEM_ACC ? 1 : NV];
  if (SMEM_ACC) {
    for (int c = lane; c < 2 * cols; c += 32) my_g[c] = 0.f;
    __syncwarp();
  } else {
#pragma unroll
    for (int i = 0; i < NV; ++i) {
#pragma unroll
      for (int j = 0; j < 8; ++j) ag[i][j] = ab[i][j] = 0.f;
      const int vi = lane + 32 * i;
      graw[i] = vi < nvec ? *reinterpret_cast<const uint4*>(gamma + vi * 8) : make_uint4(0, 0, 0, 0);
    }
  }
  auto gamma_of = [&](int i, float* g) {
    if (SMEM_ACC) unpack8(*reinterpret_cast<const uint4*>(gamma + (lane + 32 * i) * 8), g);
    else unpack8(graw[i], g);
  };

  const int stride = gridDim.x * WARPS;
  int row = blockIdx.x * WARPS + warp;
  uint4 ndy[NV], nx[NV], ndr[NV];      // the row being loaded
  float nmu = 0.f, nrs = 0.f;
  auto issue = [&](int r) {
    const long long base = static_cast<long long>(r) * cols;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int vi = lane + 32 * i;
      if (vi < nvec) {
        ndy[i] = *reinterpret_cast<const uint4*>(dy + base + vi * 8);
        nx[i] = *reinterpret_cast<const uint4*>(x + base + vi * 8);
        if (has_res) ndr[i] = *reinterpret_cast<const uint4*>(dres + base + vi * 8);
      }
    }
    nmu = mean[r];
    nrs = rstd[r];
  };
  if (PREFETCH && row < rows) issue(row);
  while (row < rows) {
    if (!PREFETCH) issue(row);
    // the current row stays packed (bf16) in registers; it is unpacked once for the row sums and once for the outputs
    uint4 cdy[NV], cx[NV], cdr[NV];
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      cdy[i] = ndy[i];
      cx[i] = nx[i];
      cdr[i] = ndr[i];
    }
    const float mu = nmu, rs = nrs;
    const int cur = row;
    row += stride;
    if (PREFETCH && row < rows) issue(row);          // next row's loads fly while this row is reduced and written
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      if (lane + 32 * i < nvec) {
        float g[8], d[8], xv[8];
        gamma_of(i, g);
        unpack8(cdy[i], d);
        unpack8(cx[i], xv);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float w = d[j] * g[j];
          s1 += w;
          s2 += w * ((xv[j] - mu) * rs);
        }
      }
    }
    s1 = warp_sum(s1) / cols;
    s2 = warp_sum(s2) / cols;
    const long long base = static_cast<long long>(cur) * cols;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int vi = lane + 32 * i;
      if (vi < nvec) {
        float g[8], d[8], xh[8], o[8];
        gamma_of(i, g);
        unpack8(cdy[i], d);
        unpack8(cx[i], xh);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          xh[j] = (xh[j] - mu) * rs;
          o[j] = rs * (d[j] * g[j] - s1 - xh[j] * s2);
        }
        if (has_res) {
          float r[8];
          unpack8(cdr[i], r);
#pragma unroll
          for (int j = 0; j < 8; ++j) o[j] += r[j];
        }
        *reinterpret_cast<uint4*>(dx + base + vi * 8) = pack8(o);
        if (SMEM_ACC) {
          float4* pg = reinterpret_cast<float4*>(my_g + vi * 8);
          float4* pb = reinterpret_cast<float4*>(my_b + vi * 8);
          float4 g0 = pg[0], g1 = pg[1], b0 = pb[0], b1 = pb[1];
          g0.x += d[0] * xh[0]; g0.y += d[1] * xh[1]; g0.z += d[2] * xh[2]; g0.w += d[3] * xh[3];
          g1.x += d[4] * xh[4]; g1.y += d[5] * xh[5]; g1.z += d[6] * xh[6]; g1.w += d[7] * xh[7];
          b0.x += d[0]; b0.y += d[1]; b0.z += d[2]; b0.w += d[3];
          b1.x += d[4]; b1.y += d[5]; b1.z += d[6]; b1.w += d[7];
          pg[0] = g0; pg[1] = g1; pb[0] = b0; pb[1] = b1;
        } else {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            ag[i][j] += d[j] * xh[j];
            ab[i][j] += d[j];
          }
        }
      }
    }
  }
  if (!SMEM_ACC) {
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int vi = lane + 32 * i;
      if (vi < nvec) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          my_g[vi * 8 + j] = ag[i][j];
          my_b[vi * 8 + j] = ab[i][j];
        }
      }
    }
  }
  __syncthreads();
  // CTA partial = sum over the warps, in warp order; row blockIdx.x of partial[parts][2][cols]
  float* out = partial + static_cast<size_t>(blockIdx.x) * 2 * cols;
  for (int c = threadIdx.x; c < 2 * cols; c += WARPS * 32) {
    float v = 0.f;
#pragma unroll
    for (int w = 0; w < WARPS; ++w) v += acc_smem[static_cast<size_t>(w) * 2 * cols + c];
    out[c] = v;
  }
}

// dgamma|dbeta[c] = sum over `nparts` partial rows, fixed order.  A block owns 32 columns; its 32 warps each add every
// 32nd partial row (loads of one warp = 128 contiguous bytes, four independent loads in flight per thread) and the
// block combines the 32 slices through shared memory.  (With 8 slices per block the kernel was a chain of ~55 dependent
// L2 round trips on 48 CTAs: 16.8 us for 2.7 MB.)
template <typename T>
__global__ void __launch_bounds__(1024)
bwd_fused_finish_kernel(const float* __restrict__ partial, int nparts, int cols, T* __restrict__ dgamma,
                        T* __restrict__ dbeta, int accumulate) {
  __shared__ float red[32][33];
  const int cx = threadIdx.x & 31, sy = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + cx;
  float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
  if (c < 2 * cols) {
    const size_t pitch = static_cast<size_t>(2) * cols;
    int p = sy;
    for (; p + 96 < nparts; p += 128) {
      s0 += partial[static_cast<size_t>(p) * pitch + c];
      s1 += partial[static_cast<size_t>(p + 32) * pitch + c];
      s2 += partial[static_cast<size_t>(p + 64) * pitch + c];
      s3 += partial[static_cast<size_t>(p + 96) * pitch + c];
    }
    for (; p < nparts; p += 32) s0 += partial[static_cast<size_t>(p) * pitch + c];
  }
  red[sy][cx] = (s0 + s1) + (s2 + s3);
  __syncthreads();
  if (sy == 0 && c < 2 * cols) {
    float v = 0.f;
#pragma unroll
    for (int k = 0; k < 32; ++k) v += red[k][cx];
    T* dst = c < cols ? dgamma + c : dbeta + (c - cols);
    if (accumulate) v += static_cast<float>(*dst);
    *dst = static_cast<T>(v);
  }
}

}  // namespace ln
}  // namespace dtb
using namespace dtb;
using namespace dtb::ln;

// Single-pass backward launch shared by dtb_layernorm_bwd and dtb_layernorm_bwd_bf16grads.
template <typename T>
static int launch_bwd_fused(const void* dy, const void* x, const void* gamma, const float* mean, const float* rstd,
                            const void* dres_in, void* dx, T* dgamma, T* dbeta, float* scratch, int rows, int cols,
                            int accumulate, cudaStream_t s) {
  int grid = (rows + WARPS - 1) / WARPS;
  const int cap = sm_count() * 3 < MAX_PARTS ? sm_count() * 3 : MAX_PARTS;
  if (grid > cap) grid = cap;
  const int nv = (cols / 8 + 31) / 32;
  const size_t smem = static_cast<size_t>(WARPS) * 2 * cols * sizeof(float);
  auto DY = static_cast<const __nv_bfloat16*>(dy);
  auto X = static_cast<const __nv_bfloat16*>(x);
  auto G = static_cast<const __nv_bfloat16*>(gamma);
  auto DR = static_cast<const __nv_bfloat16*>(dres_in);
  auto DX = static_cast<__nv_bfloat16*>(dx);
#define DTB_LN_BWD(NV_, ACC_)                                                                                     \
  do {                                                                                                            \
    static bool configured = false;                                                                               \
    if (!configured) {                                                                                            \
      DTB_CUDA_OK(cudaFuncSetAttribute(bwd_fused_kernel<NV_, ACC_>, cudaFuncAttributeMaxDynamicSharedMemorySize,  \
                                       WARPS * 2 * 2048 * 4));                                                    \
      configured = true;                                                                                          \
    }                                                                                                             \
    bwd_fused_kernel<NV_, ACC_><<<grid, WARPS * 32, smem, s>>>(DY, X, G, mean, rstd, DR, DX, scratch, rows, cols); \
  } while (0)
  if (nv <= 1) DTB_LN_BWD(1, false);
  else if (nv <= 2) DTB_LN_BWD(2, false);
  else if (nv <= 3) DTB_LN_BWD(3, false);
  else if (nv <= 4) DTB_LN_BWD(4, false);
  else if (nv <= 5) DTB_LN_BWD(5, true);
  else if (nv <= 7) DTB_LN_BWD(7, true);
  else DTB_LN_BWD(8, true);
#undef DTB_LN_BWD
  DTB_LAUNCH_OK();
  bwd_fused_finish_kernel<T><<<(2 * cols + 31) / 32, 1024, 0, s>>>(scratch, grid, cols, dgamma, dbeta, accumulate);
  DTB_LAUNCH_OK();
  return 0;
}

// The older three-kernel schedule (dx; column-blocked partial sums re-reading dy and x; finish).  Its scratch need,
// PARAM_CHUNKS x 2 x cols floats, does not depend on the grid, which is what the fixed-size built-in array of
// dtb_layernorm_bwd_noalloc requires.
static int launch_bwd_split(const void* dy, const void* x, const void* gamma, const float* mean, const float* rstd,
                            const void* dres_in, void* dx, float* dgamma, float* dbeta, float* scratch, int rows,
                            int cols, int accumulate, cudaStream_t s) {
  int grid = (rows + WARPS - 1) / WARPS;
  const int cap = sm_count() * 8;
  if (grid > cap) grid = cap;
  const int nv = (cols / 8 + 31) / 32;
  auto DY = static_cast<const __nv_bfloat16*>(dy);
  auto X = static_cast<const __nv_bfloat16*>(x);
  auto G = static_cast<const __nv_bfloat16*>(gamma);
  auto DR = static_cast<const __nv_bfloat16*>(dres_in);
  auto DX = static_cast<__nv_bfloat16*>(dx);
  if (nv <= 4) bwd_dx_kernel<4><<<grid, WARPS * 32, 0, s>>>(DY, X, G, mean, rstd, DR, DX, rows, cols);
  else bwd_dx_kernel<8><<<grid, WARPS * 32, 0, s>>>(DY, X, G, mean, rstd, DR, DX, rows, cols);
  DTB_LAUNCH_OK();
  bwd_param_kernel<<<dim3((cols + 63) / 64, PARAM_CHUNKS), 256, 0, s>>>(DY, X, mean, rstd, scratch, rows, cols);
  DTB_LAUNCH_OK();
  bwd_finish_kernel<float><<<(2 * cols + 255) / 256, 256, 0, s>>>(scratch, PARAM_CHUNKS, cols, dgamma, dbeta, accumulate);
  DTB_LAUNCH_OK();
  return 0;
}

extern "C" int dtb_layernorm_fwd(const void* x, const void* residual, const void* gamma, const void* beta, void* y,
                                 void* xres_out, float* mean, float* rstd, int rows, int cols, float eps,
                                 void* stream) {
  DTB_REQUIRE(rows > 0 && cols > 0 && cols % 8 == 0 && cols <= MAXV * 256,
              "dtb_layernorm_fwd: cols=%d must be a multiple of 8 and <= %d", cols, MAXV * 256);
  DTB_REQUIRE(residual == nullptr || xres_out != nullptr, "dtb_layernorm_fwd: residual given without xres_out");
  int grid = (rows + WARPS - 1) / WARPS;
  const int cap = sm_count() * 8;
  if (grid > cap) grid = cap;
  auto s = static_cast<cudaStream_t>(stream);
  auto X = static_cast<const __nv_bfloat16*>(x);
  auto R = static_cast<const __nv_bfloat16*>(residual);
  auto G = static_cast<const __nv_bfloat16*>(gamma);
  auto B = static_cast<const __nv_bfloat16*>(beta);
  auto Y = static_cast<__nv_bfloat16*>(y);
  auto XR = static_cast<__nv_bfloat16*>(xres_out);
  const int nv = (cols / 8 + 31) / 32;
  if (nv <= 1) fwd_kernel<1><<<grid, WARPS * 32, 0, s>>>(X, R, G, B, Y, XR, mean, rstd, rows, cols, eps);
  else if (nv <= 2) fwd_kernel<2><<<grid, WARPS * 32, 0, s>>>(X, R, G, B, Y, XR, mean, rstd, rows, cols, eps);
  else if (nv <= 3) fwd_kernel<3><<<grid, WARPS * 32, 0, s>>>(X, R, G, B, Y, XR, mean, rstd, rows, cols, eps);
  else if (nv <= 4) fwd_kernel<4><<<grid, WARPS * 32, 0, s>>>(X, R, G, B, Y, XR, mean, rstd, rows, cols, eps);
  else if (nv <= 7) fwd_kernel<7><<<grid, WARPS * 32, 0, s>>>(X, R, G, B, Y, XR, mean, rstd, rows, cols, eps);
  else if (nv <= 8) fwd_kernel<8><<<grid, WARPS * 32, 0, s>>>(X, R, G, B, Y, XR, mean, rstd, rows, cols, eps);
  else fwd_kernel<MAXV><<<grid, WARPS * 32, 0, s>>>(X, R, G, B, Y, XR, mean, rstd, rows, cols, eps);
  DTB_LAUNCH_OK();
  return 0;
}

extern "C" size_t dtb_layernorm_bwd_scratch_floats(int rows, int cols) {
  (void)rows;
  return static_cast<size_t>(MAX_PARTS) * 2 * static_cast<size_t>(cols);
}

extern "C" int dtb_layernorm_bwd(const void* dy, const void* x, const void* gamma, const float* mean,
                                 const float* rstd, const void* dres_in, void* dx, float* dgamma, float* dbeta,
                                 float* scratch, int rows, int cols, int accumulate_param_grads, void* stream) {
  DTB_REQUIRE(rows > 0 && cols > 0 && cols % 8 == 0 && cols <= 8 * 256,
              "dtb_layernorm_bwd: cols=%d must be a multiple of 8 and <= 2048", cols);
  DTB_REQUIRE(scratch != nullptr, "dtb_layernorm_bwd: scratch is required");
  return launch_bwd_fused<float>(dy, x, gamma, mean, rstd, dres_in, dx, dgamma, dbeta, scratch, rows, cols,
                                 accumulate_param_grads, static_cast<cudaStream_t>(stream));
}

// Same, with dgamma / dbeta written as bf16 (straight into bf16 gradient buffers such as DDP bucket views: saves the
// two cast launches per LayerNorm that the GPT-2 step used to make).
extern "C" int dtb_layernorm_bwd_bf16grads(const void* dy, const void* x, const void* gamma, const float* mean,
                                           const float* rstd, const void* dres_in, void* dx, void* dgamma, void* dbeta,
                                           float* scratch, int rows, int cols, int accumulate_param_grads, void* stream) {
  DTB_REQUIRE(rows > 0 && cols > 0 && cols % 8 == 0 && cols <= 8 * 256,
              "dtb_layernorm_bwd_bf16grads: cols=%d must be a multiple of 8 and <= 2048", cols);
  DTB_REQUIRE(scratch != nullptr, "dtb_layernorm_bwd_bf16grads: scratch is required");
  return launch_bwd_fused<__nv_bfloat16>(dy, x, gamma, mean, rstd, dres_in, dx, static_cast<__nv_bfloat16*>(dgamma),
                                         static_cast<__nv_bfloat16*>(dbeta), scratch, rows, cols, accumulate_param_grads,
                                         static_cast<cudaStream_t>(stream));
}

// ---------------------------------------------------------------------------------------------
// Allocation-free variant of dtb_layernorm_bwd.
//
// dtb_layernorm_bwd needs the caller to supply a scratch buffer for the per-chunk dgamma/dbeta
// partial sums.  Some users of this C ABI have no way to produce one: a plain C or Fortran host
// that links the library without a device allocator of its own, a plugin running inside a
// framework that has already claimed the whole GPU and forbids cudaMalloc from guests, or a
// latency-sensitive loop that cannot accept cudaMalloc's implicit device synchronisation and its
// out-of-memory failure mode.  For them the scratch lives in the library image itself: a file-scope
// __device__ array of DTB_LN_NOALLOC_MAX_COLS-sized capacity (32 chunks x 2 x 2048 floats = 512 KiB)
// that every GPU which loads the library holds until the process exits.
//
// In host code the array's name is only a host-side handle, not a device address, so the address
// the kernels receive is looked up with cudaGetSymbolAddress, once per device, with that device
// current, on the first call made on that device.  Make that first call outside any stream capture.
// The function calls no cudaMalloc / cudaFree or any other allocating routine.
//
// All calls on one GPU share the one array.  Nothing in it has to start at zero: every element the
// finishing kernel reads is overwritten by the partial-sum kernel of the same call, so stale
// contents from the previous call are harmless.  Two calls whose kernels overlap on the GPU (issued
// to different streams without ordering between them) race on the array and give wrong dgamma /
// dbeta; serialise them.
// No counterpart in the reference.
#include <mutex>

namespace dtb {
namespace ln {

constexpr int NOALLOC_MAX_COLS = 2048;
constexpr size_t NOALLOC_FLOATS = static_cast<size_t>(PARAM_CHUNKS) * 2 * NOALLOC_MAX_COLS;
static __device__ __align__(256) float g_ln_noalloc_scratch[NOALLOC_FLOATS];

// Device address of g_ln_noalloc_scratch on the current device (resolved once per device).
static float* noalloc_scratch_ptr() {
  constexpr int kMaxDevices = 64;
  static std::mutex mu;
  static float* resolved[kMaxDevices] = {nullptr};
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
    set_error("dtb_layernorm_bwd_noalloc: cudaGetDevice failed or device index %d unsupported (%s)", dev,
              cudaGetErrorString(e));
    return nullptr;
  }
  std::lock_guard<std::mutex> lock(mu);
  if (resolved[dev] == nullptr) {
    float* p = nullptr;
    e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_ln_noalloc_scratch);
    if (e != cudaSuccess || p == nullptr) {
      set_error("dtb_layernorm_bwd_noalloc: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
      return nullptr;
    }
    resolved[dev] = p;
  }
  return resolved[dev];
}

}  // namespace ln
}  // namespace dtb

extern "C" int dtb_layernorm_bwd_noalloc(const void* dy, const void* x, const void* gamma, const float* mean,
                                         const float* rstd, const void* dres_in, void* dx, float* dgamma,
                                         float* dbeta, int rows, int cols, int accumulate_param_grads,
                                         void* stream) {
  DTB_REQUIRE(rows > 0 && cols > 0 && cols % 8 == 0, "dtb_layernorm_bwd_noalloc: cols=%d must be a positive multiple of 8",
              cols);
  DTB_REQUIRE(cols <= NOALLOC_MAX_COLS,
              "dtb_layernorm_bwd_noalloc: cols=%d exceeds the built-in scratch capacity (max %d columns); use "
              "dtb_layernorm_bwd with a caller-provided scratch buffer",
              cols, NOALLOC_MAX_COLS);
  float* scratch = noalloc_scratch_ptr();
  if (scratch == nullptr) return 1;
  return launch_bwd_split(dy, x, gamma, mean, rstd, dres_in, dx, dgamma, dbeta, scratch, rows, cols,
                          accumulate_param_grads, static_cast<cudaStream_t>(stream));
}
