// Bandwidth-bound helpers: GELU fwd/bwd, add, casts, column sums (bias gradient),
// sum of squares (gradient-norm logging) and mean.  All use 16-byte vector
// accesses on the aligned body and a scalar tail.
//
// Reference call sites: the ATen elementwise / reduction kernels behind
// loss.backward() and torch.norm() (SURVEY K10, K15;
// /root/reference/src/playground/ddp_script.py:155-164).
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include "../../include/dtb200.h"
#include "../common/host.h"

namespace dtb {

__device__ __forceinline__ float gelu_f(float x) {
  const float k0 = 0.7978845608028654f, k1 = 0.044715f;
  float u = k0 * (x + k1 * x * x * x);
  float t;
  asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(u));
  return 0.5f * x * (1.0f + t);
}
__device__ __forceinline__ float gelu_grad_f(float x) {
  const float k0 = 0.7978845608028654f, k1 = 0.044715f;
  float x2 = x * x;
  float u = k0 * (x + k1 * x * x2);
  float t;
  asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(u));
  return 0.5f * (1.0f + t) + 0.5f * x * (1.0f - t * t) * k0 * (1.0f + 3.0f * k1 * x2);
}

struct Bf16x8 {
  uint4 raw;
  __device__ __forceinline__ void unpack(float (&f)[8]) const {
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&raw);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float2 t = __bfloat1622float2(h[i]);
      f[2 * i] = t.x;
      f[2 * i + 1] = t.y;
    }
  }
  __device__ __forceinline__ void pack(const float (&f)[8]) {
    __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&raw);
#pragma unroll
    for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
  }
};

enum { OP_GELU = 0, OP_GELU_BWD = 1, OP_ADD = 2 };

template <int OP>
__global__ void __launch_bounds__(256) eltwise_bf16_kernel(const __nv_bfloat16* __restrict__ a,
                                                            const __nv_bfloat16* __restrict__ b,
                                                            __nv_bfloat16* __restrict__ out, long long n) {
  const long long nvec = n / 8;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < nvec; i += stride) {
    Bf16x8 va, vb, vo;
    float fa[8], fb[8], fo[8];
    va.raw = reinterpret_cast<const uint4*>(a)[i];
    va.unpack(fa);
    if (OP != OP_GELU) {
      vb.raw = reinterpret_cast<const uint4*>(b)[i];
      vb.unpack(fb);
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      if (OP == OP_GELU) fo[j] = gelu_f(fa[j]);
      if (OP == OP_GELU_BWD) fo[j] = gelu_grad_f(fa[j]) * fb[j];
      if (OP == OP_ADD) fo[j] = fa[j] + fb[j];
    }
    vo.pack(fo);
    reinterpret_cast<uint4*>(out)[i] = vo.raw;
  }
  if (blockIdx.x == 0) {
    for (long long i = nvec * 8 + threadIdx.x; i < n; i += blockDim.x) {
      float x = __bfloat162float(a[i]);
      float y = (OP != OP_GELU) ? __bfloat162float(b[i]) : 0.f;
      float o = OP == OP_GELU ? gelu_f(x) : (OP == OP_GELU_BWD ? gelu_grad_f(x) * y : x + y);
      out[i] = __float2bfloat16(o);
    }
  }
}

__global__ void __launch_bounds__(256) cast_f32_bf16_kernel(const float* __restrict__ in, __nv_bfloat16* __restrict__ out,
                                                            long long n) {
  const long long nvec = n / 8;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < nvec; i += stride) {
    float4 lo = reinterpret_cast<const float4*>(in)[2 * i];
    float4 hi = reinterpret_cast<const float4*>(in)[2 * i + 1];
    Bf16x8 o;
    float f[8] = {lo.x, lo.y, lo.z, lo.w, hi.x, hi.y, hi.z, hi.w};
    o.pack(f);
    reinterpret_cast<uint4*>(out)[i] = o.raw;
  }
  if (blockIdx.x == 0)
    for (long long i = nvec * 8 + threadIdx.x; i < n; i += blockDim.x) out[i] = __float2bfloat16(in[i]);
}

__global__ void __launch_bounds__(256) cast_bf16_f32_kernel(const __nv_bfloat16* __restrict__ in, float* __restrict__ out,
                                                            long long n) {
  const long long nvec = n / 8;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < nvec; i += stride) {
    Bf16x8 v;
    v.raw = reinterpret_cast<const uint4*>(in)[i];
    float f[8];
    v.unpack(f);
    reinterpret_cast<float4*>(out)[2 * i] = make_float4(f[0], f[1], f[2], f[3]);
    reinterpret_cast<float4*>(out)[2 * i + 1] = make_float4(f[4], f[5], f[6], f[7]);
  }
  if (blockIdx.x == 0)
    for (long long i = nvec * 8 + threadIdx.x; i < n; i += blockDim.x) out[i] = __bfloat162float(in[i]);
}

// Column sums of a row-major [rows, cols] bf16 matrix.  Each CTA owns 64 columns
// (8 lanes x 8 columns) and 64 row-lanes, each row-lane keeps four independent 16-byte
// loads in flight; a shared-memory tree finishes the sum, so no scratch buffer or
// atomics are needed and the result is deterministic.
constexpr int COLSUM_THREADS = 512;
constexpr int COLSUM_ROWLANES = COLSUM_THREADS / 8;
// gridDim.y row chunks: chunk y owns rows [y*rows_per_chunk, ...); with more than one chunk the
// per-chunk sums go to `partial[y][cols]` (fp32) and colsum_finish_kernel adds them in chunk order.
__global__ void __launch_bounds__(COLSUM_THREADS) colsum_bf16_kernel(const __nv_bfloat16* __restrict__ x,
                                                                     void* __restrict__ out, int rows, int cols,
                                                                     long long ldx, int out_fp32,
                                                                     float* __restrict__ partial, int rows_per_chunk) {
  __shared__ float part[COLSUM_ROWLANES][65];
  const int cx = threadIdx.x & 7;    // column group within the CTA
  const int ry = threadIdx.x >> 3;   // row lane
  const int col = blockIdx.x * 64 + cx * 8;
  x += static_cast<long long>(blockIdx.y) * rows_per_chunk * ldx;
  rows = min(rows_per_chunk, rows - static_cast<int>(blockIdx.y) * rows_per_chunk);
  float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  if (col < cols) {
    if (col + 8 <= cols) {
      int r = ry;
      for (; r + 3 * COLSUM_ROWLANES < rows; r += 4 * COLSUM_ROWLANES) {
        Bf16x8 v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u)
          v[u].raw = *reinterpret_cast<const uint4*>(x + static_cast<long long>(r + u * COLSUM_ROWLANES) * ldx + col);
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          float f[8];
          v[u].unpack(f);
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[j] += f[j];
        }
      }
      for (; r < rows; r += COLSUM_ROWLANES) {
        Bf16x8 v;
        v.raw = *reinterpret_cast<const uint4*>(x + static_cast<long long>(r) * ldx + col);
        float f[8];
        v.unpack(f);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += f[j];
      }
    } else {
      for (int r = ry; r < rows; r += COLSUM_ROWLANES) {
        const __nv_bfloat16* p = x + static_cast<long long>(r) * ldx + col;
        for (int j = 0; j < 8 && col + j < cols; ++j) acc[j] += __bfloat162float(p[j]);
      }
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) part[ry][cx * 8 + j] = acc[j];
  __syncthreads();
  if (threadIdx.x < 64) {
    float s = 0.f;
#pragma unroll 8
    for (int r = 0; r < COLSUM_ROWLANES; ++r) s += part[r][threadIdx.x];
    const int c = blockIdx.x * 64 + threadIdx.x;
    if (c < cols) {
      if (partial != nullptr) partial[static_cast<long long>(blockIdx.y) * cols + c] = s;
      else if (out_fp32) reinterpret_cast<float*>(out)[c] = s;
      else reinterpret_cast<__nv_bfloat16*>(out)[c] = __float2bfloat16(s);
    }
  }
}

__global__ void __launch_bounds__(256) colsum_finish_kernel(const float* __restrict__ partial, int chunks, int cols,
                                                            void* __restrict__ out, int out_fp32) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= cols) return;
  float s = 0.f;
  for (int y = 0; y < chunks; ++y) s += partial[static_cast<long long>(y) * cols + c];
  if (out_fp32) reinterpret_cast<float*>(out)[c] = s;
  else reinterpret_cast<__nv_bfloat16*>(out)[c] = __float2bfloat16(s);
}

__device__ __forceinline__ float block_sum(float v, float* smem) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) smem[warp] = v;
  __syncthreads();
  const int nwarps = (blockDim.x + 31) >> 5;
  v = (threadIdx.x < nwarps) ? smem[threadIdx.x] : 0.f;
  if (warp == 0) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  }
  __syncthreads();
  return v;  // valid in warp 0
}

__global__ void __launch_bounds__(256) sumsq_kernel(const void* __restrict__ x, int is_fp32, long long n,
                                                    float* __restrict__ out) {
  __shared__ float red[32];
  float acc = 0.f;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
    float v = is_fp32 ? reinterpret_cast<const float*>(x)[i] : __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(x)[i]);
    acc += v * v;
  }
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) atomicAdd(out, acc);
}

__global__ void __launch_bounds__(1024) mean_f32_kernel(const float* __restrict__ x, long long n, float* __restrict__ out) {
  __shared__ float red[32];
  float acc = 0.f;
  for (long long i = threadIdx.x; i < n; i += blockDim.x) acc += x[i];
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) *out = acc / static_cast<float>(n);
}

static inline int grid_for(long long work_items, int block, int per_sm = 8) {
  long long blocks = (work_items + block - 1) / block;
  long long cap = static_cast<long long>(sm_count()) * per_sm;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return static_cast<int>(blocks);
}

}  // namespace dtb

using namespace dtb;

extern "C" int dtb_gelu_fwd_bf16(const void* x, void* y, long long n, void* stream) {
  if (n <= 0) return 0;
  eltwise_bf16_kernel<OP_GELU><<<grid_for(n / 8 + 1, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(x), nullptr, static_cast<__nv_bfloat16*>(y), n);
  DTB_LAUNCH_OK();
  return 0;
}

extern "C" int dtb_gelu_bwd_bf16(const void* x, const void* dy, void* dx, long long n, void* stream) {
  if (n <= 0) return 0;
  eltwise_bf16_kernel<OP_GELU_BWD><<<grid_for(n / 8 + 1, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(x), static_cast<const __nv_bfloat16*>(dy), static_cast<__nv_bfloat16*>(dx), n);
  DTB_LAUNCH_OK();
  return 0;
}

extern "C" int dtb_add_bf16(const void* a, const void* b, void* out, long long n, void* stream) {
  if (n <= 0) return 0;
  eltwise_bf16_kernel<OP_ADD><<<grid_for(n / 8 + 1, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(a), static_cast<const __nv_bfloat16*>(b), static_cast<__nv_bfloat16*>(out), n);
  DTB_LAUNCH_OK();
  return 0;
}

extern "C" int dtb_cast_f32_to_bf16(const float* in, void* out, long long n, void* stream) {
  if (n <= 0) return 0;
  cast_f32_bf16_kernel<<<grid_for(n / 8 + 1, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      in, static_cast<__nv_bfloat16*>(out), n);
  DTB_LAUNCH_OK();
  return 0;
}

// bf16(in) -> out, then in = 0 (hand-over of the fp32 bias-gradient accumulator the GEMM epilogue adds into)
__global__ void __launch_bounds__(256) take_f32_bf16_kernel(float* __restrict__ in, __nv_bfloat16* __restrict__ out,
                                                            long long n) {
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i < n) {
    out[i] = __float2bfloat16(in[i]);
    in[i] = 0.f;
  }
}

extern "C" int dtb_take_f32_to_bf16(float* in, void* out, long long n, void* stream) {
  if (n <= 0) return 0;
  DTB_REQUIRE(in != nullptr && out != nullptr, "dtb_take_f32_to_bf16: null argument");
  take_f32_bf16_kernel<<<static_cast<int>((n + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      in, static_cast<__nv_bfloat16*>(out), n);
  DTB_LAUNCH_OK();
  return 0;
}

extern "C" int dtb_cast_bf16_to_f32(const void* in, float* out, long long n, void* stream) {
  if (n <= 0) return 0;
  cast_bf16_f32_kernel<<<grid_for(n / 8 + 1, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(in), out, n);
  DTB_LAUNCH_OK();
  return 0;
}

extern "C" int dtb_colsum_bf16(const void* x, void* out, int rows, int cols, long long ldx, int out_fp32,
                               void* stream) {
  DTB_REQUIRE(rows > 0 && cols > 0, "dtb_colsum_bf16: empty input");
  DTB_REQUIRE(ldx % 8 == 0 && (reinterpret_cast<uintptr_t>(x) & 15) == 0, "dtb_colsum_bf16: x must be 16-byte aligned");
  colsum_bf16_kernel<<<(cols + 63) / 64, COLSUM_THREADS, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(x), out, rows, cols, ldx, out_fp32, nullptr, rows);
  DTB_LAUNCH_OK();
  return 0;
}

extern "C" size_t dtb_colsum_split_scratch_floats(int rows, int cols) {
  (void)rows;
  return static_cast<size_t>(DTB_COLSUM_MAX_CHUNKS) * static_cast<size_t>(cols);
}

extern "C" int dtb_colsum_bf16_split(const void* x, void* out, int rows, int cols, long long ldx, int out_fp32,
                                     float* scratch, void* stream) {
  DTB_REQUIRE(rows > 0 && cols > 0, "dtb_colsum_bf16_split: empty input");
  DTB_REQUIRE(ldx % 8 == 0 && (reinterpret_cast<uintptr_t>(x) & 15) == 0,
              "dtb_colsum_bf16_split: x must be 16-byte aligned");
  DTB_REQUIRE(scratch != nullptr, "dtb_colsum_bf16_split: scratch is required");
  // enough row chunks to put ~2 CTAs on every SM, at least 256 rows each
  const int col_blocks = (cols + 63) / 64;
  int chunks = (2 * sm_count() + col_blocks - 1) / col_blocks;
  if (chunks > DTB_COLSUM_MAX_CHUNKS) chunks = DTB_COLSUM_MAX_CHUNKS;
  if (chunks > (rows + 255) / 256) chunks = (rows + 255) / 256;
  if (chunks < 1) chunks = 1;
  const int rows_per_chunk = (rows + chunks - 1) / chunks;
  chunks = (rows + rows_per_chunk - 1) / rows_per_chunk;
  auto s = static_cast<cudaStream_t>(stream);
  if (chunks == 1) return dtb_colsum_bf16(x, out, rows, cols, ldx, out_fp32, stream);
  colsum_bf16_kernel<<<dim3(col_blocks, chunks), COLSUM_THREADS, 0, s>>>(static_cast<const __nv_bfloat16*>(x), out, rows,
                                                                         cols, ldx, out_fp32, scratch, rows_per_chunk);
  DTB_LAUNCH_OK();
  colsum_finish_kernel<<<(cols + 255) / 256, 256, 0, s>>>(scratch, chunks, cols, out, out_fp32);
  DTB_LAUNCH_OK();
  return 0;
}

extern "C" int dtb_sumsq(const void* x, int is_fp32, long long n, float* out, void* stream) {
  if (n <= 0) return 0;
  sumsq_kernel<<<grid_for(n, 256, 4), 256, 0, static_cast<cudaStream_t>(stream)>>>(x, is_fp32, n, out);
  DTB_LAUNCH_OK();
  return 0;
}

extern "C" int dtb_mean_f32(const float* x, long long n, float* out, void* stream) {
  DTB_REQUIRE(n > 0, "dtb_mean_f32: empty input");
  mean_f32_kernel<<<1, 1024, 0, static_cast<cudaStream_t>(stream)>>>(x, n, out);
  DTB_LAUNCH_OK();
  return 0;
}

// ---- row gather (input pipeline) ---------------------------------------------------------------------------------
// dst[i, :] = src[idx[i], :] for `rows` rows of `row_bytes` bytes.  `src` may be PINNED HOST memory (device-accessible
// under unified addressing): the batch is then assembled by the GPU pulling exactly the sampled rows over PCIe, with no
// host-side memcpy into a staging buffer at all.  This is the H2D leg of the reference's DataLoader(pin_memory=True) +
// .to(device) (/root/reference/src/distributed_trainer.py:176-183,204-211) done as one kernel on the copy stream.
// A handful of CTAs is enough: the transfer is PCIe-bound (~55 GB/s), not SM-bound.
namespace dtb {
namespace gather {

template <typename V>
__global__ void __launch_bounds__(256, 4)
gather_rows_kernel(const uint8_t* __restrict__ src, const long long* __restrict__ idx, uint8_t* __restrict__ dst, int rows,
                   long long row_vecs, long long src_rows) {
  const long long total = static_cast<long long>(rows) * row_vecs;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i0 = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i0 < total; i0 += stride * 4) {
    V val[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const long long i = i0 + u * stride;
      if (i < total) {
        const long long r = i / row_vecs, c = i - r * row_vecs;
        long long s = idx[r];
        if (src_rows > 0) s %= src_rows;
        val[u] = *reinterpret_cast<const V*>(src + (s * row_vecs + c) * sizeof(V));
      }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const long long i = i0 + u * stride;
      if (i < total) *reinterpret_cast<V*>(dst + i * sizeof(V)) = val[u];
    }
  }
}

}  // namespace gather
}  // namespace dtb

extern "C" int dtb_gather_rows(const void* src, const long long* idx, void* dst, int rows, long long row_bytes,
                               long long src_rows, int max_ctas, void* stream) {
  DTB_REQUIRE(src && idx && dst && rows >= 0 && row_bytes > 0 && row_bytes % 4 == 0,
              "dtb_gather_rows: row_bytes=%lld must be a positive multiple of 4", row_bytes);
  if (rows == 0) return 0;
  auto s = static_cast<cudaStream_t>(stream);
  auto S = static_cast<const uint8_t*>(src);
  auto D = static_cast<uint8_t*>(dst);
  const bool a16 = row_bytes % 16 == 0 && (reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) % 16 == 0;
  const bool a8 = row_bytes % 8 == 0 && (reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) % 8 == 0;
  const long long vec = a16 ? 16 : (a8 ? 8 : 4);
  const long long total = static_cast<long long>(rows) * (row_bytes / vec);
  long long need = (total + 256 * 4 - 1) / (256 * 4);
  const int cap = max_ctas > 0 ? max_ctas : 16;
  const int grid = static_cast<int>(need < cap ? (need < 1 ? 1 : need) : cap);
  using namespace dtb::gather;
  if (a16) gather_rows_kernel<uint4><<<grid, 256, 0, s>>>(S, idx, D, rows, row_bytes / 16, src_rows);
  else if (a8) gather_rows_kernel<uint2><<<grid, 256, 0, s>>>(S, idx, D, rows, row_bytes / 8, src_rows);
  else gather_rows_kernel<uint32_t><<<grid, 256, 0, s>>>(S, idx, D, rows, row_bytes / 4, src_rows);
  DTB_LAUNCH_OK();
  return 0;
}
