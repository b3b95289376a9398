// Causal self-attention for head_dim 64 on tcgen05 / TMEM (flash-style, no S/P in HBM).
//
// Forward, one CTA (128 threads) per (batch, head, 128-query tile):
//   S = Q K^T        tcgen05.mma  M128 N128 K64   -> TMEM cols [0,128)
//   softmax          tcgen05.ld 32x32b: thread r owns query row r, so the row max / sum need no
//                    shuffles; P (bf16) is written to shared memory in the 128B-swizzled K-major
//                    layout the next MMA reads
//   PV = P V         tcgen05.mma  M128 N64 K128   -> TMEM cols [128,192); O is rescaled and
//                    accumulated in registers
// Two CTAs fit per SM (TMEM 256 columns, ~113 KB smem), so one CTA's exponentials overlap the
// other's MMAs.  K/V tiles are double-buffered TMA loads straight out of the packed qkv tensor.
//
// Backward, one CTA per (batch, head, 128-key tile), looping over query tiles i >= j:
//   S = Q_i K_j^T, dP = dO_i V_j^T                 (issued together)
//   P = exp(S*scale - lse), dS = P*(dP - delta)*scale   -> shared memory (bf16)
//   dV += P^T dO_i, dK += dS^T Q_i  (TMEM accumulators, MN-major A operands = transposes for free)
//   dQ_i partial = dS K_j   -> TMEM -> fp32 atomics into a scratch accumulator
// A prep kernel computes delta = rowsum(dO * O) and a finish kernel casts dQ to bf16.
//
// The reference has no attention (north-star only); the baseline is torch SDPA.
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <math.h>
#include <cstdlib>

#include "../../include/dtb200.h"
#include "../common/host.h"
#include "../common/ptx.cuh"

namespace dtb {
namespace attn {

constexpr int TQ = 128, TK = 128, HD = 64;
constexpr int TILE_BYTES = 128 * 64 * 2;   // 16 KB: a [128 x 64] bf16 tile
constexpr float LOG2E = 1.4426950408889634f;

// 2^x on the SFU, flush-to-zero: exp2f() without -use_fast_math expands to a range check, a pre-scale and a
// post-square around MUFU.EX2 (4 instructions per score; ncu showed the softmax warps issue-bound)
__device__ __forceinline__ float fast_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

__device__ __forceinline__ uint32_t pack2(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}

// byte offset of the 16-byte chunk `chunk` (8 bf16 columns) of row `r` in a [128 x 128] bf16 tile
// stored as two [128 x 64] 128B-swizzled blocks
__device__ __forceinline__ uint32_t swz_chunk(int r, int chunk) {
  return static_cast<uint32_t>((chunk >> 3) * TILE_BYTES + r * 128 + (((chunk & 7) ^ (r & 7)) << 4));
}

constexpr uint32_t IDESC_S = umma_idesc_bf16(128, 128, 0, 0);      // Q K^T, dO V^T: both K-major
constexpr uint32_t IDESC_PV = umma_idesc_bf16(128, 64, 0, 1);      // P V, dS K: A K-major, B MN-major
constexpr uint32_t IDESC_TT = umma_idesc_bf16(128, 64, 1, 1);      // P^T dO, dS^T Q: both MN-major

// Descriptor words are assembled from compile-time pieces so the issuing warp only adds immediates:
//   hi = SBO (1024 B) | descriptor version 1 | 128-byte swizzle;  lo = start address >> 4 | LBO >> 4 << 16.
constexpr uint32_t DESC_HI = (1024u >> 4) | (1u << 14) | (2u << 29);
__device__ __forceinline__ uint64_t mk_desc(uint32_t lo) { return (static_cast<uint64_t>(DESC_HI) << 32) | lo; }
__device__ __forceinline__ uint32_t desc_lo(uint32_t smem_addr, uint32_t lbo_bytes) {
  return ((smem_addr >> 4) & 0x3FFFu) | (((lbo_bytes >> 4) & 0x3FFFu) << 16);
}
// K-major operand, one 64-wide K block: advance 32 B per UMMA_K
__device__ __forceinline__ uint64_t desc_k(uint32_t lo, int k16) { return mk_desc(lo + k16 * 2); }
// K-major operand spanning two 64-wide K blocks (P / dS as A with K = 128)
__device__ __forceinline__ uint64_t desc_k2(uint32_t lo, int k16) {
  return mk_desc(lo + (k16 >> 2) * (TILE_BYTES >> 4) + (k16 & 3) * 2);
}
// MN-major operand: K rows of 128 B, 8-row groups 1024 B apart (K step = 2048 B), 64-wide MN blocks LBO apart
__device__ __forceinline__ uint64_t desc_mn(uint32_t lo, int k16) { return mk_desc(lo + k16 * 128); }

__device__ __forceinline__ void red_add_v4(float* addr, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// -------------------------------------------------------------------------- forward
// Q, K x2 stages, V (single buffer), P: 96 KB + slack, so two CTAs fit on one SM
constexpr int FWD_SMEM = TILE_BYTES * (1 + 2 + 1 + 2) + 1024 + 128;

__global__ void __launch_bounds__(128, 2)
fwd_kernel(const __grid_constant__ CUtensorMap map_qkv, __nv_bfloat16* __restrict__ out, float* __restrict__ lse,
           int T, int H, float scale) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;
  uint8_t* sK = sQ + TILE_BYTES;            // 2 stages
  uint8_t* sV = sK + 2 * TILE_BYTES;        // single buffer: reloaded once P V of the previous tile is done
  uint8_t* sP = sV + TILE_BYTES;            // [128 x 128] bf16
  uint64_t* bars = reinterpret_cast<uint64_t*>(sP + 2 * TILE_BYTES);
  uint64_t* q_full = bars;
  uint64_t* k_full = bars + 1;              // [2]
  uint64_t* v_full = bars + 3;
  uint64_t* s_done = bars + 4;
  uint64_t* o_done = bars + 5;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 6);

  const int tid = threadIdx.x, warp = tid >> 5;
  // Longest-processing-time-first over the whole grid: CTAs are dispatched in linear order, so a 1-D grid whose first
  // B*H blocks are the longest query rows (n_kv = T/128 tiles), the next B*H the second longest, ... keeps the SMs
  // evenly loaded to the end.  With the (q-tile, head, batch) grid the causal 8,7,...,1 pattern repeated per head and
  // the kernel ran 20 tile-times where the average load is 14 (simulated; 47.5 us measured).
  const int n_bh = gridDim.x / (T / TQ);
  const int qt = T / TQ - 1 - static_cast<int>(blockIdx.x) / n_bh;
  const int bh = static_cast<int>(blockIdx.x) % n_bh;
  const int h = bh % H, b = bh / H;
  const int C = H * HD;
  const int n_kv = qt + 1;
  const int row_base = b * T;

  if (tid == 0) {
    prefetch_tensormap(&map_qkv);
    mbar_init(q_full, 1);
    mbar_init(&k_full[0], 1);
    mbar_init(&k_full[1], 1);
    mbar_init(v_full, 1);
    mbar_init(s_done, 1);
    mbar_init(o_done, 1);
    fence_barrier_init();
  }
  if (warp == 0) tmem_alloc<1>(tmem_ptr, 256);
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  const uint32_t tmem_row = tmem_base + (static_cast<uint32_t>(warp * 32) << 16);

  if (tid == 0) {
    mbar_arrive_expect_tx(q_full, TILE_BYTES);
    tma_load_4d(sQ, &map_qkv, q_full, h * HD, row_base + qt * TQ, 0, 0);
    mbar_arrive_expect_tx(&k_full[0], TILE_BYTES);
    tma_load_4d(sK, &map_qkv, &k_full[0], C + h * HD, row_base, 0, 0);
    mbar_arrive_expect_tx(v_full, TILE_BYTES);
    tma_load_4d(sV, &map_qkv, v_full, 2 * C + h * HD, row_base, 0, 0);
  }

  const float c = scale * LOG2E;
  float m = -INFINITY, l = 0.f;
  float o[HD];
#pragma unroll
  for (int d = 0; d < HD; ++d) o[d] = 0.f;

  const uint32_t q_lo = desc_lo(smem_u32(sQ), 16);
  const uint32_t p_lo = desc_lo(smem_u32(sP), 16);
  const uint32_t v_lo = desc_lo(smem_u32(sV), 8192);
  // Software pipeline: the S = Q K^T MMAs of tile j+1 are issued right behind the P V MMAs of tile j, so the
  // tensor core works on the next scores while the threads fold P V into O.  Prologue: S of tile 0.
  if (warp == 0) {
    mbar_wait(q_full, 0);
    mbar_wait(&k_full[0], 0);
    tcgen05_fence_after();
    const uint32_t k_lo = desc_lo(smem_u32(sK), 16);
    if (elect_one()) {
      if (n_kv > 1) {
        mbar_arrive_expect_tx(&k_full[1], TILE_BYTES);
        tma_load_4d(sK + TILE_BYTES, &map_qkv, &k_full[1], C + h * HD, row_base + TK, 0, 0);
      }
#pragma unroll
      for (int k = 0; k < HD / 16; ++k) umma_f16<1>(tmem_base, desc_k(q_lo, k), desc_k(k_lo, k), IDESC_S, k > 0);
      umma_commit(s_done);
    }
    __syncwarp();
  }
  for (int j = 0; j < n_kv; ++j) {
    const int st = j & 1;
    const bool more = (j + 1 < n_kv);
    mbar_wait(s_done, j & 1);
    tcgen05_fence_after();

    // the whole 128-column score row of this thread, read from TMEM once
    uint32_t sv[128];
#pragma unroll
    for (int cc = 0; cc < 4; ++cc) {
      uint32_t(&chunk)[32] = *reinterpret_cast<uint32_t(*)[32]>(&sv[cc * 32]);
      tmem_ld_32x32b_x32(tmem_row + cc * 32, chunk);
    }
    tmem_ld_wait();
    // Only the diagonal tile needs the causal mask; every other tile of this row block is fully visible, and a
    // per-element predicate there cost 40 % of the kernel's instructions (ISETP / PLOP3 / FSEL in the ncu source view).
    const bool diag = (j == qt);
    float tile_max = -INFINITY;
    if (diag) {
#pragma unroll
      for (int e = 0; e < 128; ++e)
        if (e <= tid) tile_max = fmaxf(tile_max, __uint_as_float(sv[e]));
    } else {
#pragma unroll
      for (int e = 0; e < 128; ++e) tile_max = fmaxf(tile_max, __uint_as_float(sv[e]));
    }
    const float m_new = fmaxf(m, tile_max);
    const float alpha = fast_exp2((m - m_new) * c);
    const float mc = m_new * c;
    float rowsum = 0.f;
    if (diag) {
#pragma unroll
      for (int q = 0; q < 16; ++q) {
        float p[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const int col = q * 8 + e;
          p[e] = (col <= tid) ? fast_exp2(fmaf(__uint_as_float(sv[col]), c, -mc)) : 0.f;
          rowsum += p[e];
        }
        uint4 pk;
        pk.x = pack2(p[0], p[1]);
        pk.y = pack2(p[2], p[3]);
        pk.z = pack2(p[4], p[5]);
        pk.w = pack2(p[6], p[7]);
        *reinterpret_cast<uint4*>(sP + swz_chunk(tid, q)) = pk;
      }
    } else {
#pragma unroll
      for (int q = 0; q < 16; ++q) {
        float p[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          p[e] = fast_exp2(fmaf(__uint_as_float(sv[q * 8 + e]), c, -mc));
          rowsum += p[e];
        }
        uint4 pk;
        pk.x = pack2(p[0], p[1]);
        pk.y = pack2(p[2], p[3]);
        pk.z = pack2(p[4], p[5]);
        pk.w = pack2(p[6], p[7]);
        *reinterpret_cast<uint4*>(sP + swz_chunk(tid, q)) = pk;
      }
    }
    l = l * alpha + rowsum;
    m = m_new;
    fence_proxy_async_smem();
    tcgen05_fence_before();
    __syncthreads();
    if (warp == 0) {
      mbar_wait(v_full, j & 1);
      if (more) mbar_wait(&k_full[st ^ 1], ((j + 1) >> 1) & 1);
      tcgen05_fence_after();
      const uint32_t kn_lo = desc_lo(smem_u32(sK + (st ^ 1) * TILE_BYTES), 16);
      if (elect_one()) {
#pragma unroll
        for (int k = 0; k < TK / 16; ++k)
          umma_f16<1>(tmem_base + 128, desc_k2(p_lo, k), desc_mn(v_lo, k), IDESC_PV, k > 0);
        umma_commit(o_done);
        if (more) {   // every thread has read S of tile j (the __syncthreads above): its TMEM columns are free
#pragma unroll
          for (int k = 0; k < HD / 16; ++k) umma_f16<1>(tmem_base, desc_k(q_lo, k), desc_k(kn_lo, k), IDESC_S, k > 0);
          umma_commit(s_done);
        }
      }
      __syncwarp();
    }
    mbar_wait(o_done, j & 1);
    tcgen05_fence_after();
    if (warp == 0) {
      // P V of tile j is complete: V's buffer and K stage `st` (read by S of tile j) can be refilled
      if (elect_one()) {
        if (more) {
          mbar_arrive_expect_tx(v_full, TILE_BYTES);
          tma_load_4d(sV, &map_qkv, v_full, 2 * C + h * HD, row_base + (j + 1) * TK, 0, 0);
        }
        if (j + 2 < n_kv) {
          mbar_arrive_expect_tx(&k_full[st], TILE_BYTES);
          tma_load_4d(sK + st * TILE_BYTES, &map_qkv, &k_full[st], C + h * HD, row_base + (j + 2) * TK, 0, 0);
        }
      }
      __syncwarp();
    }
    {
      uint32_t ov[64];
#pragma unroll
      for (int cc = 0; cc < 2; ++cc) {
        uint32_t(&chunk)[32] = *reinterpret_cast<uint32_t(*)[32]>(&ov[cc * 32]);
        tmem_ld_32x32b_x32(tmem_row + 128 + cc * 32, chunk);
      }
      tmem_ld_wait();
#pragma unroll
      for (int e = 0; e < 64; ++e) o[e] = o[e] * alpha + __uint_as_float(ov[e]);
    }
    tcgen05_fence_before();
  }

  const float inv = 1.f / l;
  const long long grow = static_cast<long long>(row_base + qt * TQ + tid);
  __nv_bfloat16* op = out + grow * C + h * HD;
#pragma unroll
  for (int q = 0; q < 8; ++q) {
    uint4 pk;
    pk.x = pack2(o[8 * q + 0] * inv, o[8 * q + 1] * inv);
    pk.y = pack2(o[8 * q + 2] * inv, o[8 * q + 3] * inv);
    pk.z = pack2(o[8 * q + 4] * inv, o[8 * q + 5] * inv);
    pk.w = pack2(o[8 * q + 6] * inv, o[8 * q + 7] * inv);
    *reinterpret_cast<uint4*>(op + 8 * q) = pk;
  }
  lse[(static_cast<long long>(b) * H + h) * T + qt * TQ + tid] = m * scale + logf(l);

  __syncthreads();
  if (warp == 0) {
    tcgen05_fence_after();
    tmem_dealloc<1>(tmem_base, 256);
  }
}

// 32 score columns (chunk cc of a 128-column row) of one query row: maximum, and exponentials against a given reference
// written to the P tile.  DIAG = the tile on the causal diagonal (column <= row only); every other tile takes the
// unmasked instantiation (a per-element predicate on all tiles cost 40 % of the forward's instructions).
template <bool DIAG>
__device__ __forceinline__ float chunk_max(const uint32_t (&x)[32], int cc, int row) {
  // four independent chains: a single 32-deep fmax chain showed up as fixed-latency ("wait") stalls with only two
  // warps per scheduler
  float v[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
  for (int e = 0; e < 32; ++e)
    if (!DIAG || cc * 32 + e <= row) v[e & 3] = fmaxf(v[e & 3], __uint_as_float(x[e]));
  return fmaxf(fmaxf(v[0], v[1]), fmaxf(v[2], v[3]));
}
template <bool DIAG>
__device__ __forceinline__ float chunk_exp(const uint32_t (&x)[32], int cc, int row, float c, float mc, uint8_t* sP) {
  float rowsum = 0.f;
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    float p[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const float v = fast_exp2(fmaf(__uint_as_float(x[q * 8 + e]), c, -mc));
      p[e] = (!DIAG || cc * 32 + q * 8 + e <= row) ? v : 0.f;
    }
    rowsum += ((p[0] + p[1]) + (p[2] + p[3])) + ((p[4] + p[5]) + (p[6] + p[7]));
    uint4 pk;
    pk.x = pack2(p[0], p[1]);
    pk.y = pack2(p[2], p[3]);
    pk.z = pack2(p[4], p[5]);
    pk.w = pack2(p[6], p[7]);
    *reinterpret_cast<uint4*>(sP + swz_chunk(row, cc * 4 + q)) = pk;
  }
  return rowsum;
}

// Same decomposition with O resident in TMEM (P V accumulates there) and lazy rescaling: per tile one mbarrier wait and
// one __syncthreads, no TMEM read-back of O (it cost 512 of the ~1500 TMEM-read cycles per tile at 64 B/clk) and no
// 64 FMAs per thread; the row maximum that P is taken against is refreshed only when it would grow by more than 2^8.
__global__ void __launch_bounds__(128, 2)
fwd_lazy_kernel(const __grid_constant__ CUtensorMap map_qkv, __nv_bfloat16* __restrict__ out, float* __restrict__ lse,
           int T, int H, float scale) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;
  uint8_t* sK = sQ + TILE_BYTES;            // 2 stages
  uint8_t* sV = sK + 2 * TILE_BYTES;        // single buffer: reloaded once P V of the previous tile is done
  uint8_t* sP = sV + TILE_BYTES;            // [128 x 128] bf16
  uint64_t* bars = reinterpret_cast<uint64_t*>(sP + 2 * TILE_BYTES);
  uint64_t* q_full = bars;
  uint64_t* k_full = bars + 1;              // [2]
  uint64_t* v_full = bars + 3;
  uint64_t* s_done = bars + 4;
  uint64_t* pv_done = bars + 5;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 6);

  const int tid = threadIdx.x, warp = tid >> 5;
  // Longest-processing-time-first over the whole grid: CTAs are dispatched in linear order, so a 1-D grid whose first
  // B*H blocks are the longest query rows (n_kv = T/128 tiles), the next B*H the second longest, ... keeps the SMs
  // evenly loaded to the end.  With the (q-tile, head, batch) grid the causal 8,7,...,1 pattern repeated per head and
  // the kernel ran 20 tile-times where the average load is 14 (simulated; 47.5 us measured).
  const int n_bh = gridDim.x / (T / TQ);
  const int qt = T / TQ - 1 - static_cast<int>(blockIdx.x) / n_bh;
  const int bh = static_cast<int>(blockIdx.x) % n_bh;
  const int h = bh % H, b = bh / H;
  const int C = H * HD;
  const int n_kv = qt + 1;
  const int row_base = b * T;

  if (tid == 0) {
    prefetch_tensormap(&map_qkv);
    mbar_init(q_full, 1);
    mbar_init(&k_full[0], 1);
    mbar_init(&k_full[1], 1);
    mbar_init(v_full, 1);
    mbar_init(s_done, 1);
    mbar_init(pv_done, 1);
    fence_barrier_init();
  }
  if (warp == 0) tmem_alloc<1>(tmem_ptr, 256);
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  const uint32_t tmem_row = tmem_base + (static_cast<uint32_t>(warp * 32) << 16);

  if (tid == 0) {
    mbar_arrive_expect_tx(q_full, TILE_BYTES);
    tma_load_4d(sQ, &map_qkv, q_full, h * HD, row_base + qt * TQ, 0, 0);
    mbar_arrive_expect_tx(&k_full[0], TILE_BYTES);
    tma_load_4d(sK, &map_qkv, &k_full[0], C + h * HD, row_base, 0, 0);
    mbar_arrive_expect_tx(v_full, TILE_BYTES);
    tma_load_4d(sV, &map_qkv, v_full, 2 * C + h * HD, row_base, 0, 0);
  }

  const float c = scale * LOG2E;
  float m = -INFINITY, l = 0.f;

  const uint32_t q_lo = desc_lo(smem_u32(sQ), 16);
  const uint32_t p_lo = desc_lo(smem_u32(sP), 16);
  const uint32_t v_lo = desc_lo(smem_u32(sV), 8192);
  // Prologue: S of tile 0.
  if (warp == 0) {
    mbar_wait(q_full, 0);
    mbar_wait(&k_full[0], 0);
    tcgen05_fence_after();
    const uint32_t k_lo = desc_lo(smem_u32(sK), 16);
    if (elect_one()) {
      if (n_kv > 1) {
        mbar_arrive_expect_tx(&k_full[1], TILE_BYTES);
        tma_load_4d(sK + TILE_BYTES, &map_qkv, &k_full[1], C + h * HD, row_base + TK, 0, 0);
      }
#pragma unroll
      for (int k = 0; k < HD / 16; ++k) umma_f16<1>(tmem_base, desc_k(q_lo, k), desc_k(k_lo, k), IDESC_S, k > 0);
      umma_commit(s_done);
    }
    __syncwarp();
  }
  for (int j = 0; j < n_kv; ++j) {
    const int st = j & 1;
    const bool more = (j + 1 < n_kv);
    mbar_wait(s_done, j & 1);
    tcgen05_fence_after();
    // tcgen05.mma ops of one thread complete in issue order: S of tile j done means P V of tile j-1 (issued before it)
    // is done too, so V's buffer, the P buffer and K stage `st` (read by S of tile j) are free from here on.
    if (warp == 0) {
      if (elect_one()) {
        if (j > 0) {
          mbar_arrive_expect_tx(v_full, TILE_BYTES);
          tma_load_4d(sV, &map_qkv, v_full, 2 * C + h * HD, row_base + j * TK, 0, 0);
        }
        if (j + 2 < n_kv) {
          mbar_arrive_expect_tx(&k_full[st], TILE_BYTES);
          tma_load_4d(sK + st * TILE_BYTES, &map_qkv, &k_full[st], C + h * HD, row_base + (j + 2) * TK, 0, 0);
        }
      }
      __syncwarp();
    }
    uint32_t sv[128];
#pragma unroll
    for (int cc = 0; cc < 4; ++cc) {
      uint32_t(&chunk)[32] = *reinterpret_cast<uint32_t(*)[32]>(&sv[cc * 32]);
      tmem_ld_32x32b_x32(tmem_row + cc * 32, chunk);
    }
    tmem_ld_wait();
    const bool diag = (j == qt);
    float tile_max = -INFINITY;
#pragma unroll
    for (int cc = 0; cc < 4; ++cc) {
      const uint32_t(&chunk)[32] = *reinterpret_cast<const uint32_t(*)[32]>(&sv[cc * 32]);
      tile_max = fmaxf(tile_max, diag ? chunk_max<true>(chunk, cc, tid) : chunk_max<false>(chunk, cc, tid));
    }
    // Lazy rescaling: O stays in TMEM and P V accumulates into it; the reference maximum of a row moves only when the
    // new maximum is more than 2^8 above it (P then stays below 256: exact enough in bf16, far from fp32 overflow),
    // so after the first tiles the read-scale-write of O is almost never taken.
    // Tried on top of this and slower at B8 H12 T1024 (43.1 us for this version): exponentials against the stale
    // maximum chunk by chunk under the next chunk's tcgen05.ld (47.1: four tcgen05.wait::ld per tile); S of tile j+1
    // issued right after the score read, before this tile's exponentials, with a second __syncthreads (46.1).
    const float m_new = fmaxf(m, tile_max);
    const bool need = (m_new - m) * c > 8.f;       // true on the first tile (m = -inf)
    float alpha = 1.f;
    if (need) {
      alpha = fast_exp2((m - m_new) * c);          // 0 on the first tile
      m = m_new;
    }
    if (j > 0 && __any_sync(0xffffffffu, need)) {
#pragma unroll
      for (int cc = 0; cc < 4; ++cc) {
        uint32_t ov[16];
        tmem_ld_32x32b_x16(tmem_row + 128 + cc * 16, ov);
        tmem_ld_wait();
#pragma unroll
        for (int e = 0; e < 16; ++e) ov[e] = __float_as_uint(__uint_as_float(ov[e]) * alpha);
        tmem_st_32x32b_x16(tmem_row + 128 + cc * 16, ov);
      }
      tmem_st_wait();
    }
    const float mc = m * c;
    float rowsum = 0.f;
#pragma unroll
    for (int cc = 0; cc < 4; ++cc) {
      const uint32_t(&chunk)[32] = *reinterpret_cast<const uint32_t(*)[32]>(&sv[cc * 32]);
      rowsum += diag ? chunk_exp<true>(chunk, cc, tid, c, mc, sP) : chunk_exp<false>(chunk, cc, tid, c, mc, sP);
    }
    l = l * alpha + rowsum;
    fence_proxy_async_smem();
    tcgen05_fence_before();
    __syncthreads();
    if (warp == 0) {
      mbar_wait(v_full, j & 1);
      if (more) mbar_wait(&k_full[st ^ 1], ((j + 1) >> 1) & 1);
      tcgen05_fence_after();
      const uint32_t kn_lo = desc_lo(smem_u32(sK + (st ^ 1) * TILE_BYTES), 16);
      if (elect_one()) {
#pragma unroll
        for (int k = 0; k < TK / 16; ++k)
          umma_f16<1>(tmem_base + 128, desc_k2(p_lo, k), desc_mn(v_lo, k), IDESC_PV, (k > 0) || (j > 0));
        if (more) {   // every thread has read S of tile j (the __syncthreads above): its TMEM columns are free
#pragma unroll
          for (int k = 0; k < HD / 16; ++k) umma_f16<1>(tmem_base, desc_k(q_lo, k), desc_k(kn_lo, k), IDESC_S, k > 0);
          umma_commit(s_done);
        } else {
          umma_commit(pv_done);
        }
      }
      __syncwarp();
    }
  }
  mbar_wait(pv_done, 0);
  tcgen05_fence_after();
  float o[HD];
  {
    uint32_t ov[64];
#pragma unroll
    for (int cc = 0; cc < 2; ++cc) {
      uint32_t(&chunk)[32] = *reinterpret_cast<uint32_t(*)[32]>(&ov[cc * 32]);
      tmem_ld_32x32b_x32(tmem_row + 128 + cc * 32, chunk);
    }
    tmem_ld_wait();
#pragma unroll
    for (int e = 0; e < 64; ++e) o[e] = __uint_as_float(ov[e]);
  }
  tcgen05_fence_before();

  const float inv = 1.f / l;
  const long long grow = static_cast<long long>(row_base + qt * TQ + tid);
  __nv_bfloat16* op = out + grow * C + h * HD;
#pragma unroll
  for (int q = 0; q < 8; ++q) {
    uint4 pk;
    pk.x = pack2(o[8 * q + 0] * inv, o[8 * q + 1] * inv);
    pk.y = pack2(o[8 * q + 2] * inv, o[8 * q + 3] * inv);
    pk.z = pack2(o[8 * q + 4] * inv, o[8 * q + 5] * inv);
    pk.w = pack2(o[8 * q + 6] * inv, o[8 * q + 7] * inv);
    *reinterpret_cast<uint4*>(op + 8 * q) = pk;
  }
  lse[(static_cast<long long>(b) * H + h) * T + qt * TQ + tid] = m * scale + logf(l);

  __syncthreads();
  if (warp == 0) {
    tcgen05_fence_after();
    tmem_dealloc<1>(tmem_base, 256);
  }
}

// -------------------------------------------------------------------------- backward
// delta[b,h,t] = sum_d dO * O.  Eight lanes own one (b,t,h) row of 64 elements (16 bytes each), so a warp covers four
// consecutive heads of a token with fully coalesced 512-byte requests and three shuffle steps (the first version used
// a whole warp and 4-byte loads per row: 18 us for 25 MB).
__global__ void __launch_bounds__(256)
bwd_prep_kernel(const __nv_bfloat16* __restrict__ o, const __nv_bfloat16* __restrict__ dout, float* __restrict__ delta,
                int BT, int T, int H) {
  const long long g = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) >> 3;   // (bt, h) row index
  const int sub = threadIdx.x & 7;
  const long long rows = static_cast<long long>(BT) * H;
  float s = 0.f;
  if (g < rows) {
    const long long off = g * HD + sub * 8;
    const uint4 a = *reinterpret_cast<const uint4*>(o + off);
    const uint4 d = *reinterpret_cast<const uint4*>(dout + off);
    const __nv_bfloat162* ah = reinterpret_cast<const __nv_bfloat162*>(&a);
    const __nv_bfloat162* dh = reinterpret_cast<const __nv_bfloat162*>(&d);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float2 x = __bfloat1622float2(ah[i]), y = __bfloat1622float2(dh[i]);
      s += x.x * y.x + x.y * y.y;
    }
  }
#pragma unroll
  for (int w = 4; w > 0; w >>= 1) s += __shfl_xor_sync(0xffffffffu, s, w);
  if (sub == 0 && g < rows) {
    const int bt = static_cast<int>(g / H), h = static_cast<int>(g % H);
    delta[(static_cast<long long>(bt / T) * H + h) * T + (bt % T)] = s;
  }
}

// dq_acc [B,H,T,64] fp32 -> dqkv[b,t,h*64+d] bf16
__global__ void __launch_bounds__(256)
bwd_finish_kernel(const float* __restrict__ dq_acc, __nv_bfloat16* __restrict__ dqkv, int B, int T, int H) {
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;   // one thread per 8 elements
  const long long total = static_cast<long long>(B) * H * T * (HD / 8);
  if (i >= total) return;
  const int d8 = static_cast<int>(i % (HD / 8));
  const long long r = i / (HD / 8);
  const int t = static_cast<int>(r % T);
  const int h = static_cast<int>((r / T) % H);
  const int b = static_cast<int>(r / (static_cast<long long>(T) * H));
  const float4 x = *reinterpret_cast<const float4*>(dq_acc + r * HD + d8 * 8);
  const float4 y = *reinterpret_cast<const float4*>(dq_acc + r * HD + d8 * 8 + 4);
  uint4 pk;
  pk.x = pack2(x.x, x.y); pk.y = pack2(x.z, x.w); pk.z = pack2(y.x, y.y); pk.w = pack2(y.z, y.w);
  *reinterpret_cast<uint4*>(dqkv + (static_cast<long long>(b) * T + t) * (3 * H * HD) + h * HD + d8 * 8) = pk;
}

constexpr int BWD_SMEM = TILE_BYTES * (1 + 1 + 2 + 2 + 2 + 2) + 1024 + 128;

__global__ void __launch_bounds__(128, 1)
bwd_kernel(const __grid_constant__ CUtensorMap map_qkv, const __grid_constant__ CUtensorMap map_do,
           const float* __restrict__ lse, const float* __restrict__ delta, float* __restrict__ dq_acc,
           __nv_bfloat16* __restrict__ dqkv, int T, int H, float scale) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sK = smem;
  uint8_t* sV = sK + TILE_BYTES;
  uint8_t* sQ = sV + TILE_BYTES;             // 2 stages
  uint8_t* sdO = sQ + 2 * TILE_BYTES;        // 2 stages
  uint8_t* sP = sdO + 2 * TILE_BYTES;        // [128 q x 128 kv]
  uint8_t* sdS = sP + 2 * TILE_BYTES;        // [128 q x 128 kv]
  uint64_t* bars = reinterpret_cast<uint64_t*>(sdS + 2 * TILE_BYTES);
  uint64_t* kv_full = bars;
  uint64_t* qdo_full = bars + 1;             // [2]
  uint64_t* sdp_done = bars + 3;
  uint64_t* grads_done = bars + 4;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 5);

  const int tid = threadIdx.x, warp = tid >> 5;
  // 1-D grid, longest key tiles first over all (batch, head) pairs (see fwd_kernel)
  const int n_bh = gridDim.x / (T / TK);
  const int jt = static_cast<int>(blockIdx.x) / n_bh;
  const int bh = static_cast<int>(blockIdx.x) % n_bh;
  const int h = bh % H, b = bh / H;
  const int C = H * HD;
  const int nT = T / TQ;
  const int n_q = nT - jt;
  const int row_base = b * T;
  constexpr uint32_t T_S = 0, T_DP = 128, T_DV = 256, T_DK = 320, T_DQ = 384;

  if (tid == 0) {
    prefetch_tensormap(&map_qkv);
    prefetch_tensormap(&map_do);
    mbar_init(kv_full, 1);
    mbar_init(&qdo_full[0], 1);
    mbar_init(&qdo_full[1], 1);
    mbar_init(sdp_done, 1);
    mbar_init(grads_done, 1);
    fence_barrier_init();
  }
  if (warp == 0) tmem_alloc<1>(tmem_ptr, 512);
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  const uint32_t tmem_row = tmem_base + (static_cast<uint32_t>(warp * 32) << 16);

  if (tid == 0) {
    mbar_arrive_expect_tx(kv_full, 2 * TILE_BYTES);
    tma_load_4d(sK, &map_qkv, kv_full, C + h * HD, row_base + jt * TK, 0, 0);
    tma_load_4d(sV, &map_qkv, kv_full, 2 * C + h * HD, row_base + jt * TK, 0, 0);
    mbar_arrive_expect_tx(&qdo_full[0], 2 * TILE_BYTES);
    tma_load_4d(sQ, &map_qkv, &qdo_full[0], h * HD, row_base + jt * TQ, 0, 0);
    tma_load_4d(sdO, &map_do, &qdo_full[0], h * HD, row_base + jt * TQ, 0, 0);
  }
  const float c = scale * LOG2E;
  const long long stat_base = (static_cast<long long>(b) * H + h) * T;

  const uint32_t k_lo = desc_lo(smem_u32(sK), 16), v_lo = desc_lo(smem_u32(sV), 16);
  const uint32_t kmn_lo = desc_lo(smem_u32(sK), 8192);
  const uint32_t p_mn_lo = desc_lo(smem_u32(sP), TILE_BYTES), ds_mn_lo = desc_lo(smem_u32(sdS), TILE_BYTES);
  const uint32_t ds_k_lo = desc_lo(smem_u32(sdS), 16);
  // Software pipeline: S and dP of query tile i+1 are issued right behind the dV/dK/dQ MMAs of tile i.
  if (warp == 0) {
    mbar_wait(kv_full, 0);
    mbar_wait(&qdo_full[0], 0);
    tcgen05_fence_after();
    const uint32_t q_lo = desc_lo(smem_u32(sQ), 16), do_lo = desc_lo(smem_u32(sdO), 16);
    if (elect_one()) {
      if (n_q > 1) {
        mbar_arrive_expect_tx(&qdo_full[1], 2 * TILE_BYTES);
        tma_load_4d(sQ + TILE_BYTES, &map_qkv, &qdo_full[1], h * HD, row_base + (jt + 1) * TQ, 0, 0);
        tma_load_4d(sdO + TILE_BYTES, &map_do, &qdo_full[1], h * HD, row_base + (jt + 1) * TQ, 0, 0);
      }
#pragma unroll
      for (int k = 0; k < HD / 16; ++k) umma_f16<1>(tmem_base + T_S, desc_k(q_lo, k), desc_k(k_lo, k), IDESC_S, k > 0);
#pragma unroll
      for (int k = 0; k < HD / 16; ++k) umma_f16<1>(tmem_base + T_DP, desc_k(do_lo, k), desc_k(v_lo, k), IDESC_S, k > 0);
      umma_commit(sdp_done);
    }
    __syncwarp();
  }
  // lse / delta of the next q-tile are fetched one iteration ahead: loaded at the point of use they were the kernel's
  // largest stall (27 % of the sampled warp cycles sat on that load)
  float next_lse = lse[stat_base + jt * TQ + tid], next_delta = delta[stat_base + jt * TQ + tid];
  for (int ii = 0; ii < n_q; ++ii) {
    const int it = jt + ii, st = ii & 1;
    const bool more = (ii + 1 < n_q);
    const float row_lse = next_lse * LOG2E;
    const float row_delta = next_delta;
    if (more) {
      next_lse = lse[stat_base + (it + 1) * TQ + tid];
      next_delta = delta[stat_base + (it + 1) * TQ + tid];
    }
    mbar_wait(sdp_done, ii & 1);
    tcgen05_fence_after();
    const bool diag = (it == jt);
#pragma unroll 1
    for (int cc = 0; cc < 4; ++cc) {
      uint32_t vs[32], vd[32];
      tmem_ld_32x32b_x32(tmem_row + T_S + cc * 32, vs);
      tmem_ld_32x32b_x32(tmem_row + T_DP + cc * 32, vd);
      tmem_ld_wait();
      float p[32], ds[32];
      if (diag) {     // causal mask on the diagonal tile only (uniform branch, see the forward kernel)
#pragma unroll
        for (int e = 0; e < 32; ++e) {
          p[e] = (cc * 32 + e <= tid) ? fast_exp2(fmaf(__uint_as_float(vs[e]), c, -row_lse)) : 0.f;
          ds[e] = p[e] * (__uint_as_float(vd[e]) - row_delta) * scale;
        }
      } else {
#pragma unroll
        for (int e = 0; e < 32; ++e) {
          p[e] = fast_exp2(fmaf(__uint_as_float(vs[e]), c, -row_lse));
          ds[e] = p[e] * (__uint_as_float(vd[e]) - row_delta) * scale;
        }
      }
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        uint4 pk, dk;
        pk.x = pack2(p[8 * q + 0], p[8 * q + 1]); pk.y = pack2(p[8 * q + 2], p[8 * q + 3]);
        pk.z = pack2(p[8 * q + 4], p[8 * q + 5]); pk.w = pack2(p[8 * q + 6], p[8 * q + 7]);
        dk.x = pack2(ds[8 * q + 0], ds[8 * q + 1]); dk.y = pack2(ds[8 * q + 2], ds[8 * q + 3]);
        dk.z = pack2(ds[8 * q + 4], ds[8 * q + 5]); dk.w = pack2(ds[8 * q + 6], ds[8 * q + 7]);
        const uint32_t off = swz_chunk(tid, cc * 4 + q);
        *reinterpret_cast<uint4*>(sP + off) = pk;
        *reinterpret_cast<uint4*>(sdS + off) = dk;
      }
    }
    fence_proxy_async_smem();
    tcgen05_fence_before();
    __syncthreads();
    if (warp == 0) {
      if (more) mbar_wait(&qdo_full[st ^ 1], ((ii + 1) >> 1) & 1);
      tcgen05_fence_after();
      const uint32_t qmn_lo = desc_lo(smem_u32(sQ + st * TILE_BYTES), 8192);
      const uint32_t domn_lo = desc_lo(smem_u32(sdO + st * TILE_BYTES), 8192);
      const uint32_t qn_lo = desc_lo(smem_u32(sQ + (st ^ 1) * TILE_BYTES), 16);
      const uint32_t don_lo = desc_lo(smem_u32(sdO + (st ^ 1) * TILE_BYTES), 16);
      if (elect_one()) {
        // dV[kv,d] += sum_q P[q,kv] dO[q,d]      A = P^T (MN-major, kv blocks 16 KB apart), B = dO (MN-major)
#pragma unroll
        for (int k = 0; k < TQ / 16; ++k)
          umma_f16<1>(tmem_base + T_DV, desc_mn(p_mn_lo, k), desc_mn(domn_lo, k), IDESC_TT, (ii | k) != 0);
        // dK[kv,d] += sum_q dS[q,kv] Q[q,d]
#pragma unroll
        for (int k = 0; k < TQ / 16; ++k)
          umma_f16<1>(tmem_base + T_DK, desc_mn(ds_mn_lo, k), desc_mn(qmn_lo, k), IDESC_TT, (ii | k) != 0);
        // dQ[q,d] = sum_kv dS[q,kv] K[kv,d]       A = dS (K-major, two K blocks), B = K (MN-major)
#pragma unroll
        for (int k = 0; k < TK / 16; ++k)
          umma_f16<1>(tmem_base + T_DQ, desc_k2(ds_k_lo, k), desc_mn(kmn_lo, k), IDESC_PV, k > 0);
        umma_commit(grads_done);
        if (more) {   // S / dP columns were fully read before the __syncthreads above
#pragma unroll
          for (int k = 0; k < HD / 16; ++k)
            umma_f16<1>(tmem_base + T_S, desc_k(qn_lo, k), desc_k(k_lo, k), IDESC_S, k > 0);
#pragma unroll
          for (int k = 0; k < HD / 16; ++k)
            umma_f16<1>(tmem_base + T_DP, desc_k(don_lo, k), desc_k(v_lo, k), IDESC_S, k > 0);
          umma_commit(sdp_done);
        }
      }
      __syncwarp();
    }
    mbar_wait(grads_done, ii & 1);
    tcgen05_fence_after();
    if (warp == 0 && ii + 2 < n_q) {
      // the MMAs that read Q / dO of tile i are done: refill this stage with tile i+2
      if (elect_one()) {
        mbar_arrive_expect_tx(&qdo_full[st], 2 * TILE_BYTES);
        tma_load_4d(sQ + st * TILE_BYTES, &map_qkv, &qdo_full[st], h * HD, row_base + (it + 2) * TQ, 0, 0);
        tma_load_4d(sdO + st * TILE_BYTES, &map_do, &qdo_full[st], h * HD, row_base + (it + 2) * TQ, 0, 0);
      }
      __syncwarp();
    }
    float* dq_row = dq_acc + (stat_base + it * TQ + tid) * HD;
#pragma unroll
    for (int cc = 0; cc < 2; ++cc) {
      uint32_t v[32];
      tmem_ld_32x32b_x32(tmem_row + T_DQ + cc * 32, v);
      tmem_ld_wait();
#pragma unroll
      for (int e = 0; e < 32; e += 4)
        red_add_v4(dq_row + cc * 32 + e, __uint_as_float(v[e]), __uint_as_float(v[e + 1]), __uint_as_float(v[e + 2]),
                   __uint_as_float(v[e + 3]));
    }
    tcgen05_fence_before();
  }

  // dK / dV of this key tile: TMEM lanes are key rows
  const long long grow = static_cast<long long>(row_base + jt * TK + tid);
  __nv_bfloat16* dk_out = dqkv + grow * (3 * C) + C + h * HD;
  __nv_bfloat16* dv_out = dqkv + grow * (3 * C) + 2 * C + h * HD;
#pragma unroll
  for (int which = 0; which < 2; ++which) {
    __nv_bfloat16* dst = which == 0 ? dv_out : dk_out;
    const uint32_t col = which == 0 ? T_DV : T_DK;
#pragma unroll
    for (int cc = 0; cc < 2; ++cc) {
      uint32_t v[32];
      tmem_ld_32x32b_x32(tmem_row + col + cc * 32, v);
      tmem_ld_wait();
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        uint4 pk;
        pk.x = pack2(__uint_as_float(v[8 * q + 0]), __uint_as_float(v[8 * q + 1]));
        pk.y = pack2(__uint_as_float(v[8 * q + 2]), __uint_as_float(v[8 * q + 3]));
        pk.z = pack2(__uint_as_float(v[8 * q + 4]), __uint_as_float(v[8 * q + 5]));
        pk.w = pack2(__uint_as_float(v[8 * q + 6]), __uint_as_float(v[8 * q + 7]));
        *reinterpret_cast<uint4*>(dst + cc * 32 + q * 8) = pk;
      }
    }
  }
  tcgen05_fence_before();
  __syncthreads();
  if (warp == 0) {
    tcgen05_fence_after();
    tmem_dealloc<1>(tmem_base, 512);
  }
}

// -------------------------------------------------------------------------- backward, warp-specialised
// Same math and the same TMEM plan as bwd_kernel, but the jobs of an iteration run in different warps and overlap
// (ncu on bwd_kernel: 6 % warps active, 16 % issue slots used, tensor pipe 16 %: one 4-warp CTA per SM spent its time
// waiting on its own MMA -> TMEM -> softmax -> MMA chain):
//   warps 0-15  "softmax": warp w owns TMEM lane quadrant w & 3 (32 query rows) and 32 of the 128 key columns (w >> 2):
//               P = exp2(S*c - lse), dS = P (dP - delta) scale, both written to shared memory as bf16
//   warps 16-19 "dQ": drain the dQ accumulator of the previous step: TMEM -> fp32 tile in shared memory -> ONE bulk
//               reduce-add (cp.reduce.async.bulk, 32 KB) into the global accumulator.  Per-thread red.global.v4 cost
//               22 us of a 134 us backward (timing experiment with the atomics disabled)
//   warp 20     issues every TMA load (3-stage Q / dO ring) and every tcgen05.mma
// Hand-offs are mbarriers (no __syncthreads in the loop):
//   sdp_done (MMA commit)       -> softmax: S, dP of this tile are in TMEM
//   sdp_read (512 arrivals)     -> MMA warp: every softmax thread holds its S / dP values in registers; the NEXT tile's
//                                  S / dP MMAs go out now, under this tile's exponentials
//   pds_ready (512 arrivals)    -> MMA warp: P, dS of this tile are in shared memory: dV, dK, dQ MMAs
//   grads_done (MMA commit)     -> dQ warps: dQ is complete; softmax: P / dS buffers may be overwritten; MMA warp: the
//                                  Q / dO stage may be refilled (tile + 3)
//   dq_free (128 arrivals)      -> MMA warp: the dQ columns have been read
// With everything but the S / dP MMAs and the hand-offs disabled the first warp-specialised version still took
// 2.8 us per iteration: the 2-stage Q / dO ring exposed the TMA latency and S / dP(i+1) waited for the stores of tile i.
constexpr int BWD_SOFTMAX_WARPS = 16;        // 4 per TMEM lane quadrant: 32 key columns each
constexpr int BWD_WS_THREADS = (BWD_SOFTMAX_WARPS + 4 + 1) * 32;
constexpr int BWD_WS_STAGES = 3;
// K, V, 3 x (Q, dO), P, dS, fp32 dQ staging tile
constexpr int BWD_WS_SMEM = TILE_BYTES * (1 + 1 + 2 * BWD_WS_STAGES + 2 + 2 + 2) + 1024 + 256;

__device__ __forceinline__ void bulk_reduce_add_f32(float* gdst, uint32_t ssrc, uint32_t bytes) {
  asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f32 [%0], [%1], %2;" ::"l"(gdst), "r"(ssrc),
               "r"(bytes)
               : "memory");
}

__global__ void __launch_bounds__(BWD_WS_THREADS, 1)
bwd_ws_kernel(const __grid_constant__ CUtensorMap map_qkv, const __grid_constant__ CUtensorMap map_do,
              const float* __restrict__ lse, const float* __restrict__ delta, float* __restrict__ dq_acc,
              __nv_bfloat16* __restrict__ dqkv, int T, int H, float scale, int dbg) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sK = smem;
  uint8_t* sV = sK + TILE_BYTES;
  uint8_t* sQ = sV + TILE_BYTES;                          // BWD_WS_STAGES stages
  uint8_t* sdO = sQ + BWD_WS_STAGES * TILE_BYTES;         // BWD_WS_STAGES stages
  uint8_t* sP = sdO + BWD_WS_STAGES * TILE_BYTES;         // [128 q x 128 kv] bf16
  uint8_t* sdS = sP + 2 * TILE_BYTES;                     // [128 q x 128 kv] bf16
  uint8_t* sdQ = sdS + 2 * TILE_BYTES;                    // [128 q x 64] fp32 staging for the bulk reduce
  uint64_t* bars = reinterpret_cast<uint64_t*>(sdQ + 2 * TILE_BYTES);
  uint64_t* kv_full = bars;
  uint64_t* qdo_full = bars + 1;             // [BWD_WS_STAGES]
  uint64_t* sdp_done = bars + 4;
  uint64_t* sdp_read = bars + 5;
  uint64_t* pds_ready = bars + 6;
  uint64_t* grads_done = bars + 7;
  uint64_t* dq_free = bars + 8;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 9);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int n_bh = gridDim.x / (T / TK);
  const int jt = static_cast<int>(blockIdx.x) / n_bh;      // longest key tiles first (see fwd_kernel)
  const int bh = static_cast<int>(blockIdx.x) % n_bh;
  const int h = bh % H, b = bh / H;
  const int C = H * HD;
  const int nT = T / TQ;
  const int n_q = nT - jt;
  const int row_base = b * T;
  constexpr uint32_t T_S = 0, T_DP = 128, T_DV = 256, T_DK = 320, T_DQ = 384;

  if (tid == 0) {
    prefetch_tensormap(&map_qkv);
    prefetch_tensormap(&map_do);
    mbar_init(kv_full, 1);
#pragma unroll
    for (int q = 0; q < BWD_WS_STAGES; ++q) mbar_init(&qdo_full[q], 1);
    mbar_init(sdp_done, 1);
    mbar_init(sdp_read, BWD_SOFTMAX_WARPS * 32);
    mbar_init(pds_ready, BWD_SOFTMAX_WARPS * 32);
    mbar_init(grads_done, 1);
    mbar_init(dq_free, 128);
    fence_barrier_init();
  }
  if (warp == 0) tmem_alloc<1>(tmem_ptr, 512);
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  const float c = scale * LOG2E;
  const long long stat_base = (static_cast<long long>(b) * H + h) * T;

  if (warp == BWD_SOFTMAX_WARPS + 4) {
    // ====================================== TMA + MMA issue ======================================
    const uint32_t k_lo = desc_lo(smem_u32(sK), 16), v_lo = desc_lo(smem_u32(sV), 16);
    const uint32_t kmn_lo = desc_lo(smem_u32(sK), 8192);
    const uint32_t p_mn_lo = desc_lo(smem_u32(sP), TILE_BYTES), ds_mn_lo = desc_lo(smem_u32(sdS), TILE_BYTES);
    const uint32_t ds_k_lo = desc_lo(smem_u32(sdS), 16);
    auto load_qdo = [&](int tile, int stage) {       // one elected lane
      mbar_arrive_expect_tx(&qdo_full[stage], 2 * TILE_BYTES);
      tma_load_4d(sQ + stage * TILE_BYTES, &map_qkv, &qdo_full[stage], h * HD, row_base + (jt + tile) * TQ, 0, 0);
      tma_load_4d(sdO + stage * TILE_BYTES, &map_do, &qdo_full[stage], h * HD, row_base + (jt + tile) * TQ, 0, 0);
    };
    auto issue_sdp = [&](int stage) {                // one elected lane: S = Q K^T, dP = dO V^T
      const uint32_t q_lo = desc_lo(smem_u32(sQ + stage * TILE_BYTES), 16);
      const uint32_t do_lo = desc_lo(smem_u32(sdO + stage * TILE_BYTES), 16);
#pragma unroll
      for (int k = 0; k < HD / 16; ++k) umma_f16<1>(tmem_base + T_S, desc_k(q_lo, k), desc_k(k_lo, k), IDESC_S, k > 0);
#pragma unroll
      for (int k = 0; k < HD / 16; ++k) umma_f16<1>(tmem_base + T_DP, desc_k(do_lo, k), desc_k(v_lo, k), IDESC_S, k > 0);
      umma_commit(sdp_done);
    };
    if (elect_one()) {
      mbar_arrive_expect_tx(kv_full, 2 * TILE_BYTES);
      tma_load_4d(sK, &map_qkv, kv_full, C + h * HD, row_base + jt * TK, 0, 0);
      tma_load_4d(sV, &map_qkv, kv_full, 2 * C + h * HD, row_base + jt * TK, 0, 0);
      for (int t = 0; t < BWD_WS_STAGES && t < n_q; ++t) load_qdo(t, t);
    }
    __syncwarp();
    mbar_wait(kv_full, 0);
    mbar_wait(&qdo_full[0], 0);
    tcgen05_fence_after();
    if (elect_one()) issue_sdp(0);
    __syncwarp();
    for (int ii = 0; ii < n_q; ++ii) {
      const int st = ii % BWD_WS_STAGES;
      const bool more = (ii + 1 < n_q);
      if (more) {
        // every softmax thread has S / dP(ii) in registers: the next tile's scores go out under this tile's softmax
        const int sn = (ii + 1) % BWD_WS_STAGES;
        mbar_wait(sdp_read, ii & 1);
        mbar_wait(&qdo_full[sn], ((ii + 1) / BWD_WS_STAGES) & 1);
        tcgen05_fence_after();
        if (elect_one()) issue_sdp(sn);
        __syncwarp();
      }
      mbar_wait(pds_ready, ii & 1);            // P / dS(ii) are in shared memory
      tcgen05_fence_after();
      const uint32_t qmn_lo = desc_lo(smem_u32(sQ + st * TILE_BYTES), 8192);
      const uint32_t domn_lo = desc_lo(smem_u32(sdO + st * TILE_BYTES), 8192);
      if (elect_one() && !(dbg & 2)) {
        // dV[kv,d] += sum_q P[q,kv] dO[q,d]      A = P^T (MN-major, kv blocks 16 KB apart), B = dO (MN-major)
#pragma unroll
        for (int k = 0; k < TQ / 16; ++k)
          umma_f16<1>(tmem_base + T_DV, desc_mn(p_mn_lo, k), desc_mn(domn_lo, k), IDESC_TT, (ii | k) != 0);
        // dK[kv,d] += sum_q dS[q,kv] Q[q,d]
#pragma unroll
        for (int k = 0; k < TQ / 16; ++k)
          umma_f16<1>(tmem_base + T_DK, desc_mn(ds_mn_lo, k), desc_mn(qmn_lo, k), IDESC_TT, (ii | k) != 0);
      }
      __syncwarp();
      if (ii > 0) {                            // the dQ warps have drained the previous tile's dQ columns
        mbar_wait(dq_free, (ii - 1) & 1);
        tcgen05_fence_after();
      }
      if (elect_one()) {
        // dQ[q,d] = sum_kv dS[q,kv] K[kv,d]       A = dS (K-major, two K blocks), B = K (MN-major)
        if (!(dbg & 4)) {
#pragma unroll
          for (int k = 0; k < TK / 16; ++k)
            umma_f16<1>(tmem_base + T_DQ, desc_k2(ds_k_lo, k), desc_mn(kmn_lo, k), IDESC_PV, k > 0);
        }
        umma_commit(grads_done);
      }
      __syncwarp();
      if (ii + BWD_WS_STAGES < n_q) {
        // the MMAs that read Q / dO of tile ii are done once grads_done flips: refill this stage
        mbar_wait(grads_done, ii & 1);
        if (elect_one()) load_qdo(ii + BWD_WS_STAGES, st);
        __syncwarp();
      }
    }
  }
  else if (warp < BWD_SOFTMAX_WARPS) {
    // ========================================= softmax / dS =========================================
    const int quad = warp & 3, part = warp >> 2;            // part: which 32 of the tile's 128 key columns
    const int row = quad * 32 + lane;                       // query row inside the tile = TMEM lane
    const uint32_t tmem_row = tmem_base + (static_cast<uint32_t>(quad * 32) << 16);
    const uint32_t sP_u32 = smem_u32(sP), sdS_u32 = smem_u32(sdS);
    const int col0 = part * 32;
    float next_lse = lse[stat_base + jt * TQ + row], next_delta = delta[stat_base + jt * TQ + row];
    for (int ii = 0; ii < n_q; ++ii) {
      const int it = jt + ii;
      const float row_lse = next_lse * LOG2E;
      const float row_delta = next_delta;
      if (ii + 1 < n_q) {
        next_lse = lse[stat_base + (it + 1) * TQ + row];
        next_delta = delta[stat_base + (it + 1) * TQ + row];
      }
      mbar_wait(sdp_done, ii & 1);
      tcgen05_fence_after();
      uint32_t vs[32], vd[32];
      tmem_ld_32x32b_x32(tmem_row + T_S + col0, vs);
      tmem_ld_32x32b_x32(tmem_row + T_DP + col0, vd);
      tmem_ld_wait();
      tcgen05_fence_before();
      mbar_arrive(sdp_read);                                 // S / dP columns may be overwritten by the next tile
      const bool diag = (it == jt);
      uint32_t pp[16], dd[16];                               // bf16 pairs
      if (dbg & 1) {
#pragma unroll
        for (int e = 0; e < 16; ++e) { pp[e] = vs[e]; dd[e] = vd[e]; }
      } else if (diag) {     // causal mask on the diagonal tile only (uniform branch)
#pragma unroll
        for (int e = 0; e < 32; e += 2) {
          const float p0 = (col0 + e <= row) ? fast_exp2(fmaf(__uint_as_float(vs[e]), c, -row_lse)) : 0.f;
          const float p1 = (col0 + e + 1 <= row) ? fast_exp2(fmaf(__uint_as_float(vs[e + 1]), c, -row_lse)) : 0.f;
          pp[e >> 1] = pack2(p0, p1);
          dd[e >> 1] = pack2(p0 * (__uint_as_float(vd[e]) - row_delta) * scale,
                             p1 * (__uint_as_float(vd[e + 1]) - row_delta) * scale);
        }
      } else {
#pragma unroll
        for (int e = 0; e < 32; e += 2) {
          const float p0 = fast_exp2(fmaf(__uint_as_float(vs[e]), c, -row_lse));
          const float p1 = fast_exp2(fmaf(__uint_as_float(vs[e + 1]), c, -row_lse));
          pp[e >> 1] = pack2(p0, p1);
          dd[e >> 1] = pack2(p0 * (__uint_as_float(vd[e]) - row_delta) * scale,
                             p1 * (__uint_as_float(vd[e + 1]) - row_delta) * scale);
        }
      }
      if (ii > 0) mbar_wait(grads_done, (ii - 1) & 1);       // the MMAs that read P / dS of tile ii-1 are done
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const uint32_t off = swz_chunk(row, part * 4 + q);
        st_shared_v4(sP_u32 + off, make_uint4(pp[4 * q], pp[4 * q + 1], pp[4 * q + 2], pp[4 * q + 3]));
        st_shared_v4(sdS_u32 + off, make_uint4(dd[4 * q], dd[4 * q + 1], dd[4 * q + 2], dd[4 * q + 3]));
      }
      fence_proxy_async_smem();
      mbar_arrive(pds_ready);
    }
    // dV (parts 0, 1) / dK (parts 2, 3) of this key tile, 32 columns per warp: TMEM lanes are key rows
    mbar_wait(grads_done, (n_q - 1) & 1);
    tcgen05_fence_after();
    const long long grow = static_cast<long long>(row_base + jt * TK + row);
    __nv_bfloat16* dst = dqkv + grow * (3 * C) + (part < 2 ? 2 * C : C) + h * HD + (part & 1) * 32;
    uint32_t v[32];
    tmem_ld_32x32b_x32(tmem_row + (part < 2 ? T_DV : T_DK) + (part & 1) * 32, v);
    tmem_ld_wait();
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      uint4 pk;
      pk.x = pack2(__uint_as_float(v[8 * q + 0]), __uint_as_float(v[8 * q + 1]));
      pk.y = pack2(__uint_as_float(v[8 * q + 2]), __uint_as_float(v[8 * q + 3]));
      pk.z = pack2(__uint_as_float(v[8 * q + 4]), __uint_as_float(v[8 * q + 5]));
      pk.w = pack2(__uint_as_float(v[8 * q + 6]), __uint_as_float(v[8 * q + 7]));
      *reinterpret_cast<uint4*>(dst + q * 8) = pk;
    }
  } else {
    // ============================================ dQ drain ============================================
    const int quad = warp & 3;
    const int row = quad * 32 + lane;
    const uint32_t tmem_row = tmem_base + (static_cast<uint32_t>(quad * 32) << 16);
    const uint32_t sdQ_u32 = smem_u32(sdQ);
    const bool issuer = (warp == BWD_SOFTMAX_WARPS) && (lane == 0);
    for (int ii = 0; ii < n_q; ++ii) {
      const int it = jt + ii;
      mbar_wait(grads_done, ii & 1);
      tcgen05_fence_after();
      uint32_t v0[32], v1[32];
      tmem_ld_32x32b_x32(tmem_row + T_DQ, v0);
      tmem_ld_32x32b_x32(tmem_row + T_DQ + 32, v1);
      tmem_ld_wait();
      tcgen05_fence_before();
      mbar_arrive(dq_free);                    // the next dQ MMA may overwrite the columns
      if (dbg & 8) continue;
      // the previous bulk reduce must have finished READING the staging tile before it is overwritten
      if (issuer) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
      asm volatile("bar.sync 1, 128;" ::: "memory");
      // plain row-major [128][64] fp32: the bulk reduce adds the tile to global memory byte for byte, so no swizzle
      // (the lanes of a warp write rows 256 B apart: bank conflicts, but in warps that are off the critical path)
      const uint32_t rbase = sdQ_u32 + row * 256;
#pragma unroll
      for (int q = 0; q < 8; ++q)
        st_shared_v4(rbase + q * 16, make_uint4(v0[4 * q], v0[4 * q + 1], v0[4 * q + 2], v0[4 * q + 3]));
#pragma unroll
      for (int q = 0; q < 8; ++q)
        st_shared_v4(rbase + (q + 8) * 16, make_uint4(v1[4 * q], v1[4 * q + 1], v1[4 * q + 2], v1[4 * q + 3]));
      fence_proxy_async_smem();
      asm volatile("bar.sync 1, 128;" ::: "memory");
      if (issuer) {
        bulk_reduce_add_f32(dq_acc + (stat_base + it * TQ) * HD, sdQ_u32, TQ * HD * 4);
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      }
    }
    if (issuer) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
  tcgen05_fence_before();
  __syncthreads();
  if (warp == 0) {
    tcgen05_fence_after();
    tmem_dealloc<1>(tmem_base, 512);
  }
}

static int make_map(CUtensorMap* map, const void* base, long long rows, long long cols) {
  const uint64_t dims[4] = {static_cast<uint64_t>(cols), static_cast<uint64_t>(rows), 1, 1};
  const uint64_t strides[3] = {static_cast<uint64_t>(cols) * 2, static_cast<uint64_t>(cols) * 2 * rows,
                               static_cast<uint64_t>(cols) * 2 * rows};
  const uint32_t box[4] = {64, 128, 1, 1};
  return get_tensor_map_bf16(map, base, dims, strides, box);
}

static int check_shape(const char* who, int batch, int seq, int heads, int head_dim) {
  DTB_REQUIRE(head_dim == HD, "%s: head_dim must be 64, got %d", who, head_dim);
  DTB_REQUIRE(batch > 0 && heads > 0 && seq > 0 && seq % TQ == 0, "%s: seq (%d) must be a positive multiple of 128", who, seq);
  return 0;
}

}  // namespace attn
}  // namespace dtb

using namespace dtb;
using namespace dtb::attn;

extern "C" int dtb_attention_fwd(const void* qkv, void* out, float* lse, int batch, int seq, int heads, int head_dim,
                                 void* stream) {
  if (check_shape("dtb_attention_fwd", batch, seq, heads, head_dim)) return 2;
  CUtensorMap map;
  if (make_map(&map, qkv, static_cast<long long>(batch) * seq, 3LL * heads * HD)) return 1;
  static bool configured = false;
  if (!configured) {
    DTB_CUDA_OK(cudaFuncSetAttribute(fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, FWD_SMEM));
    DTB_CUDA_OK(cudaFuncSetAttribute(fwd_lazy_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, FWD_SMEM));
    configured = true;
  }
  dim3 grid(static_cast<unsigned>(seq / TQ) * heads * batch);
  // DTB_ATTN_FWD_LAZY=0: the first kernel (O accumulated in registers, rescaled every tile)
  static const bool lazy = [] { const char* e = getenv("DTB_ATTN_FWD_LAZY"); return !(e && e[0] == '0'); }();
  if (lazy)
    fwd_lazy_kernel<<<grid, 128, FWD_SMEM, static_cast<cudaStream_t>(stream)>>>(
        map, static_cast<__nv_bfloat16*>(out), lse, seq, heads, 1.0f / sqrtf(static_cast<float>(HD)));
  else
    fwd_kernel<<<grid, 128, FWD_SMEM, static_cast<cudaStream_t>(stream)>>>(
        map, static_cast<__nv_bfloat16*>(out), lse, seq, heads, 1.0f / sqrtf(static_cast<float>(HD)));
  DTB_LAUNCH_OK();
  return 0;
}

extern "C" size_t dtb_attention_bwd_scratch_floats(int batch, int seq, int heads, int head_dim) {
  return static_cast<size_t>(batch) * heads * seq * (static_cast<size_t>(head_dim) + 1);
}

extern "C" int dtb_attention_bwd(const void* qkv, const void* out, const void* dout, const float* lse, void* dqkv,
                                 float* scratch, int batch, int seq, int heads, int head_dim, void* stream) {
  if (check_shape("dtb_attention_bwd", batch, seq, heads, head_dim)) return 2;
  DTB_REQUIRE(scratch != nullptr, "dtb_attention_bwd: scratch is required");
  auto s = static_cast<cudaStream_t>(stream);
  const long long bt = static_cast<long long>(batch) * seq;
  float* dq_acc = scratch;
  float* delta = scratch + bt * heads * HD;
  CUtensorMap map_qkv, map_do;
  if (make_map(&map_qkv, qkv, bt, 3LL * heads * HD)) return 1;
  if (make_map(&map_do, dout, bt, 1LL * heads * HD)) return 1;
  static bool configured = false;
  if (!configured) {
    DTB_CUDA_OK(cudaFuncSetAttribute(bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, BWD_SMEM));
    configured = true;
  }
  DTB_CUDA_OK(cudaMemsetAsync(dq_acc, 0, sizeof(float) * bt * heads * HD, s));
  const long long prep_threads = bt * heads * 8;
  bwd_prep_kernel<<<static_cast<int>((prep_threads + 255) / 256), 256, 0, s>>>(
      static_cast<const __nv_bfloat16*>(out), static_cast<const __nv_bfloat16*>(dout), delta, static_cast<int>(bt), seq,
      heads);
  DTB_LAUNCH_OK();
  dim3 grid(static_cast<unsigned>(seq / TK) * heads * batch);
  // The warp-specialised kernel is the default (validated against the fp32 oracle on a B200; B8 H12 T1024: 130 us vs
  // 138 us for bwd_kernel, B2 H12 T4096: 271 vs 345 us).  DTB_ATTN_BWD_WS=0 selects the first (4-warp) kernel.
  static const bool ws = [] { const char* e = getenv("DTB_ATTN_BWD_WS"); return !(e && e[0] == '0'); }();
  static const int dbg = [] { const char* e = getenv("DTB_ATTN_DBG"); return e ? atoi(e) : 0; }();   // timing experiments
  if (ws) {
    static bool ws_configured = false;
    if (!ws_configured) {
      DTB_CUDA_OK(cudaFuncSetAttribute(bwd_ws_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, BWD_WS_SMEM));
      ws_configured = true;
    }
    bwd_ws_kernel<<<grid, BWD_WS_THREADS, BWD_WS_SMEM, s>>>(map_qkv, map_do, lse, delta, dq_acc,
                                                         static_cast<__nv_bfloat16*>(dqkv), seq, heads,
                                                         1.0f / sqrtf(static_cast<float>(HD)), dbg);
  } else {
    bwd_kernel<<<grid, 128, BWD_SMEM, s>>>(map_qkv, map_do, lse, delta, dq_acc, static_cast<__nv_bfloat16*>(dqkv), seq,
                                           heads, 1.0f / sqrtf(static_cast<float>(HD)));
  }
  DTB_LAUNCH_OK();
  const long long vecs = bt * heads * (HD / 8);
  bwd_finish_kernel<<<static_cast<int>((vecs + 255) / 256), 256, 0, s>>>(dq_acc, static_cast<__nv_bfloat16*>(dqkv), batch,
                                                                         seq, heads);
  DTB_LAUNCH_OK();
  return 0;
}
