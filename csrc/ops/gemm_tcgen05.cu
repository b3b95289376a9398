// bf16 GEMM on 5th-generation tensor cores (tcgen05.mma, accumulators in TMEM),
// operands staged by TMA into 128B-swizzled shared memory through an mbarrier
// ring, persistent CTAs, warp-specialised:
//
//   warp 0      TMA producer (one elected lane)
//   warp 1      TMEM allocator + MMA issuer (one elected lane; leader CTA only in pair mode)
//   warps 2..5  epilogue: tcgen05.ld -> fp32 math (bias / GELU / GELU' / axpby) -> global
//
// Two TMEM accumulator stages let the epilogue of tile i overlap the MMAs of
// tile i+1.  CG=2 runs the MMA as cta_group::2 on a 2-CTA cluster (256 x BN
// tile, each CTA stages its 128 rows of A and half of B).
//
// Replaces the cuBLAS calls behind nn.Linear forward/backward in the reference
// (/root/reference/src/distributed_trainer.py:162,164).
#include <cuda.h>
#include <cstdlib>
#include <cuda_bf16.h>

#include <map>
#include <mutex>
#include <utility>
#include <cuda_runtime.h>

#include "../../include/dtb200.h"
#include "../common/host.h"
#include "../common/ptx.cuh"

namespace dtb {
namespace gemm {

constexpr int BM = 128;        // rows of the tile owned by one CTA
constexpr int BK = 64;         // one 128-byte swizzle atom of bf16 along K
constexpr int UMMA_K = 16;
constexpr int NUM_THREADS = 192;
constexpr int NUM_EPI_WARPS = 4;

struct Params {
  int M, N, K;
  int batch_inner, batch;        // batch = inner * outer
  int m_blocks, n_blocks;        // in units of (CG*BM) x BN tiles
  int group_m, group_m_log2;     // rasterisation group (power of two)
  float rcp_per_batch, rcp_group_size;   // reciprocals for div_small (host guarantees < 2^22 work units)
  void* d;
  void* d2;
  const void* aux;
  const void* bias;
  long long ldd, ld_aux;
  long long d_si, d_so, aux_si, aux_so;
  int d_fp32, aux_fp32, bias_fp32;
  int act;
  float alpha, beta;
  uint32_t lbo_a, sbo_a, lbo_b, sbo_b;   // UMMA descriptor byte offsets
  // ---- all-gather fusion (GATHER kernels): B = weights [N, K], rows sharded evenly over g_world ranks.
  // Gather warps pull 16-row pieces of every rank's shard (peer-mapped) into the local full matrix that
  // map_b describes and bump the piece's n-block counter; the TMA producer acquires the counter before it
  // loads a tile of that n-block.  Tiles are visited own-rows-first so transfers run ahead of the MMAs.
  const void* g_peer_shard[DTB_MAX_RANKS];
  void* g_full;
  int* g_flags;            // [n_blocks] arrival counters, then one work counter; zeroed before launch
  long long g_ldb;
  int g_world, g_rank, g_rows_per_rank;
  int g_rotate;            // n-blocks to rotate the visiting order by (rank * blocks_per_rank); -1 = off
  int direct_epilogue;     // 1 = per-thread row stores (no shared-memory staging); DTB_GEMM_DIRECT_EPILOGUE=1
  // ---- column sums of D (bias gradient of the layer D is the output gradient of): fp32 [N], added with vector
  // reductions from the staged epilogue's store loop, where a lane already holds 8 rows x 8 columns of the tile
  float* colsum;
  // ---- split-K: work unit = (tile, split); split s accumulates k-blocks [s*kb_per_split, ...) and writes its
  // raw fp32 accumulators to ws[s][M][N]; splitk_reduce_kernel adds the slices in order and runs the epilogue.
  int split_k, kb_per_split;
  float* ws;
  // ---- dynamic tile scheduling: clusters take work units from this global counter, so an SM slowed down by a
  // co-resident communication CTA ends up doing fewer tiles and a cluster whose SMs are held by another kernel joins
  // late (or not at all).  nullptr = static round-robin.  The last fetch of a launch resets the counter to zero.
  int* sched;
  // ---- TMA-store epilogue: [32 rows x 64 cols] boxes of D (and D2) with the staging buffers' 128-byte swizzle
  int tma_store;
  CUtensorMap map_d, map_d2;
};

constexpr int SCHED_DEPTH = 4;       // ring of published tile ids per CTA

constexpr int GATHER_ROWS = 16;         // rows per gather piece
constexpr int GATHER_CHUNK_VECS = 128;  // 16-byte vectors (2 KB) of a row per gather piece

template <int CG, int BN>
struct Config {
  static constexpr int A_BYTES = BM * BK * 2;              // 16 KB per CTA
  static constexpr int B_ROWS = BN / CG;                   // B rows staged by one CTA
  static constexpr int B_BYTES = B_ROWS * BK * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int STAGES = (STAGE_BYTES >= 48 * 1024) ? 4 : 6;
  static constexpr int EPI_STAGE_BYTES = NUM_EPI_WARPS * 8192;   // per epilogue warp: 2 x [32 rows x 128 B]
  static constexpr int TMEM_COLS = 2 * BN;                 // two accumulator stages
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + EPI_STAGE_BYTES + 1024 /*align*/ + 256 /*barriers*/;
  static_assert(TMEM_COLS <= 512, "TMEM has 512 columns");
  static_assert(TMEM_COLS == 128 || TMEM_COLS == 256 || TMEM_COLS == 512, "power of two");
};

__device__ __forceinline__ float gelu_tanh(float x) {
  const float k0 = 0.7978845608028654f, k1 = 0.044715f;
  float u = k0 * (x + k1 * x * x * x);
  float t;
  asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(u));
  return 0.5f * x * (1.0f + t);
}

__device__ __forceinline__ float gelu_tanh_grad(float x) {
  const float k0 = 0.7978845608028654f, k1 = 0.044715f;
  float x2 = x * x;
  float u = k0 * (x + k1 * x * x2);
  float t;
  asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(u));
  float du = k0 * (1.0f + 3.0f * k1 * x2);
  return 0.5f * (1.0f + t) + 0.5f * x * (1.0f - t * t) * du;
}

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}

// v[0..32) += bias[col0 .. col0+32): four 16-byte loads when the whole chunk is in range and aligned
__device__ __forceinline__ void add_bias32(const Params& p, float (&v)[32], int col0) {
  if (p.bias_fp32) {
    const float* b = reinterpret_cast<const float*>(p.bias) + col0;
    if (col0 + 32 <= p.N && (reinterpret_cast<uintptr_t>(b) & 15) == 0) {
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        const float4 t = __ldg(reinterpret_cast<const float4*>(b) + q);
        v[4 * q] += t.x; v[4 * q + 1] += t.y; v[4 * q + 2] += t.z; v[4 * q + 3] += t.w;
      }
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if (col0 + j < p.N) v[j] += __ldg(b + j);
    }
  } else {
    const __nv_bfloat16* b = reinterpret_cast<const __nv_bfloat16*>(p.bias) + col0;
    if (col0 + 32 <= p.N && (reinterpret_cast<uintptr_t>(b) & 15) == 0) {
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const uint4 t = __ldg(reinterpret_cast<const uint4*>(b) + q);
        const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&t);
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float2 f = __bfloat1622float2(h[e]);
          v[8 * q + 2 * e] += f.x;
          v[8 * q + 2 * e + 1] += f.y;
        }
      }
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if (col0 + j < p.N) v[j] += __bfloat162float(b[j]);
    }
  }
}

// n / d for 0 <= n < 2^22 with rcp = 1.0f / d: the float estimate is off by at most one.  A runtime integer division is
// ~35 dependent instructions; tile_coords used five of them per tile in every warp role, which the ncu source view
// showed as 43 % of the epilogue warps' time in an epilogue-bound (short-K) GEMM.
__device__ __forceinline__ int div_small(int n, int d, float rcp) {
  int q = __float2int_rz(__int2float_rz(n) * rcp);
  const int r = n - q * d;
  q += (r >= d) ? 1 : 0;
  q -= (r < 0) ? 1 : 0;
  return q;
}

__device__ __forceinline__ void tile_coords(const Params& p, int tile, int& bi, int& m_blk, int& n_blk) {
  if (p.split_k > 1) tile /= p.split_k;      // consecutive work units are the K-splits of one output tile
  const int per_batch = p.m_blocks * p.n_blocks;
  bi = 0;
  int t = tile;
  if (p.batch > 1) {
    bi = div_small(tile, per_batch, p.rcp_per_batch);
    t = tile - bi * per_batch;
  }
  if (p.g_rotate >= 0) {   // all m-blocks of one n-block, n-blocks starting at this rank's own rows
    const int q = t / p.m_blocks;
    m_blk = t - q * p.m_blocks;
    n_blk = (q + p.g_rotate) % p.n_blocks;
    return;
  }
  const int group_size = p.group_m * p.n_blocks;
  const int group_id = div_small(t, group_size, p.rcp_group_size);
  const int first_m = group_id * p.group_m;
  const int in_group = t - group_id * group_size;
  if (p.m_blocks - first_m >= p.group_m) {     // full group: group_m is a power of two
    m_blk = first_m + (in_group & (p.group_m - 1));
    n_blk = in_group >> p.group_m_log2;
  } else {
    const int gm = p.m_blocks - first_m;
    n_blk = in_group / gm;
    m_blk = first_m + in_group - n_blk * gm;
  }
}

// Epilogue for one 32-column chunk of one row.
__device__ __forceinline__ void epilogue_chunk(const Params& p, uint32_t (&raw)[32], int row, int col0,
                                               long long d_off, long long aux_off) {
  if (row >= p.M || col0 >= p.N) return;
  float v[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(raw[j]) * p.alpha;
  const bool full = (col0 + 32 <= p.N);

  if (p.bias != nullptr) add_bias32(p, v, col0);

  if (p.act == 1) {
    if (p.d2 != nullptr) {
      __nv_bfloat16* z = reinterpret_cast<__nv_bfloat16*>(p.d2) + d_off + static_cast<long long>(row) * p.ldd + col0;
      if (full) {
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          uint4 o;
          o.x = pack_bf16x2(v[8 * q + 0], v[8 * q + 1]);
          o.y = pack_bf16x2(v[8 * q + 2], v[8 * q + 3]);
          o.z = pack_bf16x2(v[8 * q + 4], v[8 * q + 5]);
          o.w = pack_bf16x2(v[8 * q + 6], v[8 * q + 7]);
          *reinterpret_cast<uint4*>(z + 8 * q) = o;
        }
      } else {
        for (int j = 0; j < 32 && col0 + j < p.N; ++j) z[j] = __float2bfloat16(v[j]);
      }
    }
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = gelu_tanh(v[j]);
  } else if (p.aux != nullptr) {
    float a[32];
    if (p.aux_fp32) {
      const float* ap = reinterpret_cast<const float*>(p.aux) + aux_off + static_cast<long long>(row) * p.ld_aux + col0;
      if (full) {
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          float4 t = *reinterpret_cast<const float4*>(ap + 4 * q);
          a[4 * q + 0] = t.x; a[4 * q + 1] = t.y; a[4 * q + 2] = t.z; a[4 * q + 3] = t.w;
        }
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j) a[j] = (col0 + j < p.N) ? ap[j] : 0.f;
      }
    } else {
      const __nv_bfloat16* ap =
          reinterpret_cast<const __nv_bfloat16*>(p.aux) + aux_off + static_cast<long long>(row) * p.ld_aux + col0;
      if (full) {
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          uint4 t = *reinterpret_cast<const uint4*>(ap + 8 * q);
          const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&t);
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            float2 f = __bfloat1622float2(h[e]);
            a[8 * q + 2 * e] = f.x;
            a[8 * q + 2 * e + 1] = f.y;
          }
        }
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j) a[j] = (col0 + j < p.N) ? __bfloat162float(ap[j]) : 0.f;
      }
    }
    if (p.act == 2) {
#pragma unroll
      for (int j = 0; j < 32; ++j) v[j] *= gelu_tanh_grad(a[j]);
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j) v[j] += p.beta * a[j];
    }
  }

  if (p.d_fp32) {
    float* dp = reinterpret_cast<float*>(p.d) + d_off + static_cast<long long>(row) * p.ldd + col0;
    if (full) {
#pragma unroll
      for (int q = 0; q < 8; ++q)
        *reinterpret_cast<float4*>(dp + 4 * q) = make_float4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
    } else {
      for (int j = 0; j < 32 && col0 + j < p.N; ++j) dp[j] = v[j];
    }
  } else {
    __nv_bfloat16* dp = reinterpret_cast<__nv_bfloat16*>(p.d) + d_off + static_cast<long long>(row) * p.ldd + col0;
    if (full) {
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        uint4 o;
        o.x = pack_bf16x2(v[8 * q + 0], v[8 * q + 1]);
        o.y = pack_bf16x2(v[8 * q + 2], v[8 * q + 3]);
        o.z = pack_bf16x2(v[8 * q + 4], v[8 * q + 5]);
        o.w = pack_bf16x2(v[8 * q + 6], v[8 * q + 7]);
        *reinterpret_cast<uint4*>(dp + 8 * q) = o;
      }
    } else {
      for (int j = 0; j < 32 && col0 + j < p.N; ++j) dp[j] = __float2bfloat16(v[j]);
    }
  }
}

// ---- staged epilogue ------------------------------------------------------------------------------------------
// One warp owns 32 accumulator rows (a thread owns a TMEM lane = an output row), so un-staged stores touch 32
// different cache lines per instruction.  Each 64-column group therefore goes TMEM -> registers -> this warp's
// [32 x 128 B] shared-memory buffer (16-byte chunk j of row r lives at chunk j ^ (r & 7)) -> global memory with every
// access a full 128-byte row segment.  Requirements (checked by the caller): bf16 D / aux, N % 8 == 0.
//
// The first version spent 221 SASS instructions per tcgen05.ld (ncu: the epilogue warps were issue-bound and a
// GEMM with K = 768 took 6.5 us per tile for a 4.4 us main loop): 64-bit address arithmetic and bounds checks per
// store, parameters re-read from the constant bank.  Everything that does not depend on the column group is now
// computed once per tile (EpiTile), interior tiles skip all bounds checks, and the bias is fetched once per tile
// into registers BEFORE the accumulator wait and handed out with warp shuffles.
struct EpiTile {
  uint8_t* out;          // &D[row_base + lane/8][n0 + (lane%8)*8]: this lane's first 16-byte segment
  uint8_t* out2;         // same position in D2 (pre-activation copy) or nullptr
  const uint8_t* aux;    // same position in aux or nullptr
  long long step_d;      // bytes between rows r and r + 4 of D / D2
  long long step_aux;
  int row_base, n0;
  int b_in, b_out;       // batch coordinates (TMA-store path)
  bool interior;         // all 32 rows and all BN columns are in range
  bool bias_regs;        // bias for the whole tile is held across the warp in `bias` (lane l: columns 8l .. 8l+7)
  uint4 bias;
};

// Coalesced load of this warp's [32 rows x 64 cols] bf16 aux tile of column group `c2` (8 x 16 B per lane).
__device__ __forceinline__ void load_aux_group(const Params& p, const EpiTile& t, uint4 (&areg)[8], int c2, int lane) {
  const uint8_t* ap = t.aux + c2 * 128;
  if (t.interior) {
#pragma unroll
    for (int it = 0; it < 8; ++it) areg[it] = __ldg(reinterpret_cast<const uint4*>(ap + it * t.step_aux));
  } else {
    const int sub = lane >> 3, gc = t.n0 + c2 * 64 + (lane & 7) * 8;
#pragma unroll
    for (int it = 0; it < 8; ++it) {
      areg[it] = make_uint4(0, 0, 0, 0);
      if (t.row_base + it * 4 + sub < p.M && gc < p.N) areg[it] = __ldg(reinterpret_cast<const uint4*>(ap + it * t.step_aux));
    }
  }
}

// TMEM -> registers -> epilogue math -> shared memory for column group c2.  `areg` holds this group's aux tile
// (loaded one group ahead so its latency hides behind the previous group's math); on return it holds the next one.
__device__ __forceinline__ void epilogue_group_compute(const Params& p, const EpiTile& t, uint32_t taddr, int c2,
                                                       int num_groups, uint8_t* st_out, uint8_t* st_in, int lane,
                                                       uint4 (&areg)[8]) {
  const int sub = lane >> 3, seg = lane & 7;
  const bool has_aux = (t.aux != nullptr) && (p.act != 1);
  const uint32_t xs = static_cast<uint32_t>(lane & 7) << 4;       // this row's chunk swizzle
  uint8_t* my_out = st_out + lane * 128;
  uint8_t* my_in = st_in + lane * 128;
  if (has_aux) {
#pragma unroll
    for (int it = 0; it < 8; ++it) {
      const int rl = it * 4 + sub;
      *reinterpret_cast<uint4*>(st_in + rl * 128 + ((seg ^ (rl & 7)) << 4)) = areg[it];
    }
    __syncwarp();
    if (c2 + 1 < num_groups) load_aux_group(p, t, areg, c2 + 1, lane);
  }
  const bool want_z = (p.act == 1) && (t.out2 != nullptr);
#pragma unroll 1
  for (int half = 0; half < 2; ++half) {
    const int col0 = t.n0 + c2 * 64 + half * 32;
    if (!t.interior && col0 >= p.N) break;
    uint32_t raw[32];
    tmem_ld_32x32b_x32(taddr + half * 32, raw);
    tmem_ld_wait();
    float v[32];
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(raw[j]);
    if (p.alpha != 1.0f) {
#pragma unroll
      for (int j = 0; j < 32; ++j) v[j] *= p.alpha;
    }
    if (t.bias_regs) {
      const int src0 = (c2 * 2 + half) * 4;     // lanes holding these 32 columns
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const uint32_t w0 = __shfl_sync(0xffffffffu, t.bias.x, src0 + q), w1 = __shfl_sync(0xffffffffu, t.bias.y, src0 + q);
        const uint32_t w2 = __shfl_sync(0xffffffffu, t.bias.z, src0 + q), w3 = __shfl_sync(0xffffffffu, t.bias.w, src0 + q);
        const float2 f0 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w0));
        const float2 f1 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w1));
        const float2 f2 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w2));
        const float2 f3 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w3));
        v[8 * q + 0] += f0.x; v[8 * q + 1] += f0.y; v[8 * q + 2] += f1.x; v[8 * q + 3] += f1.y;
        v[8 * q + 4] += f2.x; v[8 * q + 5] += f2.y; v[8 * q + 6] += f3.x; v[8 * q + 7] += f3.y;
      }
    } else if (p.bias != nullptr) {
      add_bias32(p, v, col0);
    }
    const uint32_t hx = xs ^ static_cast<uint32_t>(half * 64);
    if (p.act == 1) {
      if (want_z) {
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          uint4 o;
          o.x = pack_bf16x2(v[8 * q + 0], v[8 * q + 1]);
          o.y = pack_bf16x2(v[8 * q + 2], v[8 * q + 3]);
          o.z = pack_bf16x2(v[8 * q + 4], v[8 * q + 5]);
          o.w = pack_bf16x2(v[8 * q + 6], v[8 * q + 7]);
          *reinterpret_cast<uint4*>(my_in + (hx ^ (q * 16))) = o;
        }
      }
#pragma unroll
      for (int j = 0; j < 32; ++j) v[j] = gelu_tanh(v[j]);
    } else if (has_aux) {
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const uint4 a4 = *reinterpret_cast<const uint4*>(my_in + (hx ^ (q * 16)));
        const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&a4);
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float2 f = __bfloat1622float2(h[e]);
          if (p.act == 2) {
            v[8 * q + 2 * e] *= gelu_tanh_grad(f.x);
            v[8 * q + 2 * e + 1] *= gelu_tanh_grad(f.y);
          } else {
            v[8 * q + 2 * e] += p.beta * f.x;
            v[8 * q + 2 * e + 1] += p.beta * f.y;
          }
        }
      }
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      uint4 o;
      o.x = pack_bf16x2(v[8 * q + 0], v[8 * q + 1]);
      o.y = pack_bf16x2(v[8 * q + 2], v[8 * q + 3]);
      o.z = pack_bf16x2(v[8 * q + 4], v[8 * q + 5]);
      o.w = pack_bf16x2(v[8 * q + 6], v[8 * q + 7]);
      *reinterpret_cast<uint4*>(my_out + (hx ^ (q * 16))) = o;
    }
  }
  __syncwarp();
}

__device__ __forceinline__ void add_bf16x8(float (&acc)[8], const uint4& v) {
  const uint32_t u[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    acc[2 * j] += __uint_as_float(u[j] << 16);
    acc[2 * j + 1] += __uint_as_float(u[j] & 0xffff0000u);
  }
}

// shared memory -> global memory for column group c2: 8 steps of 4 rows x 128 bytes per warp
__device__ __forceinline__ void epilogue_group_store(const Params& p, const EpiTile& t, int c2, const uint8_t* st_out,
                                                     const uint8_t* st_in, int lane) {
  const int sub = lane >> 3, seg = lane & 7;
  const bool want_z = (p.act == 1) && (t.out2 != nullptr);
  // row rl = 4*it + sub sits at rl*128 + ((seg ^ (rl & 7)) << 4); (rl & 7) = sub for even it, sub + 4 for odd it
  const uint32_t even = static_cast<uint32_t>(sub * 128 + ((seg ^ sub) << 4)), odd = even ^ 64u;
  uint8_t* dp = t.out + c2 * 128;
  uint8_t* zp = t.out2 + c2 * 128;
  const bool cs_on = p.colsum != nullptr;
  float cs[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  const int gc = t.n0 + c2 * 64 + seg * 8;
  if (t.interior) {
#pragma unroll
    for (int it = 0; it < 8; ++it) {
      const uint32_t so = ((it & 1) ? odd : even) + it * 512;
      const uint4 o = *reinterpret_cast<const uint4*>(st_out + so);
      *reinterpret_cast<uint4*>(dp + it * t.step_d) = o;
      if (want_z) *reinterpret_cast<uint4*>(zp + it * t.step_d) = *reinterpret_cast<const uint4*>(st_in + so);
      if (cs_on) add_bf16x8(cs, o);
    }
  } else {
#pragma unroll
    for (int it = 0; it < 8; ++it) {
      const uint32_t so = ((it & 1) ? odd : even) + it * 512;
      if (t.row_base + it * 4 + sub < p.M && gc < p.N) {
        const uint4 o = *reinterpret_cast<const uint4*>(st_out + so);
        *reinterpret_cast<uint4*>(dp + it * t.step_d) = o;
        if (want_z) *reinterpret_cast<uint4*>(zp + it * t.step_d) = *reinterpret_cast<const uint4*>(st_in + so);
        if (cs_on) add_bf16x8(cs, o);
      }
    }
  }
  if (cs_on) {
    // lanes with the same `seg` hold the other 24 rows of these 8 columns: fold them, then one lane per segment
    // sends two 16-byte reductions.  32 x 256 of them per 128 x 256 tile: invisible next to a K >= 4096 main loop.
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      cs[j] += __shfl_xor_sync(0xffffffffu, cs[j], 8);
      cs[j] += __shfl_xor_sync(0xffffffffu, cs[j], 16);
    }
    if (sub == 0 && gc < p.N) {
      float* dst = p.colsum + gc;
      asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst), "f"(cs[0]), "f"(cs[1]), "f"(cs[2]),
                   "f"(cs[3])
                   : "memory");
      asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst + 4), "f"(cs[4]), "f"(cs[5]), "f"(cs[6]),
                   "f"(cs[7])
                   : "memory");
    }
  }
  __syncwarp();
}

// Same hand-off through the TMA unit: one elected lane issues a bulk tensor store of the [32 x 64] box (and of the
// pre-activation box).  The warps' LSU work per group drops from 8 x (LDS + STG) per lane to nothing, edges are clipped
// by the tensor map, and the store overlaps the next group's TMEM reads.
__device__ __forceinline__ void epilogue_group_store_tma(const Params& p, const EpiTile& t, int c2, const uint8_t* st_out,
                                                         const uint8_t* st_in, int lane) {
  fence_proxy_async_smem();      // every lane: my st.shared results are visible to the async proxy
  __syncwarp();
  if (lane == 0) {
    tma_store_4d(&p.map_d, st_out, t.n0 + c2 * 64, t.row_base, t.b_in, t.b_out);
    if (p.act == 1 && t.out2 != nullptr) tma_store_4d(&p.map_d2, st_in, t.n0 + c2 * 64, t.row_base, t.b_in, t.b_out);
    tma_store_commit();
  }
}

constexpr int NUM_GATHER_WARPS = 2;

template <int CG, int BN, int A_MN, int B_MN, bool GATHER = false>
__global__ void __launch_bounds__(NUM_THREADS + (GATHER ? 32 * NUM_GATHER_WARPS : 0), 1)
gemm_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
            const __grid_constant__ Params p) {
  using Cfg = Config<CG, BN>;
  constexpr int STAGES = Cfg::STAGES;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smem_a = smem;                                   // STAGES x A_BYTES
  uint8_t* smem_b = smem + STAGES * Cfg::A_BYTES;           // STAGES x B_BYTES
  uint8_t* epi_stage = smem + STAGES * Cfg::STAGE_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(epi_stage + Cfg::EPI_STAGE_BYTES);
  uint64_t* full_bar = bars;                  // [STAGES]
  uint64_t* empty_bar = bars + STAGES;        // [STAGES]
  uint64_t* tmem_full = bars + 2 * STAGES;    // [2]
  uint64_t* tmem_empty = bars + 2 * STAGES + 2;  // [2]
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 4);
  uint64_t* sched_full = bars + 2 * STAGES + 5;    // [SCHED_DEPTH] per CTA: tile id published
  uint64_t* sched_empty = bars + 2 * STAGES + 9;   // [SCHED_DEPTH] leader only: every consumer of the cluster has read it
  int* sched_tile = reinterpret_cast<int*>(bars + 2 * STAGES + 13);   // [SCHED_DEPTH]
  static_assert((2 * STAGES + 13) * 8 + SCHED_DEPTH * 4 <= 256, "barrier block is 256 bytes");

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t cta_rank = (CG == 2) ? cluster_ctarank() : 0u;
  const bool leader = (cta_rank == 0);
  const int cluster_id = (CG == 2) ? (blockIdx.x >> 1) : blockIdx.x;
  const int num_clusters = (CG == 2) ? (gridDim.x >> 1) : gridDim.x;
  const int num_tiles = p.m_blocks * p.n_blocks * p.batch * p.split_k;   // work units
  const int num_kb = (p.K + BK - 1) / BK;

  if (warp == 0 && lane == 0) {
    prefetch_tensormap(&map_a);
    prefetch_tensormap(&map_b);
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tmem_full[a], 1);
      mbar_init(&tmem_empty[a], NUM_EPI_WARPS * CG);
    }
    for (int q = 0; q < SCHED_DEPTH; ++q) {
      mbar_init(&sched_full[q], 1);
      // consumers: leader CTA = MMA warp + epilogue warps, second CTA = TMA warp + epilogue warps
      mbar_init(&sched_empty[q], (1 + NUM_EPI_WARPS) * CG);
    }
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc<CG>(tmem_ptr, Cfg::TMEM_COLS);
  }
  tcgen05_fence_before();
  if constexpr (CG == 2) {
    cluster_sync();
  } else {
    __syncthreads();
  }
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_ptr;

  // Consumer side of the tile ring (dynamic scheduling): the j-th published tile id of this cluster.  Called by whole
  // warps.  The id arrives by st.async (data + 4 bytes of transaction count on this CTA's sched_full barrier), so a
  // plain CTA-scope wait is enough; `arm` = this warp is the one that posts the barrier's single arrival + expect_tx for
  // the phase (the TMA warp of each CTA).  One lane then tells the leader CTA that this warp is done with the slot.
  const bool dyn = !GATHER && p.sched != nullptr;
  auto next_tile = [&](int tile, int& j, bool arm) -> int {
    if (!dyn) return tile + num_clusters;
    const int slot = j & (SCHED_DEPTH - 1);
    if (arm && lane == 0) mbar_arrive_expect_tx(&sched_full[slot], 4);
    mbar_wait(&sched_full[slot], (j / SCHED_DEPTH) & 1);
    const int t = *reinterpret_cast<volatile int*>(&sched_tile[slot]);
    __syncwarp();
    // the predicate depends on the loaded value, so the arrive cannot be issued before the read has returned
    if (lane == 0 && t >= 0) {
      if (CG == 1 || leader) {
        mbar_arrive(&sched_empty[slot]);
      } else {
        mbar_arrive_remote(&sched_empty[slot], 0);
      }
    }
    ++j;
    return t;
  };

  if (warp == 0) {
    // ============================ TMA producer ============================
    // The whole warp walks the loop so every address / coordinate stays in uniform registers;
    // one elected lane issues the copies (a lane-0-only region makes ptxas wrap each UTMALDG
    // in a vote / R2UR waterfall).
    int stage = 0;
    uint32_t phase = 0;
    // Dynamic scheduling: this warp of the leader CTA is the scheduler.  At the start of a tile it takes the NEXT work
    // unit from the global counter (the atomic's latency hides behind the first loads) and publishes it to both CTAs
    // one k-block later, so nobody ever waits for a tile id and a cluster is committed only one tile ahead.
    const bool scheduler = dyn && (CG == 1 || leader);
    int fetched = 0, published = 0, consumed = 0;
    auto publish = [&]() -> int {
      const int nxt = __shfl_sync(0xffffffffu, fetched, 0);
      const int slot = published & (SCHED_DEPTH - 1);
      mbar_wait(&sched_empty[slot], ((published / SCHED_DEPTH) & 1) ^ 1);
      if (lane == 0) {
        // every cluster fetches until it is handed a unit past the end: num_tiles + num_clusters fetches per launch
        if (nxt == num_tiles + num_clusters - 1) atomicExch(p.sched, 0);
        mbar_arrive_expect_tx(&sched_full[slot], 4);          // this CTA's arrival for the phase
#pragma unroll
        for (uint32_t r = 0; r < static_cast<uint32_t>(CG); ++r)
          st_async_cluster_u32(&sched_tile[slot], &sched_full[slot], r, static_cast<uint32_t>(nxt));
      }
      __syncwarp();
      ++published;
      return nxt;
    };
    int tile = cluster_id;
    if (scheduler) {           // the first tile comes from the counter too: a cluster that starts late (its SMs were
      if (lane == 0) fetched = atomicAdd(p.sched, 1);   // held by another kernel) finds only what is still undone
      tile = publish();
    } else if (dyn) {
      tile = next_tile(0, consumed, true);
    }
    while (tile < num_tiles) {
      bool pending = false;
      int following = 0;
      if (scheduler) {
        if (lane == 0) fetched = atomicAdd(p.sched, 1);
        pending = true;
      }
      int bi, m_blk, n_blk;
      tile_coords(p, tile, bi, m_blk, n_blk);
      int b_in = 0, b_out = 0;
      if (p.batch > 1) {
        b_out = bi / p.batch_inner;
        b_in = bi - b_out * p.batch_inner;
      }
      const int m0 = (m_blk * CG + static_cast<int>(cta_rank)) * BM;
      const int n0 = n_blk * BN + static_cast<int>(cta_rank) * Cfg::B_ROWS;
      if constexpr (GATHER) {
        // every 16-row piece of this n-block has landed in the local matrix (written with generic stores by
        // gather warps of any CTA): acquire the counter, then order the async-proxy (TMA) reads after it
        const int want = ((min(BN, p.N - n_blk * BN) + GATHER_ROWS - 1) / GATHER_ROWS) *
                         ((p.K / 8 + GATHER_CHUNK_VECS - 1) / GATHER_CHUNK_VECS);
        while (ld_acquire_gpu(reinterpret_cast<const uint32_t*>(p.g_flags + n_blk)) < static_cast<uint32_t>(want)) {
        }
        asm volatile("fence.proxy.async.global;" ::: "memory");
      }
      const int kb_lo = p.split_k > 1 ? (tile % p.split_k) * p.kb_per_split : 0;
      const int kb_hi = min(num_kb, kb_lo + p.kb_per_split);
      for (int kb = kb_lo; kb < kb_hi; ++kb) {
        mbar_wait(&empty_bar[stage], phase ^ 1);
        uint8_t* sa = smem_a + stage * Cfg::A_BYTES;
        uint8_t* sb = smem_b + stage * Cfg::B_BYTES;
        const int k0 = kb * BK;
        if (elect_one()) {
          if (CG == 1 || leader) mbar_arrive_expect_tx(&full_bar[stage], Cfg::STAGE_BYTES * CG);
          if constexpr (CG == 1) {
            if constexpr (A_MN) {
#pragma unroll
              for (int j = 0; j < BM / 64; ++j)
                tma_load_4d(sa + j * 8192, &map_a, &full_bar[stage], m0 + 64 * j, k0, b_in, b_out);
            } else {
              tma_load_4d(sa, &map_a, &full_bar[stage], k0, m0, b_in, b_out);
            }
            if constexpr (B_MN) {
#pragma unroll
              for (int j = 0; j < Cfg::B_ROWS / 64; ++j)
                tma_load_4d(sb + j * 8192, &map_b, &full_bar[stage], n0 + 64 * j, k0, b_in, b_out);
            } else {
              tma_load_4d(sb, &map_b, &full_bar[stage], k0, n0, b_in, b_out);
            }
          } else {
            if constexpr (A_MN) {
#pragma unroll
              for (int j = 0; j < BM / 64; ++j)
                tma_load_4d_2sm(sa + j * 8192, &map_a, &full_bar[stage], m0 + 64 * j, k0, b_in, b_out);
            } else {
              tma_load_4d_2sm(sa, &map_a, &full_bar[stage], k0, m0, b_in, b_out);
            }
            if constexpr (B_MN) {
#pragma unroll
              for (int j = 0; j < Cfg::B_ROWS / 64; ++j)
                tma_load_4d_2sm(sb + j * 8192, &map_b, &full_bar[stage], n0 + 64 * j, k0, b_in, b_out);
            } else {
              tma_load_4d_2sm(sb, &map_b, &full_bar[stage], k0, n0, b_in, b_out);
            }
          }
        }
        __syncwarp();
        if (++stage == STAGES) {
          stage = 0;
          phase ^= 1;
        }
        if (pending && kb >= kb_lo + 1) {
          following = publish();
          pending = false;
        }
      }
      if (pending) following = publish();
      tile = scheduler ? following : next_tile(tile, consumed, true);
    }
  } else if (warp == 1) {
    // ============================= MMA issuer =============================
    if (leader) {
      constexpr uint32_t idesc = umma_idesc_bf16(BM * CG, BN, A_MN, B_MN);
      // Descriptor start-address advance per UMMA_K (16 elements of K), in 16-byte units:
      //   K-major: 32 B inside the 128-byte swizzled row; MN-major: two 8-row groups of 128 B rows.
      constexpr uint32_t A_KSTEP16 = (A_MN ? 2048u : 32u) >> 4;
      constexpr uint32_t B_KSTEP16 = (B_MN ? 2048u : 32u) >> 4;
      // Descriptor halves that do not change inside the loop: hi = SBO | version 1 | 128B swizzle,
      // lo = start address (>>4) | LBO.  Stage bases differ by a constant, K steps by an immediate.
      const uint32_t a_hi = ((p.sbo_a >> 4) & 0x3FFFu) | (1u << 14) | (2u << 29);
      const uint32_t b_hi = ((p.sbo_b >> 4) & 0x3FFFu) | (1u << 14) | (2u << 29);
      const uint32_t a_lo0 = ((smem_u32(smem_a) >> 4) & 0x3FFFu) | (((p.lbo_a >> 4) & 0x3FFFu) << 16);
      const uint32_t b_lo0 = ((smem_u32(smem_b) >> 4) & 0x3FFFu) | (((p.lbo_b >> 4) & 0x3FFFu) << 16);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      int consumed = 0;
      for (int tile = dyn ? next_tile(0, consumed, false) : cluster_id; tile < num_tiles;
           tile = next_tile(tile, consumed, false)) {
        mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
        tcgen05_fence_after();
        const uint32_t tmem_d = tmem_base + static_cast<uint32_t>(acc * BN);
        const int kb_lo = p.split_k > 1 ? (tile % p.split_k) * p.kb_per_split : 0;
        const int kb_hi = min(num_kb, kb_lo + p.kb_per_split);
        for (int kb = kb_lo; kb < kb_hi; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tcgen05_fence_after();
          const uint32_t a_lo = a_lo0 + static_cast<uint32_t>(stage) * (Cfg::A_BYTES >> 4);
          const uint32_t b_lo = b_lo0 + static_cast<uint32_t>(stage) * (Cfg::B_BYTES >> 4);
          if (elect_one()) {
#pragma unroll
            for (int k = 0; k < BK / UMMA_K; ++k) {
              const uint64_t da = (static_cast<uint64_t>(a_hi) << 32) | (a_lo + k * A_KSTEP16);
              const uint64_t db = (static_cast<uint64_t>(b_hi) << 32) | (b_lo + k * B_KSTEP16);
              umma_f16<CG>(tmem_d, da, db, idesc, (kb > kb_lo || k > 0) ? 1u : 0u);
            }
            if constexpr (CG == 1) {
              umma_commit(&empty_bar[stage]);
            } else {
              umma_commit_2sm(&empty_bar[stage], 0x3);
            }
            if (kb == kb_hi - 1) {
              if constexpr (CG == 1) {
                umma_commit(&tmem_full[acc]);
              } else {
                umma_commit_2sm(&tmem_full[acc], 0x3);
              }
            }
          }
          __syncwarp();
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1;
      }
    }
  } else if (GATHER && warp >= NUM_THREADS / 32) {
    // ============================ gather warps ============================
    // Pieces are claimed from a global counter in the order the tiles will need them.
    // A piece = 16 rows x up to 2 KB of one n-block (row groups x column chunks), so that even a matrix with few,
    // long rows splits into enough pieces for every gather warp.  16 x 16 B per lane are in flight (8 KB per warp):
    // a row-by-row copy of a K=1600 row is 6 dependent load->store round trips over NVLink, and then the gather, not
    // the MMA, sets the kernel's time (measured 325 us for a 119 us GEMM before this loop was flattened).
    const int row_vecs = static_cast<int>(p.K) / 8;                       // 16-byte vectors per row (K % 8 == 0)
    const int col_chunks = (row_vecs + GATHER_CHUNK_VECS - 1) / GATHER_CHUNK_VECS;
    const int pieces_per_block = (BN / GATHER_ROWS) * col_chunks;
    const int total_pieces = p.n_blocks * pieces_per_block;
    int* work = p.g_flags + p.n_blocks;
    const long long pitch = static_cast<long long>(p.g_ldb) * 2;
    while (true) {
      int piece = 0;
      if (lane == 0) piece = atomicAdd(work, 1);
      piece = __shfl_sync(0xffffffffu, piece, 0);
      if (piece >= total_pieces) break;
      const int n_blk = (piece / pieces_per_block + p.g_rotate) % p.n_blocks;
      const int rem = piece % pieces_per_block;
      const int r0 = n_blk * BN + (rem / col_chunks) * GATHER_ROWS;
      const int c0 = (rem % col_chunks) * GATHER_CHUNK_VECS;
      if (r0 < p.N) {
        const int owner = r0 / p.g_rows_per_rank;
        const int nrows = min(GATHER_ROWS, p.N - r0);
        const int width = min(GATHER_CHUNK_VECS, row_vecs - c0);
        const uint8_t* src = static_cast<const uint8_t*>(p.g_peer_shard[owner]) +
                             static_cast<long long>(r0 - owner * p.g_rows_per_rank) * pitch + c0 * 16;
        uint8_t* dst = static_cast<uint8_t*>(p.g_full) + static_cast<long long>(r0) * pitch + c0 * 16;
#pragma unroll 1
        for (int base = lane; base < GATHER_ROWS * GATHER_CHUNK_VECS; base += 16 * 32) {
          uint4 v[16];
#pragma unroll
          for (int u = 0; u < 16; ++u) {
            const int idx = base + u * 32, r = idx / GATHER_CHUNK_VECS, c = idx % GATHER_CHUNK_VECS;
            if (r < nrows && c < width) v[u] = ld_nc_v4(src + r * pitch + c * 16);
          }
#pragma unroll
          for (int u = 0; u < 16; ++u) {
            const int idx = base + u * 32, r = idx / GATHER_CHUNK_VECS, c = idx % GATHER_CHUNK_VECS;
            if (r < nrows && c < width) st_v4(dst + r * pitch + c * 16, v[u]);
          }
        }
      }
      __syncwarp();
      if (lane == 0 && r0 < p.N) {   // pieces past the last row do not count towards the block's quota
        __threadfence();
        atomicAdd(p.g_flags + n_blk, 1);
      }
    }
  } else if (warp < NUM_THREADS / 32) {
    // ============================== epilogue ==============================
    const int quad = warp & 3;   // TMEM lane quadrant this warp may access
    int acc = 0;
    uint32_t acc_phase = 0;
    int consumed = 0;
    int store_groups = 0;        // bulk-store groups issued by this warp (TMA-store epilogue)
    for (int tile = dyn ? next_tile(0, consumed, false) : cluster_id; tile < num_tiles;
         tile = next_tile(tile, consumed, false)) {
      int bi, m_blk, n_blk;
      tile_coords(p, tile, bi, m_blk, n_blk);
      long long d_off = 0, aux_off = 0;
      int b_in = 0, b_out = 0;
      if (p.batch > 1) {
        b_out = bi / p.batch_inner;
        b_in = bi - b_out * p.batch_inner;
        d_off = b_in * p.d_si + b_out * p.d_so;
        aux_off = b_in * p.aux_si + b_out * p.aux_so;
      }
      const int row = (m_blk * CG + static_cast<int>(cta_rank)) * BM + quad * 32 + lane;
      const int n0 = n_blk * BN;
      // ---- per-tile setup, done BEFORE waiting for the accumulator so that the global loads issued here (bias, first
      // aux group) are in flight while the MMAs of this tile still run
      const bool staged = p.split_k == 1 && !p.direct_epilogue && !p.d_fp32 && !(p.aux != nullptr && p.aux_fp32) && (p.N & 7) == 0;
      EpiTile et;
      uint4 areg[8];
      uint8_t* st_out = epi_stage + (warp - 2) * 8192;
      const int num_groups = min(BN / 64, (p.N - n0 + 63) / 64);
      if (staged) {
        const int sub = lane >> 3, seg = lane & 7;
        et.row_base = row - lane;
        et.n0 = n0;
        et.b_in = b_in;
        et.b_out = b_out;
        et.interior = (et.row_base + 32 <= p.M) && (n0 + BN <= p.N);
        const long long elem0 = static_cast<long long>(et.row_base + sub) * p.ldd + n0 + seg * 8;
        et.out = reinterpret_cast<uint8_t*>(p.d) + (d_off + elem0) * 2;
        et.out2 = p.d2 != nullptr ? reinterpret_cast<uint8_t*>(p.d2) + (d_off + elem0) * 2 : nullptr;
        et.step_d = p.ldd * 8;
        et.aux = nullptr;
        et.step_aux = p.ld_aux * 8;
        if (p.aux != nullptr)
          et.aux = reinterpret_cast<const uint8_t*>(p.aux) +
                   (aux_off + static_cast<long long>(et.row_base + sub) * p.ld_aux + n0 + seg * 8) * 2;
        et.bias_regs = false;
        if (BN == 256 && p.bias != nullptr && !p.bias_fp32 && n0 + BN <= p.N &&
            (reinterpret_cast<uintptr_t>(p.bias) & 15) == 0) {
          et.bias_regs = true;       // 32 lanes x 8 columns = the tile's 256 columns
          et.bias = __ldg(reinterpret_cast<const uint4*>(reinterpret_cast<const __nv_bfloat16*>(p.bias) + n0) + lane);
        }
        if (et.aux != nullptr && p.act != 1) load_aux_group(p, et, areg, 0, lane);
      }
      mbar_wait(&tmem_full[acc], acc_phase);
      tcgen05_fence_after();
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(quad * 32) << 16) + static_cast<uint32_t>(acc * BN);
      // the accumulator stage is handed back to the MMA warp as soon as its last tcgen05.ld has completed -- before the
      // final group's global stores, which would otherwise sit in front of the arrive's release fence
      auto release_acc = [&]() {
        tcgen05_fence_before();
        __syncwarp();
        if (lane == 0) {
          if (CG == 1 || leader) {
            mbar_arrive(&tmem_empty[acc]);
          } else {
            mbar_arrive_remote(&tmem_empty[acc], 0);
          }
        }
      };
      if (p.split_k > 1) {
        // raw fp32 partial sums of this K-split: 128 contiguous bytes per thread and chunk
        float* wrow = p.ws + (static_cast<long long>(tile % p.split_k) * p.M + row) * p.N + n0;
#pragma unroll 1
        for (int c = 0; c < BN / 32; ++c) {
          uint32_t raw[32];
          tmem_ld_32x32b_x32(taddr + c * 32, raw);
          tmem_ld_wait();
          const int col0 = n0 + c * 32;
          if (row < p.M && col0 < p.N) {
            if (col0 + 32 <= p.N) {
#pragma unroll
              for (int q = 0; q < 8; ++q)
                *reinterpret_cast<uint4*>(wrow + c * 32 + 4 * q) = make_uint4(raw[4 * q], raw[4 * q + 1], raw[4 * q + 2], raw[4 * q + 3]);
            } else {
              for (int j = 0; j < 32 && col0 + j < p.N; ++j) wrow[c * 32 + j] = __uint_as_float(raw[j]);
            }
          }
        }
        release_acc();
      } else if (staged && p.tma_store) {
        // st_in doubles as the second output buffer unless this epilogue needs it (aux staging / pre-activation copy)
        const bool single = (et.aux != nullptr && p.act != 1) || (p.act == 1 && et.out2 != nullptr);
#pragma unroll 1
        for (int c2 = 0; c2 < num_groups; ++c2) {
          uint8_t* ob = single ? st_out : st_out + (store_groups & 1) * 4096;
          // the bulk store that last read this buffer (two groups ago, or the previous one in single mode) is done with it
          if (lane == 0) {
            if (single) tma_store_wait_read<0>(); else tma_store_wait_read<1>();
          }
          __syncwarp();
          epilogue_group_compute(p, et, taddr + c2 * 64, c2, num_groups, ob, st_out + 4096, lane, areg);
          if (c2 == num_groups - 1) release_acc();
          epilogue_group_store_tma(p, et, c2, ob, st_out + 4096, lane);
          ++store_groups;
        }
      } else if (staged) {
#pragma unroll 1
        for (int c2 = 0; c2 < num_groups; ++c2) {
          epilogue_group_compute(p, et, taddr + c2 * 64, c2, num_groups, st_out, st_out + 4096, lane, areg);
          if (c2 == num_groups - 1) release_acc();
          epilogue_group_store(p, et, c2, st_out, st_out + 4096, lane);
        }
      } else {
#pragma unroll 1
        for (int c = 0; c < BN / 32; ++c) {
          uint32_t raw[32];
          tmem_ld_32x32b_x32(taddr + c * 32, raw);
          tmem_ld_wait();
          epilogue_chunk(p, raw, row, n0 + c * 32, d_off, aux_off);
        }
        release_acc();
      }
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1;
    }
    if (lane == 0) tma_store_wait_read<0>();     // shared memory must outlive the bulk stores that read it
    __syncwarp();
  }

  tcgen05_fence_before();
  if constexpr (CG == 2) {
    cluster_sync();
  } else {
    __syncthreads();
  }
  if (warp == 1) {
    tcgen05_fence_after();
    tmem_dealloc<CG>(tmem_base, Cfg::TMEM_COLS);
  }
}

// Split-K finish: D = epilogue(alpha * sum_s ws[s]) for 8 consecutive columns per thread (N % 8 == 0).
__global__ void __launch_bounds__(256) splitk_reduce_kernel(const Params p) {
  const long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const int n8 = p.N >> 3;
  if (idx >= static_cast<long long>(p.M) * n8) return;
  const int row = static_cast<int>(idx / n8), col = static_cast<int>(idx % n8) * 8;
  float v[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  for (int s = 0; s < p.split_k; ++s) {
    const float* w = p.ws + (static_cast<long long>(s) * p.M + row) * p.N + col;
    const float4 a = *reinterpret_cast<const float4*>(w), b = *reinterpret_cast<const float4*>(w + 4);
    v[0] += a.x; v[1] += a.y; v[2] += a.z; v[3] += a.w; v[4] += b.x; v[5] += b.y; v[6] += b.z; v[7] += b.w;
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) v[j] *= p.alpha;
  if (p.bias != nullptr) {
#pragma unroll
    for (int j = 0; j < 8; ++j)
      v[j] += p.bias_fp32 ? reinterpret_cast<const float*>(p.bias)[col + j]
                          : __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(p.bias)[col + j]);
  }
  if (p.act == 1) {
    if (p.d2 != nullptr) {
      uint4 o = make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]), pack_bf16x2(v[6], v[7]));
      *reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(p.d2) + static_cast<long long>(row) * p.ldd + col) = o;
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = gelu_tanh(v[j]);
  } else if (p.aux != nullptr) {
    float a[8];
    if (p.aux_fp32) {
      const float* ap = reinterpret_cast<const float*>(p.aux) + static_cast<long long>(row) * p.ld_aux + col;
#pragma unroll
      for (int j = 0; j < 8; ++j) a[j] = ap[j];
    } else {
      const uint4 t = *reinterpret_cast<const uint4*>(reinterpret_cast<const __nv_bfloat16*>(p.aux) +
                                                      static_cast<long long>(row) * p.ld_aux + col);
      const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&t);
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float2 f = __bfloat1622float2(h[e]);
        a[2 * e] = f.x;
        a[2 * e + 1] = f.y;
      }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = (p.act == 2) ? v[j] * gelu_tanh_grad(a[j]) : v[j] + p.beta * a[j];
  }
  if (p.d_fp32) {
    float* dp = reinterpret_cast<float*>(p.d) + static_cast<long long>(row) * p.ldd + col;
    *reinterpret_cast<float4*>(dp) = make_float4(v[0], v[1], v[2], v[3]);
    *reinterpret_cast<float4*>(dp + 4) = make_float4(v[4], v[5], v[6], v[7]);
  } else {
    uint4 o = make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]), pack_bf16x2(v[6], v[7]));
    *reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(p.d) + static_cast<long long>(row) * p.ldd + col) = o;
  }
}

// ---------------------------------------------------------------------------
// host
// ---------------------------------------------------------------------------
static int make_operand_map(CUtensorMap* map, const void* base, int mn_major, int rows_mn, int k, long long ld,
                            int batch_inner, int batch_outer, long long stride_inner, long long stride_outer,
                            int box_mn_rows) {
  // K-major: global [MN rows, K cols], box [box_mn_rows x 64];  MN-major: global [K rows, MN cols], box [64 x 64]
  uint64_t dims[4];
  uint32_t box[4];
  if (mn_major) {
    dims[0] = static_cast<uint64_t>(rows_mn);
    dims[1] = static_cast<uint64_t>(k);
    box[0] = 64;
    box[1] = 64;
  } else {
    dims[0] = static_cast<uint64_t>(k);
    dims[1] = static_cast<uint64_t>(rows_mn);
    box[0] = 64;
    box[1] = static_cast<uint32_t>(box_mn_rows);
  }
  dims[2] = static_cast<uint64_t>(batch_inner);
  dims[3] = static_cast<uint64_t>(batch_outer);
  box[2] = 1;
  box[3] = 1;
  const uint64_t row_bytes = static_cast<uint64_t>(ld) * 2;
  uint64_t strides[3];
  strides[0] = row_bytes;
  strides[1] = batch_inner > 1 ? static_cast<uint64_t>(stride_inner) * 2 : row_bytes * dims[1];
  strides[2] = batch_outer > 1 ? static_cast<uint64_t>(stride_outer) * 2 : strides[1] * dims[2];
  return get_tensor_map_bf16(map, base, dims, strides, box);
}

template <int CG, int BN, int A_MN, int B_MN, bool GATHER = false>
static int launch(const CUtensorMap& ma, const CUtensorMap& mb, const Params& p, int grid, cudaStream_t stream) {
  using Cfg = Config<CG, BN>;
  auto kernel = gemm_kernel<CG, BN, A_MN, B_MN, GATHER>;
  static bool configured = false;
  if (!configured) {
    DTB_CUDA_OK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES));
    configured = true;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(static_cast<unsigned>(grid));
  cfg.blockDim = dim3(NUM_THREADS + (GATHER ? 32 * NUM_GATHER_WARPS : 0));
  cfg.dynamicSmemBytes = Cfg::SMEM_BYTES;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = CG;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  DTB_CUDA_OK(cudaLaunchKernelEx(&cfg, kernel, ma, mb, p));
  count_launch();
  return 0;
}

template <int CG, int BN>
static int dispatch_major(const CUtensorMap& ma, const CUtensorMap& mb, const Params& p, int a_mn, int b_mn, int grid,
                          cudaStream_t s) {
  if (!a_mn && !b_mn) return launch<CG, BN, 0, 0>(ma, mb, p, grid, s);
  if (!a_mn && b_mn) return launch<CG, BN, 0, 1>(ma, mb, p, grid, s);
  if (a_mn && !b_mn) return launch<CG, BN, 1, 0>(ma, mb, p, grid, s);
  return launch<CG, BN, 1, 1>(ma, mb, p, grid, s);
}

}  // namespace gemm
}  // namespace dtb

// One work counter per (device, stream) for dynamic tile scheduling.  Kernels on one stream run in order and every
// launch leaves its counter at zero, so a stream can share one counter between all its GEMMs; GEMMs on different
// streams may overlap and must not share.  First use on a stream allocates 256 bytes (not capturable: warm up first).
static int* sched_counter(cudaStream_t s) {
  static std::mutex mu;
  static std::map<std::pair<int, cudaStream_t>, int*> table;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return nullptr;
  std::lock_guard<std::mutex> lock(mu);
  auto it = table.find({dev, s});
  if (it != table.end()) return it->second;
  int* ptr = nullptr;
  if (cudaMalloc(&ptr, 256) != cudaSuccess) return nullptr;
  if (cudaMemset(ptr, 0, 256) != cudaSuccess) return nullptr;
  table[{dev, s}] = ptr;
  return ptr;
}

static int gemm_impl(const dtb_gemm_args* a, const dtb_gather_args* g, void* stream) {
  using namespace dtb;
  using namespace dtb::gemm;
  DTB_REQUIRE(a != nullptr, "dtb_gemm_bf16: null args");
  DTB_REQUIRE(a->m > 0 && a->n > 0 && a->k > 0, "dtb_gemm_bf16: empty problem %dx%dx%d", a->m, a->n, a->k);
  DTB_REQUIRE(a->lda % 8 == 0 && a->ldb % 8 == 0, "dtb_gemm_bf16: lda/ldb must be multiples of 8 elements");
  DTB_REQUIRE((reinterpret_cast<uintptr_t>(a->a) & 15) == 0 && (reinterpret_cast<uintptr_t>(a->b) & 15) == 0,
              "dtb_gemm_bf16: A/B must be 16-byte aligned");
  DTB_REQUIRE(a->ldd % (a->d_fp32 ? 4 : 8) == 0 && (reinterpret_cast<uintptr_t>(a->d) & 15) == 0,
              "dtb_gemm_bf16: D must be 16-byte aligned with a 16-byte-multiple pitch");
  if (a->aux)
    DTB_REQUIRE(a->ld_aux % (a->aux_fp32 ? 4 : 8) == 0 && (reinterpret_cast<uintptr_t>(a->aux) & 15) == 0,
                "dtb_gemm_bf16: aux must be 16-byte aligned with a 16-byte-multiple pitch");
  DTB_REQUIRE(a->act != 2 || a->aux != nullptr, "dtb_gemm_bf16: act=2 (gelu') needs aux");
  const int bi = a->batch_inner > 0 ? a->batch_inner : 1;
  const int bo = a->batch_outer > 0 ? a->batch_outer : 1;
  const int sms = sm_count();

  // tile selection: pair mode for big problems; 128-wide N tiles when 256 would leave SMs idle
  int cg = a->cta_pair;
  int bn = a->tile_n;
  const int num_kb_total = (a->k + BK - 1) / BK;
  // split-K needs a workspace, a plain (unbatched, non-gather) problem and 16-byte-granular rows
  const bool can_split = g == nullptr && bi * bo == 1 && a->workspace != nullptr && a->split_k != 1 && a->colsum == nullptr &&
                         (a->n % 8) == 0 && num_kb_total >= 8;
  if (bn == 0) {
    const long long tiles256 = 1LL * ((a->m + 127) / 128) * ((a->n + 255) / 256) * bi * bo;
    bn = (a->n > 128 && tiles256 >= sms) ? 256 : 128;
    if (a->n <= 128) bn = 128;
    // few output tiles but a long K: 256-wide tiles (half the operand traffic per FLOP) and split K instead
    if (can_split && a->cta_pair != 1 && a->m >= 256 && a->n >= 256) bn = 256;
  }
  DTB_REQUIRE(bn == 128 || bn == 256, "dtb_gemm_bf16: tile_n must be 128 or 256");
  if (cg == 0) {
    const long long tiles_pair = 1LL * ((a->m + 255) / 256) * ((a->n + bn - 1) / bn) * bi * bo;
    cg = (bn == 256 && a->m >= 256 && (tiles_pair >= sms / 2 || can_split)) ? 2 : 1;
    // very short K (< 8 k-blocks): the pair's extra synchronisation no longer pays; measured at 8192x2304xK, 1-CTA
    // 256-wide tiles win at K = 64 / 256 (24.6 vs 26.6, 27.4 vs 29.5 us) and lose from K = 768 up (35.5 vs 33.6 us)
    if (cg == 2 && num_kb_total < 8 && !can_split) cg = 1;
  }
  DTB_REQUIRE(cg == 1 || (cg == 2 && bn == 256), "dtb_gemm_bf16: cta_pair=2 requires tile_n=256");

  int split = 1, kb_per_split = num_kb_total;
  if (can_split) {
    const long long units = 1LL * ((a->m + BM * cg - 1) / (BM * cg)) * ((a->n + bn - 1) / bn);
    const long long slots = cg == 2 ? (sms / 2) : sms;
    // cost in k-block times: waves * (k-blocks per unit + pipeline fill / epilogue), plus the reduce pass
    const double fill = 6.0, reduce_cost = 8.0;
    double best = static_cast<double>((units + slots - 1) / slots) * (num_kb_total + fill);
    const int max_split = a->split_k > 1 ? a->split_k : 8;
    for (int sk = 2; sk <= max_split; ++sk) {
      const int kbps = (num_kb_total + sk - 1) / sk;
      if (kbps < 4) break;
      const int eff = (num_kb_total + kbps - 1) / kbps;
      if (static_cast<long long>(eff) * a->m * a->n * 4 > a->workspace_bytes) break;
      const double cost = static_cast<double>((units * eff + slots - 1) / slots) * (kbps + fill) + reduce_cost;
      if (cost < best * 0.9 || (a->split_k > 1 && sk == a->split_k)) {
        best = cost;
        split = eff;
        kb_per_split = kbps;
      }
    }
  }

  Params p = {};
  p.M = a->m; p.N = a->n; p.K = a->k;
  p.batch_inner = bi; p.batch = bi * bo;
  p.m_blocks = (a->m + BM * cg - 1) / (BM * cg);
  p.n_blocks = (a->n + bn - 1) / bn;
  p.group_m = cg == 2 ? 8 : 16;
  p.group_m_log2 = cg == 2 ? 3 : 4;
  p.rcp_per_batch = 1.0f / static_cast<float>(p.m_blocks * p.n_blocks);
  p.rcp_group_size = 1.0f / static_cast<float>(p.group_m * p.n_blocks);
  DTB_REQUIRE(1LL * p.m_blocks * p.n_blocks * bi * bo < (1LL << 22), "dtb_gemm_bf16: more than 4M output tiles");
  p.d = a->d; p.d2 = a->d2; p.aux = a->aux; p.bias = a->bias;
  p.ldd = a->ldd; p.ld_aux = a->ld_aux;
  p.d_si = a->d_stride_inner; p.d_so = a->d_stride_outer;
  p.aux_si = a->aux_stride_inner; p.aux_so = a->aux_stride_outer;
  p.d_fp32 = a->d_fp32; p.aux_fp32 = a->aux_fp32; p.bias_fp32 = a->bias_fp32;
  p.act = a->act;
  p.alpha = a->alpha; p.beta = a->beta;
  p.colsum = a->colsum;
  if (a->colsum != nullptr) {
    DTB_REQUIRE(bi * bo == 1 && !a->d_fp32 && !(a->aux != nullptr && a->aux_fp32) && a->n % 8 == 0 &&
                    (reinterpret_cast<uintptr_t>(a->colsum) & 15) == 0,
                "dtb_gemm_bf16: colsum needs an unbatched bf16 D with N %% 8 == 0 and a 16-byte aligned fp32 [N] buffer");
  }
  // canonical 128B-swizzle layouts (cute/atom/mma_traits_sm100.hpp):
  //   K-major : 8-row groups 1024 B apart (SBO); LBO unused (1)
  //   MN-major: 64-wide MN blocks 8192 B apart (LBO); 8-row K groups 1024 B apart (SBO)
  p.lbo_a = a->dbg_lbo_a ? a->dbg_lbo_a : (a->a_mn_major ? 8192u : 16u);
  p.sbo_a = a->dbg_sbo_a ? a->dbg_sbo_a : 1024u;
  p.lbo_b = a->dbg_lbo_b ? a->dbg_lbo_b : (a->b_mn_major ? 8192u : 16u);
  p.sbo_b = a->dbg_sbo_b ? a->dbg_sbo_b : 1024u;
  p.g_rotate = -1;
  p.split_k = split;
  p.kb_per_split = kb_per_split;
  p.ws = static_cast<float*>(a->workspace);
  {
    static const int direct = [] { const char* e = getenv("DTB_GEMM_DIRECT_EPILOGUE"); return (e && e[0] == '1') ? 1 : 0; }();
    p.direct_epilogue = a->colsum != nullptr ? 0 : direct;
  }
  if (g != nullptr) {
    DTB_REQUIRE(!a->a_mn_major && !a->b_mn_major && bi * bo == 1,
                "dtb_gemm_bf16_allgather: forward layout only (A, B K-major, no batch)");
    DTB_REQUIRE(g->world >= 1 && g->world <= DTB_MAX_RANKS && g->rank >= 0 && g->rank < g->world,
                "dtb_gemm_bf16_allgather: bad rank/world");
    DTB_REQUIRE(g->rows_per_rank > 0 && g->rows_per_rank % GATHER_ROWS == 0 &&
                    static_cast<long long>(g->rows_per_rank) * g->world >= a->n,
                "dtb_gemm_bf16_allgather: rows_per_rank must be a multiple of %d and cover N", GATHER_ROWS);
    DTB_REQUIRE(g->flags != nullptr, "dtb_gemm_bf16_allgather: flags scratch is required");
    DTB_REQUIRE(a->k % 8 == 0, "dtb_gemm_bf16_allgather: K must be a multiple of 8");
    for (int r = 0; r < g->world; ++r) {
      DTB_REQUIRE(g->peer_shard[r] != nullptr && (reinterpret_cast<uintptr_t>(g->peer_shard[r]) & 15) == 0,
                  "dtb_gemm_bf16_allgather: shard pointer of rank %d missing or unaligned", r);
      p.g_peer_shard[r] = g->peer_shard[r];
    }
    p.g_full = const_cast<void*>(a->b);
    p.g_flags = g->flags;
    p.g_ldb = a->ldb;
    p.g_world = g->world; p.g_rank = g->rank; p.g_rows_per_rank = g->rows_per_rank;
    // own rows first: start the n-block walk at the block holding this rank's first row
    p.g_rotate = static_cast<int>((static_cast<long long>(g->rank) * g->rows_per_rank / bn) % p.n_blocks);
    DTB_CUDA_OK(cudaMemsetAsync(g->flags, 0, sizeof(int) * (p.n_blocks + 1), static_cast<cudaStream_t>(stream)));
  }

  {
    // TMA-store epilogue (DTB_GEMM_TMA_STORE=0: per-lane LDS + STG hand-off): same conditions as the staged epilogue
    static const int tma_store = [] { const char* e = getenv("DTB_GEMM_TMA_STORE"); return (e && e[0] == '0') ? 0 : 1; }();
    const bool staged = p.split_k == 1 && !p.direct_epilogue && !p.d_fp32 && !(p.aux != nullptr && p.aux_fp32) && (p.N & 7) == 0;
    const bool strides_ok = (bi <= 1 || a->d_stride_inner % 8 == 0) && (bo <= 1 || a->d_stride_outer % 8 == 0);
    p.tma_store = 0;
    // long-K GEMMs hide their epilogue anyway, and in the power-capped wide-MLP step the per-lane hand-off measured
    // 0.7 % faster (9.745 vs 9.82 ms on the same box), so the TMA unit takes over only below K = 4096
    if (tma_store && staged && strides_ok && a->k < 4096 && a->colsum == nullptr) {
      uint64_t dims[4] = {static_cast<uint64_t>(a->n), static_cast<uint64_t>(a->m), static_cast<uint64_t>(bi),
                          static_cast<uint64_t>(bo)};
      uint32_t box[4] = {64, 32, 1, 1};
      const uint64_t row_bytes = static_cast<uint64_t>(a->ldd) * 2;
      uint64_t strides[3];
      strides[0] = row_bytes;
      strides[1] = bi > 1 ? static_cast<uint64_t>(a->d_stride_inner) * 2 : row_bytes * dims[1];
      strides[2] = bo > 1 ? static_cast<uint64_t>(a->d_stride_outer) * 2 : strides[1] * dims[2];
      if (get_tensor_map_bf16(&p.map_d, a->d, dims, strides, box)) return 1;
      if (a->d2 != nullptr && a->act == 1 && get_tensor_map_bf16(&p.map_d2, a->d2, dims, strides, box)) return 1;
      p.tma_store = 1;
    }
  }

  CUtensorMap ma, mb;
  if (make_operand_map(&ma, a->a, a->a_mn_major, a->m, a->k, a->lda, bi, bo, a->a_stride_inner, a->a_stride_outer, BM))
    return 1;
  if (make_operand_map(&mb, a->b, a->b_mn_major, a->n, a->k, a->ldb, bi, bo, a->b_stride_inner, a->b_stride_outer,
                       bn / cg))
    return 1;

  const long long tiles = 1LL * p.m_blocks * p.n_blocks * p.batch * p.split_k;
  int max_ctas = a->max_ctas > 0 ? (a->max_ctas < sms ? a->max_ctas : sms) : sms;
  int grid;
  if (cg == 2) {
    long long clusters = max_ctas / 2;
    if (clusters < 1) clusters = 1;
    if (tiles < clusters) clusters = tiles;
    grid = static_cast<int>(clusters * 2);
  } else {
    grid = static_cast<int>(tiles < max_ctas ? tiles : max_ctas);
  }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  {
    // dynamic scheduling whenever a cluster has more than one tile to do (DTB_GEMM_DYNAMIC=0: static round-robin)
    static const int dynamic = [] { const char* e = getenv("DTB_GEMM_DYNAMIC"); return (e && e[0] == '0') ? 0 : 1; }();
    const long long clusters = cg == 2 ? grid / 2 : grid;
    if (dynamic && g == nullptr && tiles > clusters) p.sched = sched_counter(s);
  }
  if (g != nullptr) {
    if (cg == 2) return launch<2, 256, 0, 0, true>(ma, mb, p, grid, s);
    if (bn == 256) return launch<1, 256, 0, 0, true>(ma, mb, p, grid, s);
    return launch<1, 128, 0, 0, true>(ma, mb, p, grid, s);
  }
  int rc;
  if (cg == 2) rc = dispatch_major<2, 256>(ma, mb, p, a->a_mn_major, a->b_mn_major, grid, s);
  else if (bn == 256) rc = dispatch_major<1, 256>(ma, mb, p, a->a_mn_major, a->b_mn_major, grid, s);
  else rc = dispatch_major<1, 128>(ma, mb, p, a->a_mn_major, a->b_mn_major, grid, s);
  if (rc != 0 || p.split_k == 1) return rc;
  const long long vecs = static_cast<long long>(p.M) * (p.N / 8);
  splitk_reduce_kernel<<<static_cast<unsigned>((vecs + 255) / 256), 256, 0, s>>>(p);
  DTB_LAUNCH_OK();
  return 0;
}

extern "C" int dtb_gemm_bf16(const dtb_gemm_args* a, void* stream) { return gemm_impl(a, nullptr, stream); }

extern "C" int dtb_gemm_bf16_allgather(const dtb_gemm_args* a, const dtb_gather_args* g, void* stream) {
  DTB_REQUIRE(g != nullptr, "dtb_gemm_bf16_allgather: null gather args");
  return gemm_impl(a, g, stream);
}

extern "C" size_t dtb_gemm_allgather_flag_ints(int n) { return static_cast<size_t>((n + 127) / 128 + 1); }
