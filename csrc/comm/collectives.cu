// Peer-memory collectives over NVLink 5 / NVSwitch.
//
// Every rank maps every other rank's symmetric heap (see symm.cu), so a
// collective is ONE kernel per rank that loads from / stores to peer addresses
// directly.  Ranks meet at flag barriers kept in a per-rank signal pad:
// rank s tells rank d "I reached step q of block b on channel c" by storing q
// (st.release.sys) into d's pad at [c][b][s]; d spins with ld.acquire.sys.
// Sequence numbers only grow (the next value is kept in the pad as well), so
// flags are never reset and back-to-back launches cannot confuse each other.
// A bounded spin turns a dead peer into an error flag instead of a GPU hang: the CTA that timed out records the
// missing peer in ERR_SLOT and the whole CTA leaves the kernel WITHOUT touching the payload; every later collective on
// that communicator sees the flag at entry and returns at once, and the host (Trainer) raises at its next check.
//
// NVLS: when the heap is VMM-backed and bound to a multicast object (symm.cu), `mc_base` is the multicast mapping of
// the same bytes.  multimem.ld_reduce on it returns the sum over all ranks computed inside the NVSwitch (fp32
// accumulation), multimem.st writes all ranks' copies with one store.  The *_nvls kernels use those instead of W peer
// loads + W peer stores per vector; the flag barriers stay on unicast peer addresses.
//
//   all-reduce (two-shot)   rank r reduces slice r from all peers in fixed rank
//                           order (fp32), scales by 1/W, casts, and stores the
//                           result into every peer's buffer: NCCL all-reduce +
//                           div_ + dtype cast in one pass.
//   all-reduce (one-shot)   small buffers: every rank reduces everything locally.
//   broadcast               root stores its bytes into every peer (K2/K3).
//   all-gather              every rank stores its shard into every peer (K6/K8).
//   reduce-scatter          owner pulls its slice from all peers (K7).
//   rs_optim                reduce-scatter fused with SGD / AdamW on the fp32
//                           master shard and the bf16 shard write-back (K7+K13).
//
// Reference call sites: dist.broadcast / dist.all_reduce in
// /root/reference/src/playground/ddp_script.py:120-121,149-154 and the implicit
// NCCL collectives of torch DDP / FSDP (ddp_strategy.py:18, fsdp_strategy.py:20).
#include <cstdlib>
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include "../../include/dtb200.h"
#include "../common/host.h"
#include "../common/ptx.cuh"

namespace dtb {
namespace comm {

constexpr int MAX_BLOCKS = 128;      // barrier slots per channel
constexpr int NUM_CHANNELS = 8;
constexpr int THREADS = 256;      // 256 threads x <= 64 registers: a comm CTA fits on an SM NEXT TO a resident GEMM CTA
constexpr int MIN_CTAS_PER_SM = 4; // launch-bounds hint that caps registers at 65536 / (256 * 4) = 64, leaving 192 x 255 for the GEMM
// signal pad layout (uint32 slots)
constexpr int FLAG_SLOTS = NUM_CHANNELS * MAX_BLOCKS * DTB_MAX_RANKS;   // 8192
constexpr int SEQ_BASE = FLAG_SLOTS;                                     // [channel][block] next sequence
constexpr int ERR_SLOT = DTB_SIGNAL_PAD_BYTES / 4 - 1;
constexpr long long DEFAULT_TIMEOUT = 40000000000LL;                     // ~20 s of SM clocks

struct Ctx {
  dtb_comm c;
  int channel;
};

__device__ __forceinline__ int flag_slot(int channel, int block, int src_rank) {
  return (channel * MAX_BLOCKS + block) * DTB_MAX_RANKS + src_rank;
}

// Block-level barrier across ranks.  All threads of the block call it; on return
// every peer's block with the same index has arrived and its earlier stores are
// visible here.  `seq` is this block's running sequence number (register copy).
__device__ __forceinline__ bool rank_barrier(const Ctx& x, uint32_t& seq) {
  __shared__ int s_dead;
  if (threadIdx.x == 0) s_dead = 0;
  __syncthreads();
  seq += 1;
  if (threadIdx.x < static_cast<unsigned>(x.c.world)) {
    const int peer = threadIdx.x;
    st_release_sys(x.c.signals[peer] + flag_slot(x.channel, blockIdx.x, x.c.rank), seq);
    const uint32_t* mine = x.c.signals[x.c.rank] + flag_slot(x.channel, blockIdx.x, peer);
    const long long limit = x.c.timeout_cycles > 0 ? x.c.timeout_cycles : DEFAULT_TIMEOUT;
    const long long t0 = clock64();
    while (static_cast<int32_t>(ld_acquire_sys(mine) - seq) < 0) {
      if (clock64() - t0 > limit) {
        x.c.signals[x.c.rank][ERR_SLOT] = 1u + static_cast<uint32_t>(peer);
        __threadfence_system();
        s_dead = 1;
        break;
      }
    }
  }
  __syncthreads();
  return s_dead == 0;
}

// True when an earlier collective on this communicator already timed out: nothing may touch the payload any more.
__device__ __forceinline__ bool comm_dead(const Ctx& x) {
  return *reinterpret_cast<const volatile uint32_t*>(x.c.signals[x.c.rank] + ERR_SLOT) != 0u;
}

__device__ __forceinline__ uint32_t load_seq(const Ctx& x) {
  return x.c.signals[x.c.rank][SEQ_BASE + x.channel * MAX_BLOCKS + blockIdx.x];
}
__device__ __forceinline__ void store_seq(const Ctx& x, uint32_t seq) {
  if (threadIdx.x == 0) x.c.signals[x.c.rank][SEQ_BASE + x.channel * MAX_BLOCKS + blockIdx.x] = seq;
}

// ---- vector helpers ---------------------------------------------------------
template <bool FP32>
struct Vec;   // 16 bytes of elements as floats

template <>
struct Vec<false> {
  static constexpr int N = 8;
  __device__ static void unpack(const uint4& r, float* f) {
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&r);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float2 t = __bfloat1622float2(h[i]);
      f[2 * i] = t.x;
      f[2 * i + 1] = t.y;
    }
  }
  __device__ static uint4 pack(const float* f) {
    uint4 r;
    __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&r);
#pragma unroll
    for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
    return r;
  }
};
template <>
struct Vec<true> {
  static constexpr int N = 4;
  __device__ static void unpack(const uint4& r, float* f) {
    f[0] = __uint_as_float(r.x); f[1] = __uint_as_float(r.y); f[2] = __uint_as_float(r.z); f[3] = __uint_as_float(r.w);
  }
  __device__ static uint4 pack(const float* f) {
    return make_uint4(__float_as_uint(f[0]), __float_as_uint(f[1]), __float_as_uint(f[2]), __float_as_uint(f[3]));
  }
};

// Sum of the same 16-byte vector on every rank, fixed order 0..W-1, all loads in flight first.
template <bool FP32, int W>
__device__ __forceinline__ void reduce_vec(const Ctx& x, size_t byte_off, float* acc) {
  uint4 raw[W];
#pragma unroll
  for (int r = 0; r < W; ++r) raw[r] = ld_volatile_v4(static_cast<const uint8_t*>(x.c.peers[r]) + byte_off);
#pragma unroll
  for (int j = 0; j < Vec<FP32>::N; ++j) acc[j] = 0.f;
#pragma unroll
  for (int r = 0; r < W; ++r) {
    float f[Vec<FP32>::N];
    Vec<FP32>::unpack(raw[r], f);
#pragma unroll
    for (int j = 0; j < Vec<FP32>::N; ++j) acc[j] += f[j];
  }
}

// ---- kernels ------------------------------------------------------------------
__global__ void __launch_bounds__(THREADS, MIN_CTAS_PER_SM) barrier_kernel(const Ctx x) {
  if (comm_dead(x)) return;
  uint32_t seq = load_seq(x);
  if (!rank_barrier(x, seq)) return;
  store_seq(x, seq);
}

// Loads of U consecutive grid-strided vectors from all W ranks are issued before any is consumed:
// U*W = 8 16-byte requests per thread x 64 CTAs x 256 threads keep ~2 MB in flight per GPU, what NVLink's ~2 us
// round trip needs to stream at full rate.
template <int W>
struct Unroll {
  static constexpr int U = W >= 8 ? 1 : (W >= 4 ? 2 : 4);
};

struct SgdArgs {
  size_t param_offset;   // params live in the heap with the same slot layout as the gradient bucket
  float lr, decay;       // p = p * decay - lr * g   (decay = 1 - lr * weight_decay)
};

// Two-shot in-place mean.  nvec = number of 16-byte vectors in the buffer.
// SGD = true: instead of writing the mean gradient back, rank r applies the SGD update to slice r of the
// parameters and stores the *new parameters* into every rank (gradient all-reduce + 1/W + optimizer step
// in one pass over NVLink; bf16 only).
template <bool FP32, int W, bool SGD>
__global__ void __launch_bounds__(THREADS, MIN_CTAS_PER_SM)
allreduce_twoshot_kernel(const Ctx x, size_t offset, long long nvec, const SgdArgs sgd) {
  constexpr int U = Unroll<W>::U;
  constexpr int N = Vec<FP32>::N;
  if (comm_dead(x)) return;
  uint32_t seq = load_seq(x);
  if (!rank_barrier(x, seq)) return;   // every rank's gradients are final (and, for SGD, nobody still reads the old params)
  const long long lo = nvec * x.c.rank / W, hi = nvec * (x.c.rank + 1) / W;
  const float scale = 1.f / W;
  const long long stride = static_cast<long long>(gridDim.x) * THREADS;
  for (long long v0 = lo + static_cast<long long>(blockIdx.x) * THREADS + threadIdx.x; v0 < hi; v0 += stride * U) {
    uint4 raw[U][W];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long v = v0 + u * stride;
      if (v < hi) {
#pragma unroll
        for (int r = 0; r < W; ++r)
          raw[u][r] = ld_volatile_v4(static_cast<const uint8_t*>(x.c.peers[r]) + offset + static_cast<size_t>(v) * 16);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long v = v0 + u * stride;
      if (v >= hi) break;
      float acc[N];
#pragma unroll
      for (int j = 0; j < N; ++j) acc[j] = 0.f;
#pragma unroll
      for (int r = 0; r < W; ++r) {
        float f[N];
        Vec<FP32>::unpack(raw[u][r], f);
#pragma unroll
        for (int j = 0; j < N; ++j) acc[j] += f[j];
      }
      size_t dst = offset + static_cast<size_t>(v) * 16;
      if (SGD) {
        dst = sgd.param_offset + static_cast<size_t>(v) * 16;
        const uint4 praw = *reinterpret_cast<const uint4*>(static_cast<const uint8_t*>(x.c.peers[x.c.rank]) + dst);
        float pf[N];
        Vec<FP32>::unpack(praw, pf);
#pragma unroll
        for (int j = 0; j < N; ++j) acc[j] = pf[j] * sgd.decay - sgd.lr * (acc[j] * scale);
      } else {
#pragma unroll
        for (int j = 0; j < N; ++j) acc[j] *= scale;
      }
      const uint4 out = Vec<FP32>::pack(acc);
#pragma unroll
      for (int r = 0; r < W; ++r) st_v4(static_cast<uint8_t*>(x.c.peers[r]) + dst, out);
    }
  }
  if (!rank_barrier(x, seq)) return;   // every slice has been delivered to this rank
  store_seq(x, seq);
}

// 32-byte accesses with "no L1 allocate, evict-first in L2" (LDG/STG.E.NA.EFL2.256): half the instructions of the
// 16-byte path, and the gradient stream does not push the concurrently running GEMM's operands out of L2.
struct V8 {
  uint32_t w[8];
};
__device__ __forceinline__ V8 ld256_stream(const void* p) {
  V8 r;
  asm volatile("ld.relaxed.sys.global.L1::no_allocate.L2::evict_first.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(r.w[0]), "=r"(r.w[1]), "=r"(r.w[2]), "=r"(r.w[3]), "=r"(r.w[4]), "=r"(r.w[5]), "=r"(r.w[6]), "=r"(r.w[7])
               : "l"(p)
               : "memory");
  return r;
}
__device__ __forceinline__ void st256_stream(void* p, const V8& v) {
  asm volatile("st.global.L1::no_allocate.L2::evict_first.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "r"(v.w[0]),
               "r"(v.w[1]), "r"(v.w[2]), "r"(v.w[3]), "r"(v.w[4]), "r"(v.w[5]), "r"(v.w[6]), "r"(v.w[7])
               : "memory");
}

// bf16 two-shot mean (optionally fused with SGD) on 32-byte vectors; same semantics and reduction order as
// allreduce_twoshot_kernel.  Peers are read four at a time to stay within the 64-register budget.
//
// NT = 256: co-resident mode (fits next to a GEMM CTA).  NT = 1024: exclusive mode -- a CTA takes a whole SM (1024
// threads x 64 registers = the entire register file), is launched in clusters of two so a pair fills one TPC, and a
// handful of them (16 by default) carry the whole reduction.  The GEMM's dynamic tile scheduler routes work around the
// SMs they occupy; measured on B200 that costs the GEMM ~11 % while the reduction runs instead of 20-35 % for
// co-resident CTAs that fight the GEMM's TMA traffic for the SM's L1/shared-memory datapath (profiles/README.md).
template <int W, bool SGD, int NT>
__global__ void __launch_bounds__(NT, NT == THREADS ? MIN_CTAS_PER_SM : 1)
allreduce_twoshot256_kernel(const Ctx x, size_t offset, long long nvec32, const SgdArgs sgd) {
  constexpr int G = W < 4 ? W : 4;          // loads in flight per batch
  constexpr int U = W <= 2 ? 2 : 1;
  if (comm_dead(x)) return;
  uint32_t seq = load_seq(x);
  if (!rank_barrier(x, seq)) return;
  const long long lo = nvec32 * x.c.rank / W, hi = nvec32 * (x.c.rank + 1) / W;
  const float scale = 1.f / W;
  const long long stride = static_cast<long long>(gridDim.x) * NT;
  for (long long v0 = lo + static_cast<long long>(blockIdx.x) * NT + threadIdx.x; v0 < hi; v0 += stride * U) {
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long v = v0 + u * stride;
      if (v >= hi) break;
      const size_t off = offset + static_cast<size_t>(v) * 32;
      float acc[16];
#pragma unroll
      for (int j = 0; j < 16; ++j) acc[j] = 0.f;
#pragma unroll
      for (int r0 = 0; r0 < W; r0 += G) {
        V8 raw[G];
#pragma unroll
        for (int g = 0; g < G; ++g) raw[g] = ld256_stream(static_cast<const uint8_t*>(x.c.peers[r0 + g]) + off);
#pragma unroll
        for (int g = 0; g < G; ++g) {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&raw[g].w[j]));
            acc[2 * j] += f.x;
            acc[2 * j + 1] += f.y;
          }
        }
      }
      size_t dst = off;
      if (SGD) {
        dst = sgd.param_offset + static_cast<size_t>(v) * 32;
        const V8 praw = ld256_stream(static_cast<const uint8_t*>(x.c.peers[x.c.rank]) + dst);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&praw.w[j]));
          acc[2 * j] = f.x * sgd.decay - sgd.lr * (acc[2 * j] * scale);
          acc[2 * j + 1] = f.y * sgd.decay - sgd.lr * (acc[2 * j + 1] * scale);
        }
      } else {
#pragma unroll
        for (int j = 0; j < 16; ++j) acc[j] *= scale;
      }
      V8 out;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        __nv_bfloat162 h = __floats2bfloat162_rn(acc[2 * j], acc[2 * j + 1]);
        out.w[j] = *reinterpret_cast<uint32_t*>(&h);
      }
#pragma unroll
      for (int r = 0; r < W; ++r) st256_stream(static_cast<uint8_t*>(x.c.peers[r]) + dst, out);
    }
  }
  if (!rank_barrier(x, seq)) return;
  store_seq(x, seq);
}

// One-shot in-place mean for small buffers: each rank reduces the whole buffer into
// registers, all ranks meet (nobody may overwrite data a peer still reads), then store.
template <bool FP32, int W, int PER_THREAD>
__global__ void __launch_bounds__(THREADS, MIN_CTAS_PER_SM) allreduce_oneshot_kernel(const Ctx x, size_t offset, long long nvec) {
  if (comm_dead(x)) return;
  uint32_t seq = load_seq(x);
  if (!rank_barrier(x, seq)) return;
  float acc[PER_THREAD][Vec<FP32>::N];
  const long long base = static_cast<long long>(blockIdx.x) * THREADS * PER_THREAD + threadIdx.x;
  const float scale = 1.f / W;
#pragma unroll
  for (int i = 0; i < PER_THREAD; ++i) {
    const long long v = base + static_cast<long long>(i) * THREADS;
    if (v < nvec) reduce_vec<FP32, W>(x, offset + static_cast<size_t>(v) * 16, acc[i]);
  }
  if (!rank_barrier(x, seq)) return;
#pragma unroll
  for (int i = 0; i < PER_THREAD; ++i) {
    const long long v = base + static_cast<long long>(i) * THREADS;
    if (v < nvec) {
#pragma unroll
      for (int j = 0; j < Vec<FP32>::N; ++j) acc[i][j] *= scale;
      st_v4(static_cast<uint8_t*>(x.c.peers[x.c.rank]) + offset + static_cast<size_t>(v) * 16, Vec<FP32>::pack(acc[i]));
    }
  }
  store_seq(x, seq);
}

__global__ void __launch_bounds__(THREADS, MIN_CTAS_PER_SM) broadcast_kernel(const Ctx x, size_t offset, long long nvec, int root) {
  if (comm_dead(x)) return;
  uint32_t seq = load_seq(x);
  if (!rank_barrier(x, seq)) return;   // destinations are free to be overwritten
  if (x.c.rank == root) {
    const uint8_t* src = static_cast<const uint8_t*>(x.c.peers[root]) + offset;
    for (long long v = static_cast<long long>(blockIdx.x) * THREADS + threadIdx.x; v < nvec;
         v += static_cast<long long>(gridDim.x) * THREADS) {
      const uint4 val = *reinterpret_cast<const uint4*>(src + v * 16);
      for (int r = 0; r < x.c.world; ++r)
        if (r != root) st_v4(static_cast<uint8_t*>(x.c.peers[r]) + offset + v * 16, val);
    }
  }
  if (!rank_barrier(x, seq)) return;
  store_seq(x, seq);
}

// Push all-gather: rank r stores its shard (nvec vectors at src_off) to dst_off + r*shard on every rank.
__global__ void __launch_bounds__(THREADS, MIN_CTAS_PER_SM) allgather_kernel(const Ctx x, size_t dst_off, size_t src_off, long long nvec) {
  if (comm_dead(x)) return;
  uint32_t seq = load_seq(x);
  if (!rank_barrier(x, seq)) return;
  const uint8_t* src = static_cast<const uint8_t*>(x.c.peers[x.c.rank]) + src_off;
  const size_t slot = dst_off + static_cast<size_t>(x.c.rank) * static_cast<size_t>(nvec) * 16;
  for (long long v = static_cast<long long>(blockIdx.x) * THREADS + threadIdx.x; v < nvec;
       v += static_cast<long long>(gridDim.x) * THREADS) {
    const uint4 val = *reinterpret_cast<const uint4*>(src + v * 16);
    for (int r = 0; r < x.c.world; ++r) st_v4(static_cast<uint8_t*>(x.c.peers[r]) + slot + v * 16, val);
  }
  if (!rank_barrier(x, seq)) return;
  store_seq(x, seq);
}


// ---- NVLS (multimem) variants ---------------------------------------------------------------------------------------
// `mc` addresses are offsets into dtb_comm::mc_base, the multicast mapping of the symmetric segment.
__device__ __forceinline__ uint4 mm_ld_reduce_bf16(const void* mc) {
  uint4 r;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(mc)
               : "memory");
  return r;
}
__device__ __forceinline__ uint4 mm_ld_reduce_f32(const void* mc) {
  uint4 r;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(mc)
               : "memory");
  return r;
}
template <bool FP32>
__device__ __forceinline__ uint4 mm_ld_reduce(const void* mc) {
  return FP32 ? mm_ld_reduce_f32(mc) : mm_ld_reduce_bf16(mc);
}
__device__ __forceinline__ void mm_st(void* mc, const uint4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "r"(v.x), "r"(v.y), "r"(v.z),
               "r"(v.w)
               : "memory");
}

// Two-shot mean (optionally fused with SGD) through the switch: rank r asks the NVSwitch for the sum of slice r over
// all ranks (one multimem.ld_reduce per 16 bytes instead of W peer loads), scales / applies the update, and writes the
// result into every rank with one multimem.st (instead of W peer stores).  Per GPU and direction that is (1 + 1/W) x
// the buffer over NVLink instead of 2 (W-1)/W x, and 2 memory instructions per vector instead of 2 W, so far fewer
// SM cycles are taken from the GEMMs that run beside it.  The sum is formed in fp32 inside the switch in an order the
// hardware chooses: all ranks receive the SAME value (one reduction, broadcast), but it may differ in the last bit
// from the fixed-rank-order sum of the peer-load kernels.
// NT = 256: co-resident CTAs; NT = 1024: exclusive whole-SM CTAs (see allreduce_twoshot256_kernel).
template <bool FP32, bool SGD, int NT>
__global__ void __launch_bounds__(NT, NT == THREADS ? MIN_CTAS_PER_SM : 1)
allreduce_nvls_kernel(const Ctx x, size_t offset, long long nvec, const SgdArgs sgd) {
  constexpr int U = 4;
  constexpr int N = Vec<FP32>::N;
  if (comm_dead(x)) return;
  uint32_t seq = load_seq(x);
  if (!rank_barrier(x, seq)) return;
  const int W = x.c.world;
  const long long lo = nvec * x.c.rank / W, hi = nvec * (x.c.rank + 1) / W;
  const float scale = 1.f / static_cast<float>(W);
  uint8_t* mc = static_cast<uint8_t*>(x.c.mc_base);
  const uint8_t* mine = static_cast<const uint8_t*>(x.c.peers[x.c.rank]);
  const long long stride = static_cast<long long>(gridDim.x) * NT;
  for (long long v0 = lo + static_cast<long long>(blockIdx.x) * NT + threadIdx.x; v0 < hi; v0 += stride * U) {
    uint4 raw[U];
    uint4 praw[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long v = v0 + u * stride;
      if (v < hi) {
        raw[u] = mm_ld_reduce<FP32>(mc + offset + static_cast<size_t>(v) * 16);
        if (SGD) praw[u] = *reinterpret_cast<const uint4*>(mine + sgd.param_offset + static_cast<size_t>(v) * 16);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long v = v0 + u * stride;
      if (v >= hi) break;
      float acc[N];
      Vec<FP32>::unpack(raw[u], acc);
      size_t dst = offset + static_cast<size_t>(v) * 16;
      if (SGD) {
        dst = sgd.param_offset + static_cast<size_t>(v) * 16;
        float pf[N];
        Vec<FP32>::unpack(praw[u], pf);
#pragma unroll
        for (int j = 0; j < N; ++j) acc[j] = pf[j] * sgd.decay - sgd.lr * (acc[j] * scale);
      } else {
#pragma unroll
        for (int j = 0; j < N; ++j) acc[j] *= scale;
      }
      mm_st(mc + dst, Vec<FP32>::pack(acc));
    }
  }
  if (!rank_barrier(x, seq)) return;
  store_seq(x, seq);
}

// Broadcast: the root writes every rank's copy (its own included) with one multimem.st per vector: the root's link
// carries the buffer once instead of W-1 times.
__global__ void __launch_bounds__(THREADS, MIN_CTAS_PER_SM)
broadcast_nvls_kernel(const Ctx x, size_t offset, long long nvec, int root) {
  if (comm_dead(x)) return;
  uint32_t seq = load_seq(x);
  if (!rank_barrier(x, seq)) return;
  if (x.c.rank == root) {
    const uint8_t* src = static_cast<const uint8_t*>(x.c.peers[root]) + offset;
    uint8_t* mc = static_cast<uint8_t*>(x.c.mc_base) + offset;
    const long long stride = static_cast<long long>(gridDim.x) * THREADS;
    for (long long v0 = static_cast<long long>(blockIdx.x) * THREADS + threadIdx.x; v0 < nvec; v0 += stride * 4) {
      uint4 val[4];
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (v0 + u * stride < nvec) val[u] = *reinterpret_cast<const uint4*>(src + (v0 + u * stride) * 16);
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (v0 + u * stride < nvec) mm_st(mc + (v0 + u * stride) * 16, val[u]);
    }
  }
  if (!rank_barrier(x, seq)) return;
  store_seq(x, seq);
}

// All-gather: rank r writes its shard into slot r of every rank with one multimem.st per vector.
__global__ void __launch_bounds__(THREADS, MIN_CTAS_PER_SM)
allgather_nvls_kernel(const Ctx x, size_t dst_off, size_t src_off, long long nvec) {
  if (comm_dead(x)) return;
  uint32_t seq = load_seq(x);
  if (!rank_barrier(x, seq)) return;
  const uint8_t* src = static_cast<const uint8_t*>(x.c.peers[x.c.rank]) + src_off;
  uint8_t* mc = static_cast<uint8_t*>(x.c.mc_base) + dst_off + static_cast<size_t>(x.c.rank) * static_cast<size_t>(nvec) * 16;
  const long long stride = static_cast<long long>(gridDim.x) * THREADS;
  for (long long v0 = static_cast<long long>(blockIdx.x) * THREADS + threadIdx.x; v0 < nvec; v0 += stride * 4) {
    uint4 val[4];
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (v0 + u * stride < nvec) val[u] = *reinterpret_cast<const uint4*>(src + (v0 + u * stride) * 16);
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (v0 + u * stride < nvec) mm_st(mc + (v0 + u * stride) * 16, val[u]);
  }
  if (!rank_barrier(x, seq)) return;
  store_seq(x, seq);
}

struct OptArgs {
  float* master; float* m; float* v;
  __nv_bfloat16* shard;       // local bf16 shard (in the symmetric heap)
  int opt;                    // 0 sgd, 1 adamw
  float lr, beta1, beta2, eps, wd, bc1, bc2, momentum;
};

// Reduce-scatter (mean) of slice `rank`; MODE 0: write fp32/bf16 result to `out`; MODE 1: fused optimizer.
// NVLS = true: the W peer loads of a vector become one multimem.ld_reduce on the multicast mapping (summed in the
// switch, W then only sets the 1/W scale).
template <bool FP32, int W, int MODE, bool NVLS = false>
__global__ void __launch_bounds__(THREADS, MIN_CTAS_PER_SM)
reduce_scatter_kernel(const Ctx x, size_t src_off, long long nvec, void* out, int out_fp32, const OptArgs o) {
  if (comm_dead(x)) return;
  uint32_t seq = load_seq(x);
  if (!rank_barrier(x, seq)) return;
  const size_t slice = src_off + static_cast<size_t>(x.c.rank) * static_cast<size_t>(nvec) * 16;
  const float scale = 1.f / W;
  constexpr int N = Vec<FP32>::N;
  for (long long v = static_cast<long long>(blockIdx.x) * THREADS + threadIdx.x; v < nvec;
       v += static_cast<long long>(gridDim.x) * THREADS) {
    float g[N];
    if (NVLS) {
      Vec<FP32>::unpack(mm_ld_reduce<FP32>(static_cast<const uint8_t*>(x.c.mc_base) + slice + static_cast<size_t>(v) * 16), g);
    } else {
      reduce_vec<FP32, W>(x, slice + static_cast<size_t>(v) * 16, g);
    }
#pragma unroll
    for (int j = 0; j < N; ++j) g[j] *= scale;
    const long long e = v * N;
    if (MODE == 0) {
      if (out_fp32) {
#pragma unroll
        for (int j = 0; j < N; j += 4)
          *reinterpret_cast<float4*>(static_cast<float*>(out) + e + j) = make_float4(g[j], g[j + 1], g[j + 2], g[j + 3]);
      } else {
        __nv_bfloat16* ob = static_cast<__nv_bfloat16*>(out) + e;
#pragma unroll
        for (int j = 0; j < N; j += 2) *reinterpret_cast<__nv_bfloat162*>(ob + j) = __floats2bfloat162_rn(g[j], g[j + 1]);
      }
    } else {
      float w[N];
#pragma unroll
      for (int j = 0; j < N; j += 4) {
        const float4 t = *reinterpret_cast<const float4*>(o.master + e + j);
        w[j] = t.x; w[j + 1] = t.y; w[j + 2] = t.z; w[j + 3] = t.w;
      }
      if (o.opt == 1) {
        // torch.optim.AdamW's update with the element-independent factors folded once per thread and approximate
        // sqrt / rcp (1-2 ulp): the IEEE form cost ~180 instructions per element and the comm CTAs share their SMs
        float m[N], vv[N];
#pragma unroll
        for (int j = 0; j < N; j += 4) {
          const float4 a = *reinterpret_cast<const float4*>(o.m + e + j);
          const float4 b = *reinterpret_cast<const float4*>(o.v + e + j);
          m[j] = a.x; m[j + 1] = a.y; m[j + 2] = a.z; m[j + 3] = a.w;
          vv[j] = b.x; vv[j + 1] = b.y; vv[j + 2] = b.z; vv[j + 3] = b.w;
        }
        const float omb1 = 1.f - o.beta1, omb2 = 1.f - o.beta2, decay = 1.f - o.lr * o.wd;
        const float step = o.lr / o.bc1, rs_bc2 = rsqrtf(o.bc2);
#pragma unroll
        for (int j = 0; j < N; ++j) {
          m[j] = fmaf(o.beta1, m[j], omb1 * g[j]);
          vv[j] = fmaf(o.beta2, vv[j], omb2 * g[j] * g[j]);
          float sq, rc;
          asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(sq) : "f"(vv[j]));
          asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(rc) : "f"(fmaf(sq, rs_bc2, o.eps)));
          w[j] = fmaf(w[j], decay, -step * m[j] * rc);
        }
#pragma unroll
        for (int j = 0; j < N; j += 4) {
          *reinterpret_cast<float4*>(o.m + e + j) = make_float4(m[j], m[j + 1], m[j + 2], m[j + 3]);
          *reinterpret_cast<float4*>(o.v + e + j) = make_float4(vv[j], vv[j + 1], vv[j + 2], vv[j + 3]);
        }
      } else {
#pragma unroll
        for (int j = 0; j < N; ++j) {
          float d = g[j] + o.wd * w[j];
          if (o.m != nullptr) {
            d = o.momentum * o.m[e + j] + d;
            o.m[e + j] = d;
          }
          w[j] -= o.lr * d;
        }
      }
#pragma unroll
      for (int j = 0; j < N; j += 4)
        *reinterpret_cast<float4*>(o.master + e + j) = make_float4(w[j], w[j + 1], w[j + 2], w[j + 3]);
#pragma unroll
      for (int j = 0; j < N; j += 2)
        *reinterpret_cast<__nv_bfloat162*>(o.shard + e + j) = __floats2bfloat162_rn(w[j], w[j + 1]);
    }
  }
  if (!rank_barrier(x, seq)) return;   // peers may reuse their gradient buffers only after every owner has read them
  store_seq(x, seq);
}

// ---- host dispatch --------------------------------------------------------------
static int pick_grid(long long nvec, int max_ctas, int dflt) {
  int cap = max_ctas > 0 ? max_ctas : dflt;
  if (cap > MAX_BLOCKS) cap = MAX_BLOCKS;
  long long need = (nvec + THREADS - 1) / THREADS;
  if (need < 1) need = 1;
  return static_cast<int>(need < cap ? need : cap);
}

// DTB_COMM_WIDE=0 falls back to the 16-byte kernels
static bool wide_enabled() {
  static const bool on = [] { const char* e = getenv("DTB_COMM_WIDE"); return !(e && e[0] == '0'); }();
  return on;
}

// DTB_COMM_NVLS: unset = the multimem kernels carry the bandwidth-bound collectives when the heap has a multicast
// mapping and world >= 4 (with 2 ranks the switch path moves 1.5x the buffer per direction, peer loads only 1x);
// 0 = never; 1 = whenever a multicast mapping exists (also at world 2).
static bool use_nvls(const dtb_comm* c) {
  static const int mode = [] { const char* e = getenv("DTB_COMM_NVLS"); return e ? atoi(e) : -1; }();
  if (c->mc_base == nullptr || c->world < 2 || mode == 0) return false;
  return mode == 1 || c->world >= 4;
}

static int check(const dtb_comm* c, int channel) {
  DTB_REQUIRE(c != nullptr && c->world >= 1 && c->world <= DTB_MAX_RANKS && c->rank >= 0 && c->rank < c->world,
              "dtb_comm: bad communicator");
  DTB_REQUIRE(channel >= 0 && channel < NUM_CHANNELS, "dtb_comm: channel %d out of range", channel);
  return 0;
}

#define DISPATCH_W(W_, ...)                 \
  switch (W_) {                             \
    case 1: { constexpr int W = 1; __VA_ARGS__; break; } \
    case 2: { constexpr int W = 2; __VA_ARGS__; break; } \
    case 4: { constexpr int W = 4; __VA_ARGS__; break; } \
    case 8: { constexpr int W = 8; __VA_ARGS__; break; } \
    default: dtb::set_error("dtb_comm: world size %d not supported (1, 2, 4, 8)", W_); return 2; \
  }

}  // namespace comm
}  // namespace dtb

using namespace dtb;
using namespace dtb::comm;

// The channel rides in the upper bits of max_ctas (bits 16..19) so the public signatures stay small.
static inline int channel_of(int max_ctas) { return (max_ctas >> 16) & 0xF; }
static inline int ctas_of(int max_ctas) { return max_ctas & 0xFFFF; }
static inline bool exclusive_of(int max_ctas) { return ((max_ctas >> 20) & 1) != 0; }
constexpr int EXCLUSIVE_THREADS = 1024;

// Exclusive-mode launch of the 256-bit two-shot kernel: `ctas` whole-SM CTAs in clusters of two.
template <int W, bool SGD>
static cudaError_t launch_exclusive(const Ctx& x, size_t offset, long long nvec32, const SgdArgs& sgd, int ctas,
                                    cudaStream_t s) {
  if (ctas < 2) ctas = 2;
  if (ctas > MAX_BLOCKS) ctas = MAX_BLOCKS;
  ctas &= ~1;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(static_cast<unsigned>(ctas));
  cfg.blockDim = dim3(EXCLUSIVE_THREADS);
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, allreduce_twoshot256_kernel<W, SGD, EXCLUSIVE_THREADS>, x, offset, nvec32, sgd);
}

template <bool FP32, bool SGD>
static cudaError_t launch_exclusive_nvls(const Ctx& x, size_t offset, long long nvec, const SgdArgs& sgd, int ctas,
                                         cudaStream_t s) {
  if (ctas < 2) ctas = 2;
  if (ctas > MAX_BLOCKS) ctas = MAX_BLOCKS;
  ctas &= ~1;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(static_cast<unsigned>(ctas));
  cfg.blockDim = dim3(EXCLUSIVE_THREADS);
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, allreduce_nvls_kernel<FP32, SGD, EXCLUSIVE_THREADS>, x, offset, nvec, sgd);
}

extern "C" int dtb_comm_barrier(const dtb_comm* c, int channel, void* stream) {
  if (check(c, channel)) return 2;
  Ctx x = {*c, channel};
  barrier_kernel<<<1, THREADS, 0, static_cast<cudaStream_t>(stream)>>>(x);
  DTB_LAUNCH_OK();
  return 0;
}

extern "C" int dtb_comm_broadcast(const dtb_comm* c, size_t offset, size_t bytes, int root, int max_ctas,
                                  void* stream) {
  if (check(c, channel_of(max_ctas))) return 2;
  DTB_REQUIRE(offset % 16 == 0 && bytes % 16 == 0, "dtb_comm_broadcast: offset/bytes must be multiples of 16");
  if (bytes == 0) return 0;
  Ctx x = {*c, channel_of(max_ctas)};
  const long long nvec = static_cast<long long>(bytes / 16);
  const int grid = pick_grid(nvec, ctas_of(max_ctas), 64);
  if (use_nvls(c)) broadcast_nvls_kernel<<<grid, THREADS, 0, static_cast<cudaStream_t>(stream)>>>(x, offset, nvec, root);
  else broadcast_kernel<<<grid, THREADS, 0, static_cast<cudaStream_t>(stream)>>>(x, offset, nvec, root);
  DTB_LAUNCH_OK();
  return 0;
}

extern "C" int dtb_comm_all_reduce_avg(const dtb_comm* c, size_t offset, long long n, int dtype, int algo, int max_ctas,
                                       void* stream) {
  if (check(c, channel_of(max_ctas))) return 2;
  const int esize = dtype == 1 ? 4 : 2;
  DTB_REQUIRE(offset % 16 == 0 && (n * esize) % 16 == 0, "dtb_comm_all_reduce_avg: buffer must be 16-byte granular");
  DTB_REQUIRE(algo != 3 || (c->mc_base != nullptr && c->world >= 2),
              "dtb_comm_all_reduce_avg: algo 3 (NVLS) needs a multicast mapping of the heap (dtb_comm.mc_base) and >= 2 ranks");
  if (n == 0) return 0;
  Ctx x = {*c, channel_of(max_ctas)};
  const long long nvec = n * esize / 16;
  auto s = static_cast<cudaStream_t>(stream);
  constexpr int PT = 4;
  const long long oneshot_cap = 64LL * 1024 / 16;   // <= 64 KiB
  if (algo == 0) algo = (nvec <= oneshot_cap) ? 1 : (use_nvls(c) ? 3 : 2);
  if (algo == 3) {
    const int grid = pick_grid(nvec / c->world / 4 + 1, ctas_of(max_ctas), 64);
    const SgdArgs none = {0, 0.f, 1.f};
    if (exclusive_of(max_ctas)) {
      if (dtype == 1) DTB_CUDA_OK((launch_exclusive_nvls<true, false>(x, offset, nvec, none, ctas_of(max_ctas), s)));
      else DTB_CUDA_OK((launch_exclusive_nvls<false, false>(x, offset, nvec, none, ctas_of(max_ctas), s)));
    } else if (dtype == 1) {
      allreduce_nvls_kernel<true, false, THREADS><<<grid, THREADS, 0, s>>>(x, offset, nvec, none);
    } else {
      allreduce_nvls_kernel<false, false, THREADS><<<grid, THREADS, 0, s>>>(x, offset, nvec, none);
    }
  } else if (algo == 1) {
    const long long per_block = static_cast<long long>(THREADS) * PT;
    const long long blocks = (nvec + per_block - 1) / per_block;
    DTB_REQUIRE(blocks <= MAX_BLOCKS, "dtb_comm_all_reduce_avg: one-shot limited to %lld bytes",
                per_block * MAX_BLOCKS * 16);
    DISPATCH_W(c->world, {
      if (dtype == 1) allreduce_oneshot_kernel<true, W, PT><<<static_cast<int>(blocks), THREADS, 0, s>>>(x, offset, nvec);
      else allreduce_oneshot_kernel<false, W, PT><<<static_cast<int>(blocks), THREADS, 0, s>>>(x, offset, nvec);
    });
  } else {
    const int grid = pick_grid(nvec / c->world + 1, ctas_of(max_ctas), 64);
    const SgdArgs none = {0, 0.f, 1.f};
    const bool wide = dtype == 0 && offset % 32 == 0 && (n * esize) % 32 == 0 && wide_enabled();
    DISPATCH_W(c->world, {
      if (dtype == 1) allreduce_twoshot_kernel<true, W, false><<<grid, THREADS, 0, s>>>(x, offset, nvec, none);
      else if (wide && exclusive_of(max_ctas)) DTB_CUDA_OK((launch_exclusive<W, false>(x, offset, nvec / 2, none, ctas_of(max_ctas), s)));
      else if (wide) allreduce_twoshot256_kernel<W, false, THREADS><<<grid, THREADS, 0, s>>>(x, offset, nvec / 2, none);
      else allreduce_twoshot_kernel<false, W, false><<<grid, THREADS, 0, s>>>(x, offset, nvec, none);
    });
  }
  DTB_LAUNCH_OK();
  return 0;
}

extern "C" int dtb_comm_all_reduce_sgd(const dtb_comm* c, size_t grad_offset, size_t param_offset, long long n,
                                       float lr, float weight_decay, int max_ctas, void* stream) {
  if (check(c, channel_of(max_ctas))) return 2;
  DTB_REQUIRE(grad_offset % 16 == 0 && param_offset % 16 == 0 && (n * 2) % 16 == 0,
              "dtb_comm_all_reduce_sgd: buffers must be 16-byte granular");
  if (n == 0) return 0;
  Ctx x = {*c, channel_of(max_ctas)};
  const long long nvec = n * 2 / 16;
  const int grid = pick_grid(nvec / c->world + 1, ctas_of(max_ctas), 64);
  const SgdArgs sgd = {param_offset, lr, 1.f - lr * weight_decay};
  auto s = static_cast<cudaStream_t>(stream);
  const bool wide = grad_offset % 32 == 0 && param_offset % 32 == 0 && (n * 2) % 32 == 0 && wide_enabled();
  if (use_nvls(c)) {
    if (exclusive_of(max_ctas)) DTB_CUDA_OK((launch_exclusive_nvls<false, true>(x, grad_offset, nvec, sgd, ctas_of(max_ctas), s)));
    else allreduce_nvls_kernel<false, true, THREADS><<<pick_grid(nvec / c->world / 4 + 1, ctas_of(max_ctas), 64), THREADS, 0, s>>>(x, grad_offset, nvec, sgd);
    DTB_LAUNCH_OK();
    return 0;
  }
  DISPATCH_W(c->world, {
    if (wide && exclusive_of(max_ctas)) DTB_CUDA_OK((launch_exclusive<W, true>(x, grad_offset, nvec / 2, sgd, ctas_of(max_ctas), s)));
    else if (wide) allreduce_twoshot256_kernel<W, true, THREADS><<<grid, THREADS, 0, s>>>(x, grad_offset, nvec / 2, sgd);
    else allreduce_twoshot_kernel<false, W, true><<<grid, THREADS, 0, s>>>(x, grad_offset, nvec, sgd);
  });
  DTB_LAUNCH_OK();
  return 0;
}

extern "C" int dtb_comm_all_gather(const dtb_comm* c, size_t dst_offset, size_t src_offset, long long n, int esize,
                                   int max_ctas, void* stream) {
  if (check(c, channel_of(max_ctas))) return 2;
  DTB_REQUIRE(dst_offset % 16 == 0 && src_offset % 16 == 0 && (n * esize) % 16 == 0,
              "dtb_comm_all_gather: shard must be 16-byte granular");
  if (n == 0) return 0;
  Ctx x = {*c, channel_of(max_ctas)};
  const long long nvec = n * esize / 16;
  const int grid = pick_grid(nvec, ctas_of(max_ctas), 64);
  if (use_nvls(c)) allgather_nvls_kernel<<<grid, THREADS, 0, static_cast<cudaStream_t>(stream)>>>(x, dst_offset, src_offset, nvec);
  else allgather_kernel<<<grid, THREADS, 0, static_cast<cudaStream_t>(stream)>>>(x, dst_offset, src_offset, nvec);
  DTB_LAUNCH_OK();
  return 0;
}

extern "C" int dtb_comm_reduce_scatter_avg(const dtb_comm* c, void* out, int out_fp32, size_t src_offset, long long n,
                                           int src_dtype, int max_ctas, void* stream) {
  if (check(c, channel_of(max_ctas))) return 2;
  const int esize = src_dtype == 1 ? 4 : 2;
  DTB_REQUIRE(src_offset % 16 == 0 && (n * esize) % 16 == 0, "dtb_comm_reduce_scatter_avg: shard must be 16-byte granular");
  if (n == 0) return 0;
  Ctx x = {*c, channel_of(max_ctas)};
  const long long nvec = n * esize / 16;
  const int grid = pick_grid(nvec, ctas_of(max_ctas), 64);
  auto s = static_cast<cudaStream_t>(stream);
  OptArgs o = {};
  const bool nvls = use_nvls(c);
  DISPATCH_W(c->world, {
    if (src_dtype == 1 && nvls) reduce_scatter_kernel<true, W, 0, true><<<grid, THREADS, 0, s>>>(x, src_offset, nvec, out, out_fp32, o);
    else if (src_dtype == 1) reduce_scatter_kernel<true, W, 0><<<grid, THREADS, 0, s>>>(x, src_offset, nvec, out, out_fp32, o);
    else if (nvls) reduce_scatter_kernel<false, W, 0, true><<<grid, THREADS, 0, s>>>(x, src_offset, nvec, out, out_fp32, o);
    else reduce_scatter_kernel<false, W, 0><<<grid, THREADS, 0, s>>>(x, src_offset, nvec, out, out_fp32, o);
  });
  DTB_LAUNCH_OK();
  return 0;
}

extern "C" int dtb_comm_rs_optim(const dtb_comm* c, size_t grad_offset, size_t shard_offset, float* master, float* m,
                                 float* v, long long n, int opt, float lr, float beta1, float beta2, float eps,
                                 float weight_decay, float bias_corr1, float bias_corr2, float momentum, int max_ctas,
                                 void* stream) {
  if (check(c, channel_of(max_ctas))) return 2;
  DTB_REQUIRE(grad_offset % 16 == 0 && shard_offset % 16 == 0 && n % 8 == 0, "dtb_comm_rs_optim: shard must be 16-byte granular");
  DTB_REQUIRE(master != nullptr, "dtb_comm_rs_optim: fp32 master shard is required");
  DTB_REQUIRE(opt == 0 || (m != nullptr && v != nullptr), "dtb_comm_rs_optim: AdamW needs both moments");
  DTB_REQUIRE(((reinterpret_cast<uintptr_t>(master) | reinterpret_cast<uintptr_t>(m) | reinterpret_cast<uintptr_t>(v)) & 15) == 0,
              "dtb_comm_rs_optim: master and moment shards must be 16-byte aligned");
  if (n == 0) return 0;
  Ctx x = {*c, channel_of(max_ctas)};
  const long long nvec = n * 2 / 16;
  const int grid = pick_grid(nvec, ctas_of(max_ctas), 64);
  OptArgs o = {master, m, v,
               reinterpret_cast<__nv_bfloat16*>(static_cast<uint8_t*>(c->peers[c->rank]) + shard_offset),
               opt, lr, beta1, beta2, eps, weight_decay, bias_corr1, bias_corr2, momentum};
  auto s = static_cast<cudaStream_t>(stream);
  const bool nvls = use_nvls(c);
  DISPATCH_W(c->world, {
    if (nvls) reduce_scatter_kernel<false, W, 1, true><<<grid, THREADS, 0, s>>>(x, grad_offset, nvec, nullptr, 0, o);
    else reduce_scatter_kernel<false, W, 1><<<grid, THREADS, 0, s>>>(x, grad_offset, nvec, nullptr, 0, o);
  });
  DTB_LAUNCH_OK();
  return 0;
}

// ---- copy-engine two-shot ----------------------------------------------------------------------------------------
// The NVLink legs of a large all-reduce are plain copies; issued from SM threads they share an SM's L1/shared-memory
// datapath with the resident GEMM CTA's TMA traffic and cost the GEMM 20-35 % (bench/interference.py), while the same
// bytes moved by the copy engines cost it < 8 %.  So for big buckets: barrier kernel -> DMA the W-1 peer copies of my
// slice into a local staging buffer -> ONE local kernel (sum in rank order, 1/W, cast, optional SGD) -> DMA the result
// slice to every peer -> barrier kernel.  Only the short local kernel competes with the GEMM for SM resources.
struct SliceArgs {
  const uint8_t* own;       // my own copy of the slice (gradients)
  const uint8_t* staged;    // W-1 peer copies, in rank order with my rank skipped
  uint8_t* dst;             // mean gradient (in place) or, with sgd, the parameter slice
  int rank;
  int sgd;
  float lr, decay;
};

template <int W>
__global__ void __launch_bounds__(THREADS, MIN_CTAS_PER_SM)
slice_reduce_kernel(const SliceArgs a, long long nvec32, long long slice_bytes) {
  const float scale = 1.f / W;
  const long long stride = static_cast<long long>(gridDim.x) * THREADS;
  for (long long v = static_cast<long long>(blockIdx.x) * THREADS + threadIdx.x; v < nvec32; v += stride) {
    const size_t off = static_cast<size_t>(v) * 32;
    float acc[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) acc[j] = 0.f;
#pragma unroll
    for (int r = 0; r < W; ++r) {       // fixed rank order: identical rounding no matter which rank owns the slice
      const uint8_t* src = (r == a.rank) ? a.own + off : a.staged + static_cast<size_t>(r < a.rank ? r : r - 1) * slice_bytes + off;
      const V8 raw = ld256_stream(src);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&raw.w[j]));
        acc[2 * j] += f.x;
        acc[2 * j + 1] += f.y;
      }
    }
    if (a.sgd) {
      const V8 praw = ld256_stream(a.dst + off);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&praw.w[j]));
        acc[2 * j] = f.x * a.decay - a.lr * (acc[2 * j] * scale);
        acc[2 * j + 1] = f.y * a.decay - a.lr * (acc[2 * j + 1] * scale);
      }
    } else {
#pragma unroll
      for (int j = 0; j < 16; ++j) acc[j] *= scale;
    }
    V8 out;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      __nv_bfloat162 h = __floats2bfloat162_rn(acc[2 * j], acc[2 * j + 1]);
      out.w[j] = *reinterpret_cast<uint32_t*>(&h);
    }
    st256_stream(a.dst + off, out);
  }
}

// Single-GPU stand-in for allreduce_twoshot256_kernel<2, true>: same instruction mix and bytes per element (two gradient
// reads, one parameter read, two parameter writes), all on local memory and without flags.  bench/interference.py runs
// it beside a GEMM to separate SM/L2/HBM interference from NVLink effects.
__global__ void __launch_bounds__(THREADS, MIN_CTAS_PER_SM)
stream_probe_kernel(const uint8_t* a, const uint8_t* b, uint8_t* p, uint8_t* q, long long nvec32, float lr) {
  const long long stride = static_cast<long long>(gridDim.x) * THREADS;
  for (long long v0 = static_cast<long long>(blockIdx.x) * THREADS + threadIdx.x; v0 < nvec32; v0 += stride * 2) {
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const long long v = v0 + u * stride;
      if (v >= nvec32) break;
      const size_t off = static_cast<size_t>(v) * 32;
      const V8 ra = ld256_stream(a + off), rb = ld256_stream(b + off), rp = ld256_stream(p + off);
      V8 out;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float2 fa = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&ra.w[j]));
        const float2 fb = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&rb.w[j]));
        const float2 fp = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&rp.w[j]));
        __nv_bfloat162 h = __floats2bfloat162_rn(fp.x - lr * 0.5f * (fa.x + fb.x), fp.y - lr * 0.5f * (fa.y + fb.y));
        out.w[j] = *reinterpret_cast<uint32_t*>(&h);
      }
      st256_stream(p + off, out);
      st256_stream(q + off, out);
    }
  }
}

extern "C" int dtb_comm_all_reduce_ce(const dtb_comm* c, size_t grad_offset, size_t param_offset, int sgd, long long n,
                                      void* staging, size_t staging_bytes, float lr, float weight_decay, int max_ctas,
                                      void* stream) {
  if (check(c, channel_of(max_ctas))) return 2;
  const int W = c->world;
  const long long total = n * 2;
  DTB_REQUIRE(W >= 2, "dtb_comm_all_reduce_ce: needs at least two ranks");
  DTB_REQUIRE(n > 0 && total % (static_cast<long long>(W) * 32) == 0 && grad_offset % 32 == 0 && param_offset % 32 == 0,
              "dtb_comm_all_reduce_ce: the buffer must split into 32-byte-granular slices per rank");
  const long long slice = total / W;
  DTB_REQUIRE(staging != nullptr && staging_bytes >= static_cast<size_t>(slice) * (W - 1) &&
                  (reinterpret_cast<uintptr_t>(staging) & 31) == 0,
              "dtb_comm_all_reduce_ce: staging must hold world-1 slices (%lld bytes)", slice * (W - 1));
  Ctx x = {*c, channel_of(max_ctas)};
  auto s = static_cast<cudaStream_t>(stream);
  const int r = c->rank;
  uint8_t* mine = static_cast<uint8_t*>(c->peers[r]);
  // every rank's bucket is complete (each rank's comm stream waited for its own producer before this point)
  barrier_kernel<<<1, THREADS, 0, s>>>(x);
  DTB_LAUNCH_OK();
  int j = 0;
  for (int p = 0; p < W; ++p) {
    if (p == r) continue;
    DTB_CUDA_OK(cudaMemcpyAsync(static_cast<uint8_t*>(staging) + static_cast<size_t>(j) * slice,
                                static_cast<const uint8_t*>(c->peers[p]) + grad_offset + static_cast<size_t>(r) * slice,
                                static_cast<size_t>(slice), cudaMemcpyDefault, s));
    ++j;
  }
  SliceArgs a = {};
  a.own = mine + grad_offset + static_cast<size_t>(r) * slice;
  a.staged = static_cast<const uint8_t*>(staging);
  const size_t dst_off = (sgd ? param_offset : grad_offset) + static_cast<size_t>(r) * slice;
  a.dst = mine + dst_off;
  a.rank = r;
  a.sgd = sgd;
  a.lr = lr;
  a.decay = 1.f - lr * weight_decay;
  const int grid = pick_grid(slice / 32, ctas_of(max_ctas), 128);
  DISPATCH_W(W, { slice_reduce_kernel<W><<<grid, THREADS, 0, s>>>(a, slice / 32, slice); });
  DTB_LAUNCH_OK();
  for (int p = 0; p < W; ++p) {
    if (p == r) continue;
    DTB_CUDA_OK(cudaMemcpyAsync(static_cast<uint8_t*>(c->peers[p]) + dst_off, mine + dst_off, static_cast<size_t>(slice),
                                cudaMemcpyDefault, s));
  }
  // once every peer has passed this barrier, all their pushes into my buffer have completed
  barrier_kernel<<<1, THREADS, 0, s>>>(x);
  DTB_LAUNCH_OK();
  return 0;
}

extern "C" int dtb_debug_stream_probe(const void* a, const void* b, void* p, void* q, long long bytes, int ctas,
                                      void* stream) {
  DTB_REQUIRE(bytes % 32 == 0 && ctas > 0 && ctas <= 1024, "dtb_debug_stream_probe: bytes must be a multiple of 32");
  stream_probe_kernel<<<ctas, THREADS, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const uint8_t*>(a), static_cast<const uint8_t*>(b), static_cast<uint8_t*>(p), static_cast<uint8_t*>(q),
      bytes / 32, 0.f);
  DTB_LAUNCH_OK();
  return 0;
}

extern "C" int dtb_comm_error_flag(const dtb_comm* c) {
  if (c == nullptr) return -1;
  uint32_t v = 0;
  if (cudaMemcpy(&v, c->signals[c->rank] + ERR_SLOT, sizeof(v), cudaMemcpyDeviceToHost) != cudaSuccess) return -1;
  return static_cast<int>(v);
}
