/*
 * dtb200 — C ABI of the B200 (sm_100a) kernel library behind
 * distributed_training_b200.  Every entry point:
 *   - takes raw device pointers and a cudaStream_t passed as void*;
 *   - is asynchronous with respect to the host unless stated otherwise;
 *   - returns 0 on success, non-zero on failure (dtb_last_error() has the text);
 *   - allocates no device memory unless its comment says so.
 *
 * Where a function replaces something the reference executes implicitly through
 * PyTorch, the comment cites the reference call site (SURVEY.md §2.5 numbering).
 */
#ifndef DTB200_H_
#define DTB200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- library ------------------------------------------------------------ */
const char* dtb_last_error(void);
int dtb_version(void);
/* Kernel launches issued by this library since load / the last reset (bench.py's gpu_launches). */
long long dtb_launch_count(void);
void dtb_reset_launch_count(void);

/* ---- GEMM (K9/K10: nn.Linear forward, dgrad, wgrad → cuBLAS in the reference,
 *      /root/reference/src/distributed_trainer.py:162,164) ----------------------
 * D[m,n] = epilogue( alpha * sum_k A(m,k) * B(n,k) )   bf16 inputs, fp32 accumulation in TMEM.
 *   a_mn_major = 0: A stored [M,K] row-major (pitch lda);  1: stored [K,M] row-major.
 *   b_mn_major = 0: B stored [N,K] row-major (pitch ldb);  1: stored [K,N] row-major.
 * Linear forward  Y = X W^T           : A = X  (K-major), B = W  (K-major)
 * Linear dgrad    dX = dY W           : A = dY (K-major), B = W  (MN-major)
 * Linear wgrad    dW = dY^T X         : A = dY (MN-major), B = X (MN-major)
 * Epilogue, in order:  v = alpha*acc; v += bias[n]; act==1: (d2 = v if d2) v = gelu_tanh(v);
 *   act==2: v *= gelu_tanh'(aux[m,n]);  otherwise if aux: v += beta*aux[m,n];  D = v.
 * Batched over batch_outer x batch_inner problems with independent element strides
 * (attention heads: inner = head, outer = batch).  All pitches/strides in elements;
 * pitches must be multiples of 8 elements, base pointers 16-byte aligned. */
typedef struct dtb_gemm_args {
  const void* a; const void* b; void* d;
  void* d2;              /* optional second output (pre-activation), same layout as d, bf16 */
  const void* aux;       /* optional [M,N] operand (residual / pre-activation / accumulate-into) */
  const void* bias;      /* optional [N] */
  long long lda, ldb, ldd, ld_aux;
  long long a_stride_inner, a_stride_outer;
  long long b_stride_inner, b_stride_outer;
  long long d_stride_inner, d_stride_outer;
  long long aux_stride_inner, aux_stride_outer;
  int m, n, k;
  int batch_inner, batch_outer;
  int a_mn_major, b_mn_major;
  int d_fp32, aux_fp32, bias_fp32;
  int act;               /* 0 none, 1 gelu(tanh), 2 multiply by gelu'(aux) */
  float alpha, beta;
  int tile_n;            /* 0 = auto, else 128 or 256 */
  int cta_pair;          /* 0 = auto, 1 = force 1-CTA MMA, 2 = force cta_group::2 */
  int max_ctas;          /* 0 = all SMs; otherwise cap the persistent grid (leave SMs to a comm kernel) */
  unsigned dbg_lbo_a, dbg_sbo_a, dbg_lbo_b, dbg_sbo_b;   /* 0 = default; bring-up overrides */
  /* split-K (problems with few output tiles and a long K, e.g. transformer wgrads): fp32 partial sums of the
   * K slices go through `workspace` and a second kernel adds them in slice order and applies the epilogue.
   * workspace == NULL or split_k == 1 disables it; split_k == 0 lets the library choose (cost model);
   * split_k >= 2 forces that many slices.  Needs split * m * n * 4 bytes. */
  void* workspace;
  long long workspace_bytes;
  int split_k;
  /* optional fp32 [N], 16-byte aligned: the column sums of D (as rounded to bf16) are ADDED to it from the epilogue
   * with vector reductions -- the bias gradient when D is an output gradient (replaces a dtb_colsum_bf16 pass over
   * D).  The caller zeroes it (dtb_take_f32_to_bf16 re-zeroes while converting).  Needs bf16 D, no batching; turns
   * split-K and the direct epilogue off.  Summation order is not fixed (fp32 atomics). */
  float* colsum;
} dtb_gemm_args;
int dtb_gemm_bf16(const dtb_gemm_args* args, void* stream);

/* All-gather fused into the GEMM that consumes the gathered weights (K6+K9: FSDP's all_gather_into_tensor
 * followed by the Linear forward in the reference, /root/reference/src/dist_strategy/fsdp_strategy.py:20).
 * B (= args->b, the local FULL weight matrix [N, K], K-major) is row-sharded over `world` ranks; rank r's
 * shard [rows_per_rank, K] (same pitch ldb) is reachable at peer_shard[r] (symmetric heap, peer-mapped).
 * Inside ONE kernel, extra warps of every CTA pull 16-row pieces of all shards over NVLink into B in the
 * order the tiles need them (own rows first) and the TMA producer of a tile waits only for its own n-block,
 * so the transfer overlaps the MMAs tile by tile and every byte crosses NVLink once.  After the call B holds
 * the complete matrix (it is reused by the backward dgrad).  `flags`: dtb_gemm_allgather_flag_ints(N) ints of
 * scratch, zeroed by the call.  A must be K-major; no batching.  The caller orders the call after the peers'
 * shard updates (the reduce-scatter+optimizer kernel ends with a rank barrier). */
#define DTB_MAX_RANKS 8
typedef struct dtb_gather_args {
  const void* peer_shard[DTB_MAX_RANKS];
  int* flags;
  int world, rank, rows_per_rank;
} dtb_gather_args;
size_t dtb_gemm_allgather_flag_ints(int n);
int dtb_gemm_bf16_allgather(const dtb_gemm_args* args, const dtb_gather_args* gather, void* stream);

/* Tiny-shape Linear for the reference's own model sizes (K=20,N=1; K=10,N=1): fp32 SIMT kernels. */
int dtb_linear_small_fwd(const float* x, const float* w, const float* b, float* y, int rows, int in_f, int out_f,
                         void* stream);
int dtb_linear_small_bwd(const float* x, const float* w, const float* dy, float* dx, float* dw, float* db, int rows,
                         int in_f, int out_f, void* stream);

/* ---- elementwise / reductions -------------------------------------------- */
int dtb_gelu_fwd_bf16(const void* x, void* y, long long n, void* stream);
int dtb_gelu_bwd_bf16(const void* x, const void* dy, void* dx, long long n, void* stream);
/* out[n] = sum_m x[m,n]  (bias gradient). bf16 in, out bf16 or fp32. Uses no scratch. */
int dtb_colsum_bf16(const void* x, void* out, int rows, int cols, long long ldx, int out_fp32, void* stream);
/* Same result with the rows split over up to DTB_COLSUM_MAX_CHUNKS CTAs per column block (narrow, tall
 * inputs such as [8192, 768] would otherwise occupy 12 CTAs); partial sums go through `scratch`
 * (dtb_colsum_split_scratch_floats floats) and are added in a fixed order (deterministic). */
#define DTB_COLSUM_MAX_CHUNKS 64
size_t dtb_colsum_split_scratch_floats(int rows, int cols);
int dtb_colsum_bf16_split(const void* x, void* out, int rows, int cols, long long ldx, int out_fp32, float* scratch,
                          void* stream);
int dtb_add_bf16(const void* a, const void* b, void* out, long long n, void* stream);
int dtb_cast_f32_to_bf16(const float* in, void* out, long long n, void* stream);
int dtb_cast_bf16_to_f32(const void* in, float* out, long long n, void* stream);
/* out[i] = bf16(in[i]); in[i] = 0: hands an fp32 accumulator (dtb_gemm_args.colsum) over to a bf16 gradient and
 * leaves it ready for the next step. */
int dtb_take_f32_to_bf16(float* in, void* out, long long n, void* stream);
/* sum of squares of n elements into *out (fp32, accumulated with atomicAdd; caller zeroes). K15. */
int dtb_sumsq(const void* x, int is_fp32, long long n, float* out, void* stream);

/* Input pipeline (K14: the H2D copy behind DataLoader(pin_memory=True) + .to(device),
 * /root/reference/src/distributed_trainer.py:176-183): dst[i,:] = src[idx[i] (% src_rows if src_rows > 0), :] for
 * `rows` rows of `row_bytes` bytes (multiple of 4).  `src` may be pinned HOST memory: the GPU then pulls exactly the
 * sampled rows over PCIe and no host-side staging copy exists.  idx and dst are device memory.  max_ctas: 0 = 16. */
int dtb_gather_rows(const void* src, const long long* idx, void* dst, int rows, long long row_bytes,
                    long long src_rows, int max_ctas, void* stream);

/* ---- LayerNorm (K16) ------------------------------------------------------- */
/* y = LN(x [+ residual]) * gamma + beta; when residual != NULL, xres_out receives x + residual. */
int dtb_layernorm_fwd(const void* x, const void* residual, const void* gamma, const void* beta, void* y,
                      void* xres_out, float* mean, float* rstd, int rows, int cols, float eps, void* stream);
/* dx = LN backward (+ dres_in if given).  dgamma/dbeta are fp32 [cols]; partial sums go through
 * `scratch` (fp32, at least dtb_layernorm_bwd_scratch_floats(rows, cols) floats). */
size_t dtb_layernorm_bwd_scratch_floats(int rows, int cols);
int dtb_layernorm_bwd(const void* dy, const void* x, const void* gamma, const float* mean, const float* rstd,
                      const void* dres_in, void* dx, float* dgamma, float* dbeta, float* scratch, int rows, int cols,
                      int accumulate_param_grads, void* stream);

/* dtb_layernorm_bwd with dgamma / dbeta written (or accumulated) as bf16, e.g. straight into bf16 gradient buckets. */
int dtb_layernorm_bwd_bf16grads(const void* dy, const void* x, const void* gamma, const float* mean, const float* rstd,
                                const void* dres_in, void* dx, void* dgamma, void* dbeta, float* scratch, int rows,
                                int cols, int accumulate_param_grads, void* stream);
/* Same computation as dtb_layernorm_bwd, for callers that cannot hand over a scratch buffer because they
 * cannot allocate device memory at all: a bare C/Fortran host with no device allocator, a plugin inside a
 * framework that owns the whole GPU and forbids cudaMalloc, or a loop that cannot tolerate cudaMalloc's
 * implicit synchronisation.  The scratch is a 512 KiB __device__ array inside the library (held by every GPU
 * that loads it, until exit), which limits this entry point to cols <= DTB_LN_NOALLOC_MAX_COLS; larger
 * rows are rejected with an error.  No allocating call is made.  The first call on each device resolves the
 * array's device address (cudaGetSymbolAddress) — issue it outside stream capture.  All calls on a GPU share
 * the array: calls whose kernels overlap on the GPU race; nothing needs zeroing between calls.
 * No counterpart in the reference. */
#define DTB_LN_NOALLOC_MAX_COLS 2048
int dtb_layernorm_bwd_noalloc(const void* dy, const void* x, const void* gamma, const float* mean, const float* rstd,
                              const void* dres_in, void* dx, float* dgamma, float* dbeta, int rows, int cols,
                              int accumulate_param_grads, void* stream);

/* ---- losses (K11/K12: F.cross_entropy / nn.MSELoss,
 *      /root/reference/src/distributed_trainer.py:163, playground/ddp_script.py:135) */
/* Fused softmax cross-entropy with index targets: per-row loss (fp32) and, in place over
 * `logits`, dlogits = (softmax - onehot) * grad_scale.  Columns >= vocab (padding) get 0. */
int dtb_cross_entropy_fwd_bwd(void* logits, const long long* targets, float* row_loss, int rows, int vocab,
                              long long ld, float grad_scale, int write_grad, void* stream);
/* Soft-target form the reference's trainer actually exercises (float targets, same shape). */
int dtb_soft_cross_entropy_fwd_bwd(const float* logits, const float* targets, float* row_loss, float* dlogits,
                                   int rows, int cols, float grad_scale, void* stream);
int dtb_mse_fwd_bwd(const float* pred, const float* target, float* loss_sum, float* dpred, long long n,
                    float grad_scale, void* stream);
/* mean of n floats into *out */
int dtb_mean_f32(const float* x, long long n, float* out, void* stream);

/* ---- embedding (K16) -------------------------------------------------------- */
int dtb_embedding_fwd(const long long* tokens, const void* wte, const void* wpe, void* out, int batch, int seq,
                      int channels, void* stream);
/* dwte[token] += dout (bf16x2 atomics), dwpe[t] = sum_b dout[b,t] */
int dtb_embedding_bwd(const void* dout, const long long* tokens, void* dwte, void* dwpe, int batch, int seq,
                      int channels, int accumulate_wte, void* stream);

/* ---- attention (K16) ----------------------------------------------------------
 * Causal self-attention over a packed qkv activation [batch, seq, 3*heads*64] (bf16).
 * Forward writes out [batch, seq, heads*64] and the log-sum-exp per row (fp32 [batch,heads,seq]). */
int dtb_attention_fwd(const void* qkv, void* out, float* lse, int batch, int seq, int heads, int head_dim,
                      void* stream);
/* Backward: dqkv [batch, seq, 3*heads*64].  `scratch` must hold
 * dtb_attention_bwd_scratch_floats(...) floats (fp32 dQ accumulator + row deltas). */
size_t dtb_attention_bwd_scratch_floats(int batch, int seq, int heads, int head_dim);
int dtb_attention_bwd(const void* qkv, const void* out, const void* dout, const float* lse, void* dqkv,
                      float* scratch, int batch, int seq, int heads, int head_dim, void* stream);

/* ---- optimizers (K13: torch.optim.SGD foreach kernels,
 *      /root/reference/src/distributed_trainer.py:165,200) -------------------- */
#define DTB_MT_MAX_TENSORS 48
typedef struct dtb_mt_args {
  void* param[DTB_MT_MAX_TENSORS];       /* bf16 or fp32 */
  const void* grad[DTB_MT_MAX_TENSORS];  /* bf16 or fp32 */
  float* master[DTB_MT_MAX_TENSORS];     /* optional fp32 master weights (NULL entries allowed if master_on=0) */
  float* m[DTB_MT_MAX_TENSORS];          /* momentum / exp_avg (fp32) or NULL */
  float* v[DTB_MT_MAX_TENSORS];          /* exp_avg_sq (fp32) or NULL */
  long long numel[DTB_MT_MAX_TENSORS];
  int count;
  int param_fp32, grad_fp32, master_on;
} dtb_mt_args;
int dtb_multi_tensor_sgd(const dtb_mt_args* args, float lr, float momentum, float weight_decay, void* stream);
int dtb_multi_tensor_adamw(const dtb_mt_args* args, float lr, float beta1, float beta2, float eps, float weight_decay,
                           float bias_corr1, float bias_corr2, void* stream);

/* ---- symmetric-memory collectives (K2-K8: NCCL broadcast / all-reduce / all-gather /
 *      reduce-scatter issued by torch DDP / FSDP in the reference) ----------------
 * A communicator is described by the table of every rank's mapping of the same
 * symmetric allocation.  `peers[r]` is rank r's heap base as mapped in THIS process;
 * `signals[r]` likewise for the signal pad (DTB_SIGNAL_PAD_BYTES, zero-initialised).
 * Operands are given as byte offsets into the heap, identical on all ranks.
 * `max_ctas` arguments: low 16 bits cap the grid (0 = default 64, at most 128 CTAs of 256 threads — the
 * SM budget left to concurrently running GEMMs); bits 16..19 select the barrier channel (0..7).
 * Kernels that may run concurrently on different streams must use different channels; every rank
 * must issue the same sequence of collectives with the same arguments on a given channel. */
#define DTB_SIGNAL_PAD_BYTES (64 * 1024)
typedef struct dtb_comm {
  void* peers[DTB_MAX_RANKS];
  uint32_t* signals[DTB_MAX_RANKS];
  void* mc_base;          /* NVLS multicast mapping of the same segment (dtb_symm_mc_*), or NULL */
  int rank, world;
  long long timeout_cycles; /* spin watchdog in SM clocks (0 = default, ~20 s).  On expiry the waiting CTA records the
                             * missing peer in the communicator's error slot and leaves the kernel without touching the
                             * payload; later collectives on the communicator return at entry; dtb_comm_error_flag()
                             * reports it to the host. */
} dtb_comm;
/* all ranks enter; returns when every rank's prior work on `stream` is visible to peers */
int dtb_comm_barrier(const dtb_comm* c, int channel, void* stream);
/* buf[offset .. +bytes) on every rank = root's */
int dtb_comm_broadcast(const dtb_comm* c, size_t offset, size_t bytes, int root, int max_ctas, void* stream);
/* in-place mean over ranks of n elements at `offset` (dtype: 0 bf16, 1 fp32); algo: 0 auto, 1 one-shot (peer loads),
 * 2 two-shot (peer loads + peer stores, fixed rank order: bit-identical to a sequential fp32 sum), 3 NVLS two-shot
 * (multimem.ld_reduce + multimem.st through the NVSwitch; needs c->mc_base; identical on all ranks, summation order
 * chosen by the switch).  auto = one-shot up to 64 KiB, else NVLS when c->mc_base is mapped and world >= 4
 * (DTB_COMM_NVLS=0/1 overrides), else two-shot.  The same rule selects the multimem variants of broadcast,
 * all-gather, reduce-scatter and rs_optim.  OR 1<<20 into max_ctas to run a large two-shot / NVLS reduction on
 * whole-SM 1024-thread CTAs (the caller must have left that many SMs free). */
int dtb_comm_all_reduce_avg(const dtb_comm* c, size_t offset, long long n, int dtype, int algo, int max_ctas,
                            void* stream);
/* Gradient all-reduce fused with the optimizer step (bf16): rank r averages slice r of the gradients over
 * all ranks, applies p = p*(1 - lr*wd) - lr*g to slice r of the parameters (same layout at param_offset)
 * and stores the NEW PARAMETERS into every rank.  The gradient buffers are left untouched. */
int dtb_comm_all_reduce_sgd(const dtb_comm* c, size_t grad_offset, size_t param_offset, long long n, float lr,
                            float weight_decay, int max_ctas, void* stream);
/* Large-bucket variant of the two calls above that moves the NVLink legs with the copy engines instead of SM
 * threads (a resident GEMM loses < 8 % to DMA traffic but 20-35 % to co-resident load/store CTAs): barrier kernel,
 * W-1 peer->staging copies of this rank's slice, one local reduce(+SGD when sgd != 0, else the mean gradient is
 * written in place) kernel, W-1 slice pushes, barrier kernel -- all enqueued on `stream` by this one call.
 * n*2 bytes must split into 32-byte-granular slices per rank; staging: local scratch of (world-1) * n*2/world bytes. */
int dtb_comm_all_reduce_ce(const dtb_comm* c, size_t grad_offset, size_t param_offset, int sgd, long long n,
                           void* staging, size_t staging_bytes, float lr, float weight_decay, int max_ctas,
                           void* stream);
/* dst[r*n .. (r+1)*n) = rank r's src[0..n)  (element size esize) */
int dtb_comm_all_gather(const dtb_comm* c, size_t dst_offset, size_t src_offset, long long n, int esize,
                        int max_ctas, void* stream);
/* out[0..n) (local, fp32 or bf16) = mean_r(rank r's src[rank*n .. (rank+1)*n)) */
int dtb_comm_reduce_scatter_avg(const dtb_comm* c, void* out, int out_fp32, size_t src_offset, long long n,
                                int src_dtype, int max_ctas, void* stream);
/* fused reduce-scatter + optimizer on the local shard (K7+K13):
 * g = mean_r(grad@r[rank*n..]); update master/moments; write bf16 shard at shard_offset. opt: 0 sgd, 1 adamw */
int dtb_comm_rs_optim(const dtb_comm* c, size_t grad_offset, size_t shard_offset, float* master, float* m,
                      float* v, long long n, int opt, float lr, float beta1, float beta2, float eps,
                      float weight_decay, float bias_corr1, float bias_corr2, float momentum, int max_ctas,
                      void* stream);
/* error flag raised by a spin watchdog (0 = none); host-synchronous */
/* Debug/benchmark only: single-GPU kernel with the memory traffic of a 2-rank fused all-reduce+SGD (3 reads, 2 writes
 * per 32 bytes, no flags); p and q are updated in place from a, b.  Used by bench/interference.py. */
int dtb_debug_stream_probe(const void* a, const void* b, void* p, void* q, long long bytes, int ctas, void* stream);

int dtb_comm_error_flag(const dtb_comm* c);
/* Debug/benchmark only: cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDefault) on `stream` (peer-copy rate measurement). */
int dtb_debug_memcpy_async(void* dst, const void* src, size_t bytes, void* stream);

/* ---- host-side helpers for the symmetric heap --------------------------------- */
/* CUDA IPC back end: cudaMalloc + cudaIpcGetMemHandle; handle_out must hold 64 bytes. Allocates device memory. */
int dtb_symm_alloc(size_t bytes, void** ptr_out, void* handle_out);
int dtb_symm_open(const void* handle, void** ptr_out);
int dtb_symm_close(void* ptr);
int dtb_symm_free(void* ptr);
/* VMM + NVLS back end (what NCCL's NVLS transport sets up for the reference, /root/reference/src/distributed_trainer.py:60-62).
 * All calls act on the CURRENT device.  Bootstrap of one segment (parallel/symm.py):
 *   every rank : granularity -> create (physical memory + POSIX fd) -> send the fd to every peer (SCM_RIGHTS)
 *                -> import each peer's fd -> map own and peers' handles (unicast addresses = dtb_comm.peers[])
 *   rank 0     : mc_create (multicast object + fd) -> send the fd to every peer;  peers: import
 *   every rank : mc_add_device -> [host barrier] -> mc_bind(own memory) -> [host barrier] -> map(mc handle)
 *                = dtb_comm.mc_base.
 * probe: *vmm_ok = the driver supports VMM with fd handles; *multicast_ok = and multicast objects. */
int dtb_symm_vmm_probe(int* vmm_ok, int* multicast_ok);
int dtb_symm_vmm_granularity(int world, int multicast, size_t* gran_out);
/* cuMemCreate (allocates device memory) + export; the caller closes the fd once the peers have imported it */
int dtb_symm_vmm_create(size_t bytes, unsigned long long* handle_out, int* fd_out);
int dtb_symm_vmm_import(int fd, unsigned long long* handle_out);
/* reserve a VA range, map `handle` (memory or multicast) into it, grant this device read/write */
int dtb_symm_vmm_map(unsigned long long handle, size_t bytes, size_t alignment, void** ptr_out);
int dtb_symm_vmm_unmap(void* ptr, size_t bytes);
int dtb_symm_vmm_release(unsigned long long handle);
int dtb_symm_mc_create(size_t bytes, int world, unsigned long long* handle_out, int* fd_out);
int dtb_symm_mc_add_device(unsigned long long mc_handle);
int dtb_symm_mc_bind(unsigned long long mc_handle, unsigned long long mem_handle, size_t bytes);
int dtb_symm_mc_unbind(unsigned long long mc_handle, size_t bytes);

#ifdef __cplusplus
}
#endif
#endif /* DTB200_H_ */
