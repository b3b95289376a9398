"""ctypes bindings for ``libdtb200.so`` (C ABI in ``include/dtb200.h``).

The wrappers take torch tensors, pass raw device pointers plus the *current*
CUDA stream, and never fall back to PyTorch: if the library is missing on a
machine that has a GPU, :func:`lib` raises, so a silent eager fallback cannot
masquerade as the native path.

The reference has no native layer of its own; these entry points stand in for the torch internals it executes
(DDP init broadcast and Reducer, c10d collectives, ATen/cuBLAS GEMMs, loss and optimizer kernels).
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path
from typing import Optional, Sequence, Tuple

import torch

_LIB: Optional[C.CDLL] = None
_LIB_ERR: Optional[str] = None
LIB_PATH = Path(__file__).resolve().parent.parent / "_native" / "libdtb200.so"

MT_MAX = 48
MAX_RANKS = 8
SIGNAL_PAD_BYTES = 64 * 1024


class GemmArgs(C.Structure):
    _fields_ = [
        ("a", C.c_void_p), ("b", C.c_void_p), ("d", C.c_void_p),
        ("d2", C.c_void_p), ("aux", C.c_void_p), ("bias", C.c_void_p),
        ("lda", C.c_longlong), ("ldb", C.c_longlong), ("ldd", C.c_longlong), ("ld_aux", C.c_longlong),
        ("a_stride_inner", C.c_longlong), ("a_stride_outer", C.c_longlong),
        ("b_stride_inner", C.c_longlong), ("b_stride_outer", C.c_longlong),
        ("d_stride_inner", C.c_longlong), ("d_stride_outer", C.c_longlong),
        ("aux_stride_inner", C.c_longlong), ("aux_stride_outer", C.c_longlong),
        ("m", C.c_int), ("n", C.c_int), ("k", C.c_int),
        ("batch_inner", C.c_int), ("batch_outer", C.c_int),
        ("a_mn_major", C.c_int), ("b_mn_major", C.c_int),
        ("d_fp32", C.c_int), ("aux_fp32", C.c_int), ("bias_fp32", C.c_int),
        ("act", C.c_int),
        ("alpha", C.c_float), ("beta", C.c_float),
        ("tile_n", C.c_int), ("cta_pair", C.c_int), ("max_ctas", C.c_int),
        ("dbg_lbo_a", C.c_uint), ("dbg_sbo_a", C.c_uint), ("dbg_lbo_b", C.c_uint), ("dbg_sbo_b", C.c_uint),
        ("workspace", C.c_void_p), ("workspace_bytes", C.c_longlong), ("split_k", C.c_int),
        ("colsum", C.c_void_p),
    ]


class MtArgs(C.Structure):
    _fields_ = [
        ("param", C.c_void_p * MT_MAX), ("grad", C.c_void_p * MT_MAX), ("master", C.c_void_p * MT_MAX),
        ("m", C.c_void_p * MT_MAX), ("v", C.c_void_p * MT_MAX), ("numel", C.c_longlong * MT_MAX),
        ("count", C.c_int), ("param_fp32", C.c_int), ("grad_fp32", C.c_int), ("master_on", C.c_int),
    ]


class CommDesc(C.Structure):
    _fields_ = [
        ("peers", C.c_void_p * MAX_RANKS), ("signals", C.c_void_p * MAX_RANKS), ("mc_base", C.c_void_p),
        ("rank", C.c_int), ("world", C.c_int), ("timeout_cycles", C.c_longlong),
    ]


def _load() -> Optional[C.CDLL]:
    global _LIB, _LIB_ERR
    if _LIB is not None or _LIB_ERR is not None:
        return _LIB
    path = Path(os.environ.get("DTB200_LIB", str(LIB_PATH)))
    if not path.exists():
        _LIB_ERR = f"{path} not found; run `python -m distributed_training_b200._native` to build it"
        return None
    try:
        lib = C.CDLL(str(path))
    except OSError as exc:  # pragma: no cover
        _LIB_ERR = f"cannot load {path}: {exc}"
        return None
    lib.dtb_last_error.restype = C.c_char_p
    lib.dtb_launch_count.restype = C.c_longlong
    lib.dtb_layernorm_bwd_scratch_floats.restype = C.c_size_t
    lib.dtb_attention_bwd_scratch_floats.restype = C.c_size_t
    _LIB = lib
    return lib


def lib() -> C.CDLL:
    handle = _load()
    if handle is None:
        raise RuntimeError(f"dtb200 native library unavailable: {_LIB_ERR}")
    return handle


def available() -> bool:
    """True when the kernels can run here: library loaded and a CUDA device present."""
    return torch.cuda.is_available() and _load() is not None


def require() -> None:
    """Fail loudly on a GPU machine whose extension is missing (never fall back silently)."""
    if torch.cuda.is_available():
        lib()


def _ok(rc: int) -> None:
    if rc != 0:
        raise RuntimeError(f"dtb200: {lib().dtb_last_error().decode()} (rc={rc})")


def _stream(handle: Optional[int] = None) -> C.c_void_p:
    """The stream a launch goes to: torch's current stream, or an explicit raw handle (callers that launch on a side
    stream pass it directly -- entering a ``torch.cuda.stream`` context costs ~10 us of host time per launch)."""
    return C.c_void_p(torch.cuda.current_stream().cuda_stream if handle is None else handle)


def _p(t: Optional[torch.Tensor]) -> C.c_void_p:
    return C.c_void_p(t.data_ptr() if t is not None else 0)


def launch_count() -> int:
    return int(lib().dtb_launch_count())


def reset_launch_count() -> None:
    lib().dtb_reset_launch_count()


# ---------------------------------------------------------------------------
# GEMM
# ---------------------------------------------------------------------------
_GEMM_WS = {}


def _gemm_workspace(device: torch.device) -> torch.Tensor:
    """fp32 scratch for split-K partial sums (64 MiB: 8 slices of a 1024 x 2048 output), one per (device, stream):
    GEMMs queued on different streams (dgrad / wgrad of the native models) may run at the same time."""
    key = (device, torch.cuda.current_stream(device).cuda_stream)
    ws = _GEMM_WS.get(key)
    if ws is None:
        ws = torch.empty(16 << 20, dtype=torch.float32, device=device)
        _GEMM_WS[key] = ws
    return ws


def gemm(a: torch.Tensor, b: torch.Tensor, out: torch.Tensor, *, m: int, n: int, k: int,
         a_mn: bool = False, b_mn: bool = False, lda: Optional[int] = None, ldb: Optional[int] = None,
         ldd: Optional[int] = None, bias: Optional[torch.Tensor] = None, aux: Optional[torch.Tensor] = None,
         ld_aux: Optional[int] = None, act: int = 0, alpha: float = 1.0, beta: float = 1.0,
         d2: Optional[torch.Tensor] = None, batch: Tuple[int, int] = (1, 1),
         a_strides: Tuple[int, int] = (0, 0), b_strides: Tuple[int, int] = (0, 0),
         d_strides: Tuple[int, int] = (0, 0), aux_strides: Tuple[int, int] = (0, 0),
         tile_n: int = 0, cta_pair: int = 0, max_ctas: int = 0, dbg: Tuple[int, int, int, int] = (0, 0, 0, 0),
         split_k: int = 0, colsum: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``out[m,n] = epilogue(alpha * sum_k A(m,k) B(n,k))`` — see ``dtb_gemm_bf16`` in the header.

    ``colsum`` (fp32 ``[n]``, zeroed by the caller): the epilogue adds the column sums of ``out`` to it.

    ``a``/``b`` are bf16; layouts are described by ``*_mn`` and the leading
    dimensions (default: dense rows of the logical storage shape).
    """
    assert a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16
    args = GemmArgs()
    args.a, args.b, args.d = a.data_ptr(), b.data_ptr(), out.data_ptr()
    args.d2 = d2.data_ptr() if d2 is not None else None
    args.aux = aux.data_ptr() if aux is not None else None
    args.bias = bias.data_ptr() if bias is not None else None
    args.lda = lda if lda is not None else (m if a_mn else k)
    args.ldb = ldb if ldb is not None else (n if b_mn else k)
    args.ldd = ldd if ldd is not None else n
    args.ld_aux = ld_aux if ld_aux is not None else n
    args.a_stride_inner, args.a_stride_outer = a_strides
    args.b_stride_inner, args.b_stride_outer = b_strides
    args.d_stride_inner, args.d_stride_outer = d_strides
    args.aux_stride_inner, args.aux_stride_outer = aux_strides
    args.m, args.n, args.k = m, n, k
    args.batch_inner, args.batch_outer = batch
    args.a_mn_major, args.b_mn_major = int(a_mn), int(b_mn)
    args.d_fp32 = int(out.dtype == torch.float32)
    args.aux_fp32 = int(aux is not None and aux.dtype == torch.float32)
    args.bias_fp32 = int(bias is not None and bias.dtype == torch.float32)
    args.act = act
    args.alpha, args.beta = alpha, beta
    args.tile_n, args.cta_pair, args.max_ctas = tile_n, cta_pair, max_ctas
    args.dbg_lbo_a, args.dbg_sbo_a, args.dbg_lbo_b, args.dbg_sbo_b = dbg
    if split_k != 1 and batch == (1, 1) and k >= 512:
        ws = _gemm_workspace(a.device)
        args.workspace, args.workspace_bytes = ws.data_ptr(), ws.numel() * 4
    args.split_k = split_k
    if colsum is not None:
        assert colsum.dtype == torch.float32 and colsum.numel() >= n
        args.colsum = colsum.data_ptr()
    _ok(lib().dtb_gemm_bf16(C.byref(args), _stream()))
    return out


class GatherArgs(C.Structure):
    _fields_ = [("peer_shard", C.c_void_p * MAX_RANKS), ("flags", C.c_void_p),
                ("world", C.c_int), ("rank", C.c_int), ("rows_per_rank", C.c_int)]


def gemm_allgather_flag_ints(n: int) -> int:
    return (n + 127) // 128 + 1


def gemm_allgather(a: torch.Tensor, b_full: torch.Tensor, out: torch.Tensor, peer_shard_ptrs: Sequence[int],
                   flags: torch.Tensor, rank: int, rows_per_rank: int, *, m: int, n: int, k: int,
                   bias: Optional[torch.Tensor] = None, aux: Optional[torch.Tensor] = None, act: int = 0,
                   d2: Optional[torch.Tensor] = None, beta: float = 1.0, max_ctas: int = 0, tile_n: int = 0,
                   cta_pair: int = 0, gather_args: Optional["GatherArgs"] = None) -> torch.Tensor:
    """Linear forward whose weight all-gather runs inside the GEMM kernel (``dtb_gemm_bf16_allgather``).

    ``b_full`` is the local [n, k] weight buffer that receives the gathered rows; ``peer_shard_ptrs[r]`` is
    rank r's [rows_per_rank, k] shard as mapped in this process."""
    args = GemmArgs()
    args.a, args.b, args.d = a.data_ptr(), b_full.data_ptr(), out.data_ptr()
    args.d2 = d2.data_ptr() if d2 is not None else None
    args.aux = aux.data_ptr() if aux is not None else None
    args.bias = bias.data_ptr() if bias is not None else None
    args.lda, args.ldb, args.ldd, args.ld_aux = k, k, n, n
    args.m, args.n, args.k = m, n, k
    args.batch_inner, args.batch_outer = 1, 1
    args.d_fp32 = int(out.dtype == torch.float32)
    args.aux_fp32 = int(aux is not None and aux.dtype == torch.float32)
    args.bias_fp32 = int(bias is not None and bias.dtype == torch.float32)
    args.act = act
    args.alpha, args.beta = 1.0, beta
    args.tile_n, args.cta_pair, args.max_ctas = tile_n, cta_pair, max_ctas
    g = gather_args if gather_args is not None else make_gather_args(peer_shard_ptrs, flags, rank, rows_per_rank)
    _ok(lib().dtb_gemm_bf16_allgather(C.byref(args), C.byref(g), _stream()))
    return out


def make_gather_args(peer_shard_ptrs: Sequence[int], flags: torch.Tensor, rank: int,
                     rows_per_rank: int) -> "GatherArgs":
    """The per-weight constant half of a ``gemm_allgather`` call (build once, pass as ``gather_args``)."""
    g = GatherArgs()
    for r, ptr in enumerate(peer_shard_ptrs):
        g.peer_shard[r] = ptr
    g.flags = flags.data_ptr()
    g.world, g.rank, g.rows_per_rank = len(peer_shard_ptrs), rank, rows_per_rank
    return g


def linear_small_fwd(x, w, b, y):
    rows, in_f = x.shape
    _ok(lib().dtb_linear_small_fwd(_p(x), _p(w), _p(b), _p(y), rows, in_f, w.shape[0], _stream()))
    return y


def linear_small_bwd(x, w, dy, dx, dw, db):
    rows, in_f = x.shape
    _ok(lib().dtb_linear_small_bwd(_p(x), _p(w), _p(dy), _p(dx), _p(dw), _p(db), rows, in_f, w.shape[0], _stream()))


# ---------------------------------------------------------------------------
# elementwise / reductions
# ---------------------------------------------------------------------------
def gelu_fwd(x, y):
    _ok(lib().dtb_gelu_fwd_bf16(_p(x), _p(y), C.c_longlong(x.numel()), _stream()))
    return y


def gelu_bwd(x, dy, dx):
    _ok(lib().dtb_gelu_bwd_bf16(_p(x), _p(dy), _p(dx), C.c_longlong(x.numel()), _stream()))
    return dx


def add(a, b, out):
    _ok(lib().dtb_add_bf16(_p(a), _p(b), _p(out), C.c_longlong(a.numel()), _stream()))
    return out


def colsum(x2d: torch.Tensor, out: torch.Tensor, ldx: Optional[int] = None):
    rows, cols = x2d.shape
    _ok(lib().dtb_colsum_bf16(_p(x2d), _p(out), rows, cols, C.c_longlong(ldx if ldx is not None else cols),
                              int(out.dtype == torch.float32), _stream()))
    return out


_COLSUM_SCRATCH = {}


def colsum_split(x2d: torch.Tensor, out: torch.Tensor, ldx: Optional[int] = None):
    """Column sums with the rows split across CTAs (tall, narrow inputs); scratch is cached per device."""
    rows, cols = x2d.shape
    need = 64 * cols
    key = x2d.device
    buf = _COLSUM_SCRATCH.get(key)
    if buf is None or buf.numel() < need:
        buf = torch.empty(max(need, 64 * 8192), dtype=torch.float32, device=x2d.device)
        _COLSUM_SCRATCH[key] = buf
    _ok(lib().dtb_colsum_bf16_split(_p(x2d), _p(out), rows, cols, C.c_longlong(ldx if ldx is not None else cols),
                                    int(out.dtype == torch.float32), _p(buf), _stream()))
    return out


def cast_f32_to_bf16(src, dst):
    _ok(lib().dtb_cast_f32_to_bf16(_p(src), _p(dst), C.c_longlong(src.numel()), _stream()))
    return dst


def take_f32_to_bf16(src, dst):
    """``dst = bf16(src); src = 0`` (hand-over of a GEMM ``colsum`` accumulator to a bf16 gradient)."""
    assert src.dtype == torch.float32 and dst.dtype == torch.bfloat16 and dst.numel() >= src.numel()
    _ok(lib().dtb_take_f32_to_bf16(_p(src), _p(dst), C.c_longlong(src.numel()), _stream()))
    return dst


def cast_bf16_to_f32(src, dst):
    _ok(lib().dtb_cast_bf16_to_f32(_p(src), _p(dst), C.c_longlong(src.numel()), _stream()))
    return dst


def gather_rows(src: torch.Tensor, idx: torch.Tensor, dst: torch.Tensor, src_rows: int = 0, max_ctas: int = 0,
                stream: Optional[int] = None) -> torch.Tensor:
    """``dst[i] = src[idx[i] % src_rows]`` (rows of the first dimension); ``src`` may be a PINNED host tensor, which
    the kernel reads over PCIe.  ``idx`` (int64) and ``dst`` live on the device."""
    assert idx.dtype == torch.int64 and idx.is_cuda and dst.is_cuda and src.is_contiguous() and dst.is_contiguous()
    assert src.is_cuda or src.is_pinned(), "gather_rows: host sources must be pinned"
    row_bytes = (src.numel() // max(src.shape[0], 1)) * src.element_size()
    _ok(lib().dtb_gather_rows(_p(src), _p(idx), _p(dst), int(idx.numel()), C.c_longlong(row_bytes),
                              C.c_longlong(src_rows), max_ctas, _stream(stream)))
    return dst


def sumsq(x, out):
    _ok(lib().dtb_sumsq(_p(x), int(x.dtype == torch.float32), C.c_longlong(x.numel()), _p(out), _stream()))
    return out


def mean_f32(x, out):
    _ok(lib().dtb_mean_f32(_p(x), C.c_longlong(x.numel()), _p(out), _stream()))
    return out


# ---------------------------------------------------------------------------
# layernorm / losses / embedding
# ---------------------------------------------------------------------------
def layernorm_fwd(x, residual, gamma, beta, y, xres_out, mean, rstd, eps=1e-5):
    rows, cols = x.numel() // x.shape[-1], x.shape[-1]
    _ok(lib().dtb_layernorm_fwd(_p(x), _p(residual), _p(gamma), _p(beta), _p(y), _p(xres_out), _p(mean), _p(rstd),
                                rows, cols, C.c_float(eps), _stream()))


def layernorm_bwd_scratch_floats(rows: int, cols: int) -> int:
    return int(lib().dtb_layernorm_bwd_scratch_floats(rows, cols))


def layernorm_bwd(dy, x, gamma, mean, rstd, dres_in, dx, dgamma, dbeta, scratch, accumulate=False):
    rows, cols = x.numel() // x.shape[-1], x.shape[-1]
    _ok(lib().dtb_layernorm_bwd(_p(dy), _p(x), _p(gamma), _p(mean), _p(rstd), _p(dres_in), _p(dx), _p(dgamma),
                                _p(dbeta), _p(scratch), rows, cols, int(accumulate), _stream()))


def layernorm_bwd_bf16grads(dy, x, gamma, mean, rstd, dres_in, dx, dgamma, dbeta, scratch, accumulate=False):
    """``layernorm_bwd`` writing dgamma / dbeta as bf16 (no separate cast launches)."""
    assert dgamma.dtype == torch.bfloat16 and dbeta.dtype == torch.bfloat16
    rows, cols = x.numel() // x.shape[-1], x.shape[-1]
    _ok(lib().dtb_layernorm_bwd_bf16grads(_p(dy), _p(x), _p(gamma), _p(mean), _p(rstd), _p(dres_in), _p(dx),
                                          _p(dgamma), _p(dbeta), _p(scratch), rows, cols, int(accumulate), _stream()))


def cross_entropy_fwd_bwd(logits, targets, row_loss, vocab, grad_scale, write_grad=True):
    rows, ld = logits.numel() // logits.shape[-1], logits.shape[-1]
    _ok(lib().dtb_cross_entropy_fwd_bwd(_p(logits), _p(targets), _p(row_loss), rows, vocab, C.c_longlong(ld),
                                        C.c_float(grad_scale), int(write_grad), _stream()))


def soft_cross_entropy_fwd_bwd(logits, targets, row_loss, dlogits, grad_scale):
    rows, cols = logits.shape
    _ok(lib().dtb_soft_cross_entropy_fwd_bwd(_p(logits), _p(targets), _p(row_loss), _p(dlogits), rows, cols,
                                             C.c_float(grad_scale), _stream()))


def mse_fwd_bwd(pred, target, loss_sum, dpred, grad_scale):
    _ok(lib().dtb_mse_fwd_bwd(_p(pred), _p(target), _p(loss_sum), _p(dpred), C.c_longlong(pred.numel()),
                              C.c_float(grad_scale), _stream()))


def embedding_fwd(tokens, wte, wpe, out):
    batch, seq = tokens.shape
    _ok(lib().dtb_embedding_fwd(_p(tokens), _p(wte), _p(wpe), _p(out), batch, seq, wte.shape[1], _stream()))


def embedding_bwd(dout, tokens, dwte, dwpe):
    batch, seq = tokens.shape
    _ok(lib().dtb_embedding_bwd(_p(dout), _p(tokens), _p(dwte), _p(dwpe), batch, seq, dwte.shape[1], 1, _stream()))


def attention_fwd(qkv, out, lse, heads: int):
    batch, seq, c3 = qkv.shape
    _ok(lib().dtb_attention_fwd(_p(qkv), _p(out), _p(lse), batch, seq, heads, c3 // (3 * heads), _stream()))


def attention_bwd_scratch_floats(batch: int, seq: int, heads: int, head_dim: int = 64) -> int:
    return int(lib().dtb_attention_bwd_scratch_floats(batch, seq, heads, head_dim))


def attention_bwd(qkv, out, dout, lse, dqkv, scratch, heads: int):
    batch, seq, c3 = qkv.shape
    _ok(lib().dtb_attention_bwd(_p(qkv), _p(out), _p(dout), _p(lse), _p(dqkv), _p(scratch), batch, seq, heads,
                                c3 // (3 * heads), _stream()))


# ---------------------------------------------------------------------------
# optimizers
# ---------------------------------------------------------------------------
def _mt_chunks(params, grads, masters, ms, vs):
    by_kind = {}
    for i, (p, g) in enumerate(zip(params, grads)):
        by_kind.setdefault((p.dtype, g.dtype), []).append(i)
    for (pd, gd), idxs in by_kind.items():
        for lo in range(0, len(idxs), MT_MAX):
            chunk = idxs[lo: lo + MT_MAX]
            a = MtArgs()
            a.count = len(chunk)
            a.param_fp32 = int(pd == torch.float32)
            a.grad_fp32 = int(gd == torch.float32)
            a.master_on = int(masters is not None and pd != torch.float32)
            for j, i in enumerate(chunk):
                a.param[j] = params[i].data_ptr()
                a.grad[j] = grads[i].data_ptr()
                a.master[j] = masters[i].data_ptr() if (a.master_on and masters[i] is not None) else None
                a.m[j] = ms[i].data_ptr() if ms is not None else None
                a.v[j] = vs[i].data_ptr() if vs is not None else None
                a.numel[j] = params[i].numel()
            yield a


def multi_tensor_sgd(params, grads, masters, bufs, lr, momentum, weight_decay):
    for a in _mt_chunks(params, grads, masters, bufs, None):
        _ok(lib().dtb_multi_tensor_sgd(C.byref(a), C.c_float(lr), C.c_float(momentum), C.c_float(weight_decay),
                                       _stream()))


def multi_tensor_adamw(params, grads, ms, vs, masters, lr, b1, b2, eps, wd, bc1, bc2):
    for a in _mt_chunks(params, grads, masters, ms, vs):
        _ok(lib().dtb_multi_tensor_adamw(C.byref(a), C.c_float(lr), C.c_float(b1), C.c_float(b2), C.c_float(eps),
                                         C.c_float(wd), C.c_float(bc1), C.c_float(bc2), _stream()))


def shard_update(shard, master, grad, state, name, lr, wd, momentum, betas, eps, step):
    """Optimizer update of one local fp32 master shard + bf16 shard (sharded strategy, unfused path)."""
    if name == "adamw":
        multi_tensor_adamw([shard], [grad], [state["exp_avg"]], [state["exp_avg_sq"]],
                           [master] if shard.dtype != torch.float32 else None,
                           lr, betas[0], betas[1], eps, wd, 1 - betas[0] ** step, 1 - betas[1] ** step)
    else:
        bufs = [state["momentum_buffer"]] if momentum else None
        multi_tensor_sgd([shard], [grad], [master] if shard.dtype != torch.float32 else None, bufs, lr, momentum, wd)
    if shard.dtype == torch.float32:
        master.copy_(shard)


# ---------------------------------------------------------------------------
# symmetric memory + collectives
# ---------------------------------------------------------------------------
def symm_alloc(nbytes: int) -> Tuple[int, bytes]:
    ptr = C.c_void_p()
    handle = (C.c_ubyte * 64)()
    _ok(lib().dtb_symm_alloc(C.c_size_t(nbytes), C.byref(ptr), handle))
    return int(ptr.value), bytes(handle)


def symm_open(handle: bytes) -> int:
    ptr = C.c_void_p()
    buf = (C.c_ubyte * 64).from_buffer_copy(handle)
    _ok(lib().dtb_symm_open(buf, C.byref(ptr)))
    return int(ptr.value)


def symm_close(ptr: int) -> None:
    _ok(lib().dtb_symm_close(C.c_void_p(ptr)))


def symm_free(ptr: int) -> None:
    _ok(lib().dtb_symm_free(C.c_void_p(ptr)))


def symm_vmm_probe() -> Tuple[bool, bool]:
    """(VMM with fd handles supported, multicast objects supported) on the current device."""
    v, m = C.c_int(0), C.c_int(0)
    _ok(lib().dtb_symm_vmm_probe(C.byref(v), C.byref(m)))
    return bool(v.value), bool(m.value)


def symm_vmm_granularity(world: int, multicast: bool) -> int:
    g = C.c_size_t(0)
    _ok(lib().dtb_symm_vmm_granularity(world, int(multicast), C.byref(g)))
    return int(g.value)


def symm_vmm_create(nbytes: int) -> Tuple[int, int]:
    """(allocation handle, exported POSIX fd) of ``nbytes`` of device memory on the current device."""
    h, fd = C.c_ulonglong(0), C.c_int(-1)
    _ok(lib().dtb_symm_vmm_create(C.c_size_t(nbytes), C.byref(h), C.byref(fd)))
    return int(h.value), int(fd.value)


def symm_vmm_import(fd: int) -> int:
    h = C.c_ulonglong(0)
    _ok(lib().dtb_symm_vmm_import(int(fd), C.byref(h)))
    return int(h.value)


def symm_vmm_map(handle: int, nbytes: int, alignment: int = 0) -> int:
    ptr = C.c_void_p()
    _ok(lib().dtb_symm_vmm_map(C.c_ulonglong(handle), C.c_size_t(nbytes), C.c_size_t(alignment), C.byref(ptr)))
    return int(ptr.value)


def symm_vmm_unmap(ptr: int, nbytes: int) -> None:
    _ok(lib().dtb_symm_vmm_unmap(C.c_void_p(ptr), C.c_size_t(nbytes)))


def symm_vmm_release(handle: int) -> None:
    _ok(lib().dtb_symm_vmm_release(C.c_ulonglong(handle)))


def symm_mc_create(nbytes: int, world: int) -> Tuple[int, int]:
    h, fd = C.c_ulonglong(0), C.c_int(-1)
    _ok(lib().dtb_symm_mc_create(C.c_size_t(nbytes), world, C.byref(h), C.byref(fd)))
    return int(h.value), int(fd.value)


def symm_mc_add_device(mc_handle: int) -> None:
    _ok(lib().dtb_symm_mc_add_device(C.c_ulonglong(mc_handle)))


def symm_mc_bind(mc_handle: int, mem_handle: int, nbytes: int) -> None:
    _ok(lib().dtb_symm_mc_bind(C.c_ulonglong(mc_handle), C.c_ulonglong(mem_handle), C.c_size_t(nbytes)))


EXCLUSIVE = 1 << 20     # OR into max_ctas: whole-SM (1024-thread) comm CTAs instead of co-resident 256-thread ones


def _cta_arg(max_ctas: int, channel: int) -> int:
    return (max_ctas & 0xFFFF) | ((channel & 0xF) << 16) | (max_ctas & EXCLUSIVE)


def comm_barrier(desc: CommDesc, channel: int = 0, stream: Optional[int] = None):
    _ok(lib().dtb_comm_barrier(C.byref(desc), channel, _stream(stream)))


def comm_broadcast(desc: CommDesc, offset: int, nbytes: int, root: int, max_ctas: int = 0, channel: int = 0,
                   stream: Optional[int] = None):
    _ok(lib().dtb_comm_broadcast(C.byref(desc), C.c_size_t(offset), C.c_size_t(nbytes), root,
                                 _cta_arg(max_ctas, channel), _stream(stream)))


def comm_all_reduce_avg(desc: CommDesc, offset: int, numel: int, fp32: bool, algo: int = 0, max_ctas: int = 0,
                        channel: int = 0, stream: Optional[int] = None):
    _ok(lib().dtb_comm_all_reduce_avg(C.byref(desc), C.c_size_t(offset), C.c_longlong(numel), int(fp32), algo,
                                      _cta_arg(max_ctas, channel), _stream(stream)))


def comm_all_reduce_sgd(desc: CommDesc, grad_offset: int, param_offset: int, numel: int, lr: float,
                        weight_decay: float, max_ctas: int = 0, channel: int = 0, stream: Optional[int] = None):
    _ok(lib().dtb_comm_all_reduce_sgd(C.byref(desc), C.c_size_t(grad_offset), C.c_size_t(param_offset),
                                      C.c_longlong(numel), C.c_float(lr), C.c_float(weight_decay),
                                      _cta_arg(max_ctas, channel), _stream(stream)))


def comm_all_reduce_ce(desc: CommDesc, grad_offset: int, param_offset: int, sgd: bool, numel: int,
                       staging: torch.Tensor, lr: float = 0.0, weight_decay: float = 0.0, max_ctas: int = 0,
                       channel: int = 0, stream: Optional[int] = None):
    _ok(lib().dtb_comm_all_reduce_ce(C.byref(desc), C.c_size_t(grad_offset), C.c_size_t(param_offset), int(sgd),
                                     C.c_longlong(numel), _p(staging),
                                     C.c_size_t(staging.numel() * staging.element_size()), C.c_float(lr),
                                     C.c_float(weight_decay), _cta_arg(max_ctas, channel), _stream(stream)))


def comm_all_gather(desc: CommDesc, dst_offset: int, src_offset: int, numel: int, esize: int, max_ctas: int = 0,
                    channel: int = 0, stream: Optional[int] = None):
    _ok(lib().dtb_comm_all_gather(C.byref(desc), C.c_size_t(dst_offset), C.c_size_t(src_offset), C.c_longlong(numel),
                                  esize, _cta_arg(max_ctas, channel), _stream(stream)))


def comm_reduce_scatter_avg(desc: CommDesc, out: torch.Tensor, src_offset: int, numel: int, src_fp32: bool,
                            max_ctas: int = 0, channel: int = 0, stream: Optional[int] = None):
    _ok(lib().dtb_comm_reduce_scatter_avg(C.byref(desc), _p(out), int(out.dtype == torch.float32),
                                          C.c_size_t(src_offset), C.c_longlong(numel), int(src_fp32),
                                          _cta_arg(max_ctas, channel), _stream(stream)))


def comm_rs_optim(desc: CommDesc, grad_offset: int, shard_offset: int, master, m, v, numel: int, opt: int,
                  lr: float, b1: float, b2: float, eps: float, wd: float, bc1: float, bc2: float, momentum: float,
                  max_ctas: int = 0, channel: int = 0, stream: Optional[int] = None):
    _ok(lib().dtb_comm_rs_optim(C.byref(desc), C.c_size_t(grad_offset), C.c_size_t(shard_offset), _p(master), _p(m),
                                _p(v), C.c_longlong(numel), opt, C.c_float(lr), C.c_float(b1), C.c_float(b2),
                                C.c_float(eps), C.c_float(wd), C.c_float(bc1), C.c_float(bc2), C.c_float(momentum),
                                _cta_arg(max_ctas, channel), _stream(stream)))


def debug_stream_probe(a, b, p, q, nbytes: int, ctas: int):
    _ok(lib().dtb_debug_stream_probe(_p(a), _p(b), _p(p), _p(q), C.c_longlong(nbytes), ctas, _stream()))


def debug_memcpy_async(dst_ptr: int, src_ptr: int, nbytes: int, stream: Optional[int] = None) -> None:
    _ok(lib().dtb_debug_memcpy_async(C.c_void_p(dst_ptr), C.c_void_p(src_ptr), C.c_size_t(nbytes), _stream(stream)))


def comm_error_flag(desc: CommDesc) -> int:
    return int(lib().dtb_comm_error_flag(C.byref(desc)))
