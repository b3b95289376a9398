"""Hand-scheduled GPT-2 training step on the sm_100a kernels (no autograd).

Forward, per block (residual stream ``x``; every matmul is ``native.gemm``):

    a    = LN1(x)                                   layernorm_fwd
    qkv  = a Wqkv^T + b                             GEMM + bias epilogue
    att  = causal_attention(qkv)                    tcgen05 flash attention (S/P never leave the SM)
    xm   = x + att Wo^T + bo                        GEMM + bias + residual epilogue
    m    = LN2(xm)
    h    = gelu(m Wfc^T + bfc)   (z kept)           GEMM + bias + GELU epilogue, pre-activation as 2nd output
    x'   = xm + h Wpr^T + bpr                       GEMM + bias + residual epilogue

then ``LN_f``, the tied LM head GEMM over the padded vocabulary and the fused softmax-CE kernel that
overwrites the logits with their gradient.  Backward walks the same graph in reverse: each Linear is
a dgrad GEMM (GELU' fused where a GELU preceded) followed by a wgrad GEMM that writes straight into
the caller's gradient bucket, bias gradients are column sums, LayerNorm backward folds the residual
branch's gradient in.  A weight's dgrad is always queued before its wgrad is handed to the reducer
(``grad_sink.mark_ready``), because a fused reduce+optimizer kernel may overwrite the weight.

The reference has no transformer; the baseline for this path is eager PyTorch + SDPA.
"""
from __future__ import annotations

import os
from typing import Any, Dict

import torch
from torch import nn

from ..ops import native

#: weight gradients on a side stream beside their dgrad (DTB_ASYNC_WGRAD=0: one stream)
_ASYNC_WGRAD = os.environ.get("DTB_ASYNC_WGRAD", "1") != "0"
#: the caller's stream waits for a weight gradient only before a buffer it reads is overwritten
#: (DTB_ASYNC_WGRAD_DEFER=0: right after queuing it)
_DEFER_JOIN = os.environ.get("DTB_ASYNC_WGRAD_DEFER", "1") != "0"


def _side_stream(model, device: torch.device) -> "torch.cuda.Stream":
    st = getattr(model, "_wgrad_stream", None)
    if st is None or st.device != device:
        st = torch.cuda.Stream(device)
        model._wgrad_stream = st
    return st


def _workspace(model, B: int, T: int, device: torch.device) -> Dict[str, Any]:
    key = (B, T, device)
    ws = model._ws.get(key)
    if ws is not None:
        return ws
    cfg = model.cfg
    C, H, V, L = cfg.n_embd, cfg.n_head, cfg.padded_vocab, cfg.n_layer
    R = B * T
    bf = dict(dtype=torch.bfloat16, device=device)
    f32 = dict(dtype=torch.float32, device=device)

    def act(cols):
        return torch.empty(R, cols, **bf)

    # recompute_gelu: h = gelu(z) is not kept per block (it is 4 of the 16 C-wide activations a block saves); forward
    # writes it into one shared buffer and backward rebuilds it from z with one elementwise kernel per block
    recompute = bool(getattr(model, "recompute_gelu", False))
    h_shared = act(4 * C) if recompute else None
    blocks = []
    for _ in range(L):
        blocks.append({
            "a": act(C), "qkv": torch.empty(B, T, 3 * C, **bf), "att": act(C), "xm": act(C), "m": act(C),
            "z": act(4 * C), "h": h_shared if recompute else act(4 * C), "x_out": act(C),
            "lse": torch.empty(B, H, T, **f32),
            "mean1": torch.empty(R, **f32), "rstd1": torch.empty(R, **f32),
            "mean2": torch.empty(R, **f32), "rstd2": torch.empty(R, **f32),
        })
    ws = {
        "blocks": blocks, "recompute_gelu": recompute,
        "x0": act(C), "xf": act(C), "meanf": torch.empty(R, **f32), "rstdf": torch.empty(R, **f32),
        "logits": torch.empty(R, V, **bf),
        "row_loss": torch.empty(R, **f32), "loss": torch.empty((), **f32),
        # backward temporaries (reused by every block)
        "dx": [act(C), act(C)], "dC": act(C), "d4C": act(4 * C), "dqkv": torch.empty(B, T, 3 * C, **bf),
        "ln_scratch": torch.empty(native.layernorm_bwd_scratch_floats(R, C), **f32),
        "ln_dg": torch.empty(C, **f32), "ln_db": torch.empty(C, **f32),
        "attn_scratch": torch.empty(native.attention_bwd_scratch_floats(B, T, H), **f32),
    }
    if device.type == "cuda":
        ws["ev_fork"], ws["ev_join"] = torch.cuda.Event(), torch.cuda.Event()
        ws["ev_pool_w"] = [torch.cuda.Event() for _ in range(16)]
        ws["ev_pool_d"] = [torch.cuda.Event() for _ in range(4)]
    model._ws = {key: ws}
    return ws


def gpt2_forward_backward(model, tokens: torch.Tensor, targets: torch.Tensor, grad_sink: Any = None,
                          max_ctas: int = 0) -> torch.Tensor:
    cfg = model.cfg
    B, T = tokens.shape
    C, H, V = cfg.n_embd, cfg.n_head, cfg.padded_vocab
    R = B * T
    dev = tokens.device
    ws = _workspace(model, B, T, dev)
    tokens = tokens.contiguous()
    targets = targets.contiguous().view(-1)
    eps = cfg.layer_norm_eps

    def grad_of(p: nn.Parameter) -> torch.Tensor:
        if grad_sink is not None:
            return grad_sink.grad_view(p)
        if p.grad is None:
            p.grad = torch.empty_like(p)
        return p.grad

    def ready(*params: nn.Parameter) -> None:
        if grad_sink is None:
            return
        if deferred:
            # handed over at the next fork, on the side stream, once that stream is ordered behind this one: a bucket
            # is reduced behind the stream that is current when its LAST parameter is marked, so every mark has to
            # come from the stream that also carries the weight gradients
            pending_ready.extend(params)
            return
        for p in params:
            grad_sink.mark_ready(p)

    # a sharded grad_sink gathers each weight inside the forward GEMM that consumes it (gemm_forward) and hands out
    # the transiently gathered matrix wherever else it is read (weight_view: embedding lookup, LM head, dgrad)
    gemm_fwd = getattr(grad_sink, "gemm_forward", None) or native.gemm
    wv = getattr(grad_sink, "weight_view", None) or (lambda p: p)

    def linear_fwd(x, lin: nn.Linear, out, *, act=0, d2=None, residual=None):
        out_f, in_f = lin.weight.shape
        gemm_fwd(x, lin.weight, out, m=R, n=out_f, k=in_f, bias=lin.bias, act=act, d2=d2, aux=residual,
                 beta=1.0, max_ctas=max_ctas)

    # dgrad and wgrad of a Linear read the same inputs and write different outputs: the wgrad (+ bias column sums)
    # goes to a side stream and the two GEMMs share the GPU.  Each of them alone leaves most SMs idle in its last
    # wave (8192 x 768 x 768: 96 tiles on 74 CTA pairs = 1.3 waves), the other one's CTAs take those SMs.
    # Two forms: fork-join (the caller's stream waits for the wgrad right away; sharded sinks, whose slot pools are
    # ordered on the caller's stream) and deferred (plain gradients / ReplicatedModule): the caller's stream goes on
    # with LayerNorm and attention backward and waits for a wgrad only before one of the buffers it reads is
    # overwritten, so the weight gradients also run under those latency-bound kernels.
    fork = _ASYNC_WGRAD and tokens.is_cuda
    deferred = fork and _DEFER_JOIN and (grad_sink is None or hasattr(grad_sink, "async_wgrad_mode"))
    main = torch.cuda.current_stream() if fork else None
    side = _side_stream(model, dev) if fork else None
    ev_fork, ev_join = (ws["ev_fork"], ws["ev_join"]) if fork else (None, None)
    pending_ready: list = []
    last_read: Dict[int, Any] = {}          # buffer -> event of the last side-stream kernel that reads it
    # two rotating pools.  Side-stream events: a re-recorded one stands for a LATER kernel of that in-order stream,
    # so a stale entry in `last_read` can only wait for more than it needs.  Caller-stream events are consumed by
    # `side.wait_event` right where they are recorded.
    pools = {"w": ws.get("ev_pool_w", []), "d": ws.get("ev_pool_d", [])}
    pool_at = {"w": 0, "d": 0}

    def next_event(kind: str):
        ev = pools[kind][pool_at[kind] % len(pools[kind])]
        pool_at[kind] += 1
        return ev

    def before_write(*bufs: torch.Tensor) -> None:
        """The caller's stream is about to overwrite ``bufs``: wait for the weight gradients that still read them."""
        if deferred:
            for buf in bufs:
                ev = last_read.pop(buf.data_ptr(), None)
                if ev is not None:
                    main.wait_event(ev)

    def wgrad_pair(dy, x_in, w: nn.Parameter, bias, out_f, in_f, dgrad, dx_out=None, mark=True):
        """Run ``dgrad()`` on the caller's stream and dW = dy^T x_in (+ db) beside it."""
        gw = grad_of(w)
        gb = grad_of(bias) if bias is not None else None
        if fork:
            ev_fork.record(main)              # dy, x_in and the gradient slots are ready
            side.wait_event(ev_fork)
        if dx_out is not None:
            before_write(dx_out)
        dgrad()
        ev_d = None
        if deferred:
            ev_d = next_event("d")
            ev_d.record(main)                 # the weight has been read: its bucket may be stepped
        if fork:
            torch.cuda.set_stream(side)
        try:
            if deferred and pending_ready:    # LayerNorm parameters finished before this fork (side is behind them now)
                for p in pending_ready:
                    grad_sink.mark_ready(p)
                pending_ready.clear()
            native.gemm(dy, x_in, gw, m=out_f, n=in_f, k=R, a_mn=True, b_mn=True, max_ctas=max_ctas)
            if gb is not None:
                native.colsum_split(dy, gb)
            if deferred:
                ev_w = next_event("w")
                ev_w.record(side)
                last_read[dy.data_ptr()] = ev_w
                last_read[x_in.data_ptr()] = ev_w
                if grad_sink is not None and mark:
                    side.wait_event(ev_d)
                    grad_sink.mark_ready(w)
                    if bias is not None:
                        grad_sink.mark_ready(bias)
            elif fork:
                ev_join.record(side)
        finally:
            if fork:
                torch.cuda.set_stream(main)
        if fork and not deferred:
            main.wait_event(ev_join)

    def linear_bwd(dy, x_in, lin: nn.Linear, dx_out, *, gelu_pre=None):
        """dx = dy W (optionally * gelu'(pre)); dW = dy^T x_in, db = colsum(dy) into the buckets."""
        out_f, in_f = lin.weight.shape

        def dgrad():
            if dx_out is not None:
                native.gemm(dy, wv(lin.weight), dx_out, m=R, n=in_f, k=out_f, b_mn=True, aux=gelu_pre,
                            act=2 if gelu_pre is not None else 0, max_ctas=max_ctas)

        wgrad_pair(dy, x_in, lin.weight, lin.bias, out_f, in_f, dgrad, dx_out)
        if not deferred:
            ready(lin.weight, lin.bias)

    def ln_bwd(dy, x, ln: nn.LayerNorm, mean, rstd, dres, dx):
        before_write(dx)
        native.layernorm_bwd_bf16grads(dy, x, ln.weight, mean, rstd, dres, dx, grad_of(ln.weight), grad_of(ln.bias),
                                       ws["ln_scratch"])
        ready(ln.weight, ln.bias)

    # ------------------------------------------------------------------ forward
    native.embedding_fwd(tokens, wv(model.wte.weight), wv(model.wpe.weight), ws["x0"])
    x = ws["x0"]
    for blk, a in zip(model.h, ws["blocks"]):
        native.layernorm_fwd(x, None, blk.ln_1.weight, blk.ln_1.bias, a["a"], None, a["mean1"], a["rstd1"], eps)
        linear_fwd(a["a"], blk.attn_qkv, a["qkv"])
        native.attention_fwd(a["qkv"], a["att"], a["lse"], H)
        linear_fwd(a["att"], blk.attn_proj, a["xm"], residual=x)
        native.layernorm_fwd(a["xm"], None, blk.ln_2.weight, blk.ln_2.bias, a["m"], None, a["mean2"], a["rstd2"], eps)
        linear_fwd(a["m"], blk.mlp_fc, a["h"], act=1, d2=a["z"])
        linear_fwd(a["h"], blk.mlp_proj, a["x_out"], residual=a["xm"])
        a["x_in"] = x
        x = a["x_out"]
    native.layernorm_fwd(x, None, model.ln_f.weight, model.ln_f.bias, ws["xf"], None, ws["meanf"], ws["rstdf"], eps)
    native.gemm(ws["xf"], wv(model.wte.weight), ws["logits"], m=R, n=V, k=C, max_ctas=max_ctas)
    native.cross_entropy_fwd_bwd(ws["logits"], targets, ws["row_loss"], cfg.vocab_size, 1.0 / R)
    native.mean_f32(ws["row_loss"], ws["loss"])

    # ----------------------------------------------------------------- backward
    dlogits = ws["logits"]
    dxa, dxb = ws["dx"]
    # LM head (tied): dxf = dlogits wte ; dwte = dlogits^T xf  (embedding gradient is added at the very end)
    # (tied weight: handed to the reducer at the very end, after the embedding gradient has been added)
    wgrad_pair(dlogits, ws["xf"], model.wte.weight, None, V, C,
               lambda: native.gemm(dlogits, wv(model.wte.weight), ws["dC"], m=R, n=C, k=V, b_mn=True,
                                   max_ctas=max_ctas), mark=False)
    ln_bwd(ws["dC"], x, model.ln_f, ws["meanf"], ws["rstdf"], None, dxa)
    dx = dxa                                   # gradient w.r.t. the block output
    other = dxb
    for blk, a in zip(reversed(list(model.h)), reversed(ws["blocks"])):
        # x_out = xm + h Wpr^T + b   -> dh (with GELU') ; dWpr, dbpr
        if ws["recompute_gelu"]:
            before_write(a["h"])
            native.gelu_fwd(a["z"], a["h"])
        linear_bwd(dx, a["h"], blk.mlp_proj, ws["d4C"], gelu_pre=a["z"])
        # z = m Wfc^T + b            -> dm ; dWfc, dbfc
        linear_bwd(ws["d4C"], a["m"], blk.mlp_fc, ws["dC"])
        # m = LN2(xm); residual branch adds dx
        ln_bwd(ws["dC"], a["xm"], blk.ln_2, a["mean2"], a["rstd2"], dx, other)
        dxm = other
        # xm = x_in + att Wo^T + b   -> datt ; dWo, dbo
        linear_bwd(dxm, a["att"], blk.attn_proj, ws["dC"])
        before_write(ws["dqkv"])
        native.attention_bwd(a["qkv"], a["att"], ws["dC"], a["lse"], ws["dqkv"], ws["attn_scratch"], H)
        # qkv = a Wqkv^T + b         -> da ; dWqkv, dbqkv
        linear_bwd(ws["dqkv"].view(R, 3 * C), a["a"], blk.attn_qkv, ws["dC"])
        # a = LN1(x_in); residual branch adds dxm
        ln_bwd(ws["dC"], a["x_in"], blk.ln_1, a["mean1"], a["rstd1"], dxm, dx)
        # dx now holds the gradient w.r.t. this block's input; `other` is free again
    if fork:
        main.wait_stream(side)     # the LM head's weight gradient is complete: the embedding gradient is added to it
    native.embedding_bwd(dx.view(B, T, C), tokens, grad_of(model.wte.weight), grad_of(model.wpe.weight))
    ready(model.wpe.weight, model.wte.weight)
    if deferred and grad_sink is not None:      # both streams have joined: the rest goes out from here
        for p in pending_ready:
            grad_sink.mark_ready(p)
        pending_ready.clear()
    return ws["loss"]
