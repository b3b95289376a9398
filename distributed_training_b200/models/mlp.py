"""MLP family: the toy regression MLP (BASELINE.json config 1) and the wide
8 x 8192 bf16 MLP (config 2).

The reference's only model is ``torch.nn.Linear(20, 1)``
(/root/reference/src/distributed_trainer.py:199); these are the north-star
extensions of the same idea.  ``forward`` is ordinary autograd-visible PyTorch
(CPU tests, arbitrary wrappers).  ``forward_backward`` is the B200 path: every
Linear forward / dgrad / wgrad is one ``tcgen05`` GEMM with its bias, GELU or
GELU' fused into the epilogue, the loss is a fused softmax-CE kernel, gradients
are written straight into the caller's gradient buckets and the bucket is
handed to the reducer the moment a layer's wgrad has been queued.
"""
from __future__ import annotations

import os
from typing import Any, Dict, List, Optional, Tuple

import torch
import torch.nn.functional as F
from torch import nn

from ..ops import native

_FUSED_COLSUM = os.environ.get("DTB_FUSED_COLSUM", "1") != "0"
#: weight gradients on a side stream so that consecutive GEMMs fill each other's last wave
#: (DTB_ASYNC_WGRAD=0: one stream)
_ASYNC_WGRAD = os.environ.get("DTB_ASYNC_WGRAD", "1") != "0"


class MLP(nn.Module):
    """``num_layers`` Linear layers with tanh-GELU between them."""

    supports_sharded_native = True

    def __init__(self, input_size: int, hidden_size: int, output_size: int, num_layers: int = 2,
                 dtype: torch.dtype = torch.float32, bias: bool = True) -> None:
        super().__init__()
        if num_layers < 1:
            raise ValueError("num_layers must be >= 1")
        dims = [input_size] + [hidden_size] * (num_layers - 1) + [output_size]
        self.layers = nn.ModuleList(
            [nn.Linear(dims[i], dims[i + 1], bias=bias, dtype=dtype) for i in range(num_layers)])
        self.dims = dims
        self._ws: Dict[Tuple[int, torch.device], Dict[str, Any]] = {}

    def fsdp_units(self) -> List[nn.Module]:
        return list(self.layers)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        h = x
        last = len(self.layers) - 1
        for i, layer in enumerate(self.layers):
            h = layer(h)
            if i != last:
                h = F.gelu(h, approximate="tanh")
        return h

    # ------------------------------------------------------------------ native
    def native_ready(self) -> bool:
        """The tcgen05 path needs CUDA bf16 parameters and 8-element-aligned widths."""
        p = self.layers[0].weight
        return (p.is_cuda and p.dtype == torch.bfloat16 and native.available()
                and all(d % 8 == 0 for d in self.dims))

    def _workspace(self, rows: int, device: torch.device) -> Dict[str, Any]:
        key = (rows, device)
        ws = self._ws.get(key)
        if ws is None:
            bf = dict(dtype=torch.bfloat16, device=device)
            n = len(self.layers)
            ws = {
                "z": [torch.empty(rows, self.dims[i + 1], **bf) for i in range(n - 1)],   # pre-activations
                "h": [torch.empty(rows, self.dims[i + 1], **bf) for i in range(n - 1)],   # activations
                "logits": torch.empty(rows, self.dims[-1], **bf),
                # output gradients: D_i writes slot i % 3 while W_{i+1} (side stream) may still read slot (i+2) % 3
                "dz": [torch.empty(rows, max(self.dims[1:-1] or [8]), **bf) for _ in range(3 if n > 2 else 2)],
                "row_loss": torch.empty(rows, dtype=torch.float32, device=device),
                "loss": torch.empty((), dtype=torch.float32, device=device),
            }
            # fp32 bias-gradient accumulators the dgrad epilogues add into: views of one buffer that is cleared with
            # one launch at the start of every backward (a step that died half-way must not leave sums behind)
            widths = [(self.dims[i + 1] + 7) // 8 * 8 for i in range(n - 1)]
            ws["bsum_flat"] = torch.zeros(max(sum(widths), 8), dtype=torch.float32, device=device)
            ws["bsum"], at = [], 0
            for i, w in enumerate(widths):
                ws["bsum"].append(ws["bsum_flat"][at:at + self.dims[i + 1]])
                at += w
            if device.type == "cuda":
                ws["ev_d"] = [torch.cuda.Event() for _ in range(n)]      # dz_i complete (dgrad stream)
                ws["ev_w"] = [torch.cuda.Event() for _ in range(n)]      # layer i's gradients queued (wgrad stream)
            self._ws = {key: ws}      # one live batch shape at a time
        return ws

    def _wgrad_stream(self, device: torch.device) -> "torch.cuda.Stream":
        st = getattr(self, "_side_stream", None)
        if st is None or st.device != device:
            st = torch.cuda.Stream(device)
            self._side_stream = st
        return st

    def _fuse_bias_grad(self, i: int, k_of_dgrad: int) -> bool:
        """Layer ``i``'s bias gradient comes out of the epilogue of the dgrad GEMM that produces its output gradient
        (``dtb_gemm_args.colsum``) when that GEMM's K is long enough to hide the extra epilogue work and the
        gradient is bf16; DTB_FUSED_COLSUM=0 keeps the separate column-sum kernel."""
        bias = self.layers[i].bias
        return (_FUSED_COLSUM and bias is not None and bias.dtype == torch.bfloat16 and k_of_dgrad >= 2048
                and self.dims[i + 1] % 8 == 0)

    #: forward_backward accepts ``bwd_max_ctas`` (see ReplicatedModule: exclusive-mode reduction kernels)
    supports_bwd_max_ctas = True

    def forward_backward(self, inputs: torch.Tensor, targets: torch.Tensor, loss: str = "cross_entropy",
                         grad_sink: Any = None, max_ctas: int = 0,
                         fused_sgd: Optional[Tuple[float, float]] = None, on_updated: Any = None,
                         bwd_max_ctas: int = 0) -> torch.Tensor:
        """Fused training step (no autograd).  Returns the mean loss (0-dim fp32 tensor).

        ``grad_sink`` (a ``ReplicatedModule``) supplies ``grad_view(param)`` and
        ``mark_ready(param)``; without one, gradients land in ``param.grad``.
        ``fused_sgd=(lr, weight_decay)`` makes the wgrad GEMM epilogue apply the SGD update to the
        weight in place (single replica only); ``on_updated(params)`` is told which parameters were
        stepped that way so the optimizer skips them.
        ``bwd_max_ctas`` caps the GEMM grids of backward once the first gradient has been handed to the reducer:
        the data-parallel wrapper uses it to leave whole SMs to its exclusive-mode reduction kernels.
        """
        if loss != "cross_entropy" or targets.dtype != torch.int64:
            raise NotImplementedError("native MLP step supports cross_entropy with int64 class targets")
        x = inputs.contiguous()
        rows = x.shape[0]
        ws = self._workspace(rows, x.device)
        n = len(self.layers)

        def grad_of(p: nn.Parameter) -> torch.Tensor:
            if grad_sink is not None:
                return grad_sink.grad_view(p)
            if p.grad is None:
                p.grad = torch.empty_like(p)
            return p.grad

        # a sharded grad_sink gathers the weight inside the forward GEMM (all-gather -> GEMM fusion)
        gemm_fwd = getattr(grad_sink, "gemm_forward", None) or native.gemm
        wv = getattr(grad_sink, "weight_view", None) or (lambda p: p)     # sharded: transiently re-gathered matrix
        # ---- forward: h_{i+1} = gelu(h_i W_i^T + b_i), pre-activation kept for backward
        h = x
        for i, layer in enumerate(self.layers):
            out_f, in_f = layer.weight.shape
            if i != n - 1:
                gemm_fwd(h, layer.weight, ws["h"][i], m=rows, n=out_f, k=in_f, bias=layer.bias, act=1,
                         d2=ws["z"][i], max_ctas=max_ctas)
                h = ws["h"][i]
            else:
                gemm_fwd(h, layer.weight, ws["logits"], m=rows, n=out_f, k=in_f, bias=layer.bias,
                         max_ctas=max_ctas)
        # ---- loss + dlogits in place
        native.cross_entropy_fwd_bwd(ws["logits"], targets, ws["row_loss"], self.dims[-1], 1.0 / rows)
        native.mean_f32(ws["row_loss"], ws["loss"])
        # ---- backward, last layer first; each layer's grads go to the reducer as soon as queued.
        # Two streams: the dgrad chain (D_i: dz_i -> dz_{i-1}) stays on the caller's stream, the weight gradients
        # (W_i: needs dz_i only) go to a side stream.  A persistent GEMM ends with a partly filled wave -- 512 tiles
        # over the 72 CTA pairs that are left while 4 SMs carry the gradient reduction are 7.1 waves, i.e. 8 -- and
        # on ONE stream the next kernel cannot start before the last CTA has exited.  With two independent GEMMs
        # queued the SMs a finishing kernel frees are taken by the other one's CTAs (the tile scheduler is a global
        # counter per stream, so late CTAs simply join).
        # Modes (the grad sink decides, ``ReplicatedModule.async_wgrad_mode``): "full" = the two chains never join (one
        # replica); None = one stream (several replicas: a second queued GEMM takes the SMs that are left to the
        # whole-SM reduction CTAs, measured slower); "forkjoin" = W_i beside D_i with a join per layer (opt-in, also
        # measured slower for the same reason).
        mode = None
        if _ASYNC_WGRAD and x.is_cuda:
            mode = "full" if grad_sink is None else getattr(grad_sink, "async_wgrad_mode", lambda: None)()
        async_w = mode is not None
        forkjoin = mode == "forkjoin"
        tail_layers = getattr(grad_sink, "TAIL_LAYERS", 1)
        main = torch.cuda.current_stream() if async_w else None
        side = self._wgrad_stream(x.device) if async_w else None
        nbuf = len(ws["dz"]) if async_w else 2
        dz = ws["logits"]
        cap = max_ctas
        if n > 1 and any(self._fuse_bias_grad(i, self.dims[i + 2]) for i in range(n - 1)):
            ws["bsum_flat"].zero_()
        ev_dz = None            # dz_i (and layer i's bias-gradient sums) are complete on the dgrad stream
        if async_w:
            ev_dz = ws["ev_d"][n - 1]
            ev_dz.record(main)
        for i in range(n - 1, -1, -1):
            layer = self.layers[i]
            out_f, in_f = layer.weight.shape
            h_in = x if i == 0 else ws["h"][i - 1]
            nxt = None
            ev_next = None
            if i > 0:
                # dgrad fused with GELU': dz_prev = (dz W_i) * gelu'(z_{i-1}).  It runs BEFORE this layer's
                # wgrad is handed to the reducer: a fused reduce+SGD (or the fused epilogue below) overwrites W_i.
                # The same epilogue adds dz_prev's column sums (= the bias gradient of layer i-1) to an fp32
                # accumulator: no separate pass over dz_prev.
                nxt = ws["dz"][i % nbuf]
                if async_w and i + nbuf - 1 <= n - 1:
                    main.wait_event(ws["ev_w"][i + nbuf - 1])      # the wgrad that read this buffer's last content
                fuse_bias = self._fuse_bias_grad(i - 1, out_f)
                native.gemm(dz, wv(layer.weight), nxt, m=rows, n=in_f, k=out_f, b_mn=True, lda=dz.shape[1], ldb=in_f,
                            ldd=nxt.shape[1], aux=ws["z"][i - 1], ld_aux=in_f, act=2, max_ctas=cap,
                            colsum=ws["bsum"][i - 1] if fuse_bias else None)
                if async_w:
                    ev_next = ws["ev_d"][i - 1]
                    ev_next.record(main)
            if async_w:
                side.wait_event(ev_dz)
                if ev_next is not None and (fused_sgd is not None or (forkjoin and i < tail_layers)):
                    # fused SGD: the epilogue below rewrites W_i, which D_i is reading.  Tail of a multi-replica
                    # step: the pieces of W_i and their reductions must not queue behind a full-size D_i
                    side.wait_event(ev_next)
                torch.cuda.set_stream(side)
            try:
                if fused_sgd is not None:
                    # single replica: W = (1 - lr*wd) W - lr * (dz^T h_in), applied in the wgrad epilogue (no dW in HBM)
                    lr, wd = fused_sgd
                    native.gemm(dz, h_in, layer.weight.data, m=out_f, n=in_f, k=rows, a_mn=True, b_mn=True,
                                lda=dz.shape[1], ldb=in_f, aux=layer.weight.data, alpha=-lr, beta=1.0 - lr * wd,
                                max_ctas=cap)
                    if on_updated is not None:
                        on_updated([layer.weight])
                elif i < tail_layers and getattr(grad_sink, "tail_chunks", lambda p: 1)(layer.weight) > 1:
                    # the last wgrads of backward: produce (and reduce) them in row chunks so that only the final
                    # chunk's reduction is exposed; the reductions get more SMs here and the GEMM pieces leave them
                    nch = grad_sink.tail_chunks(layer.weight)
                    gw, per = grad_of(layer.weight), out_f // nch
                    if hasattr(grad_sink, "begin_tail"):
                        grad_sink.begin_tail()
                        cap = grad_sink.tail_gemm_ctas(cap)
                    for c in range(nch):
                        native.gemm(dz[:, c * per:], h_in, gw[c * per:], m=per, n=in_f, k=rows, a_mn=True, b_mn=True,
                                    lda=dz.shape[1], ldb=in_f, max_ctas=cap)
                        grad_sink.mark_ready_chunk(layer.weight, c, nch)
                else:
                    # wgrad: dW[out,in] = dz^T h_in   (both operands MN-major), straight into the gradient bucket
                    native.gemm(dz, h_in, grad_of(layer.weight), m=out_f, n=in_f, k=rows, a_mn=True, b_mn=True,
                                lda=dz.shape[1], ldb=in_f, max_ctas=cap)
                if layer.bias is not None:
                    if i < n - 1 and self._fuse_bias_grad(i, self.dims[i + 2]):
                        gb = grad_of(layer.bias)
                        if gb.dtype == torch.bfloat16:
                            native.take_f32_to_bf16(ws["bsum"][i], gb)
                        else:
                            gb.copy_(ws["bsum"][i])
                            ws["bsum"][i].zero_()
                    else:
                        native.colsum_split(dz[:, :out_f], grad_of(layer.bias), ldx=dz.shape[1])
                if grad_sink is not None:
                    if async_w and ev_next is not None:
                        # the reducer may rewrite W_i (fused reduce + SGD): D_i must be done
                        side.wait_event(ev_next)
                    grad_sink.mark_ready(layer.weight)
                    if layer.bias is not None:
                        grad_sink.mark_ready(layer.bias)
                    if bwd_max_ctas:
                        cap = bwd_max_ctas
                if async_w:
                    ws["ev_w"][i].record(side)
            finally:
                if async_w:
                    torch.cuda.set_stream(main)
            if forkjoin:
                main.wait_event(ws["ev_w"][i])
            if nxt is not None:
                dz = nxt
                ev_dz = ev_next
        if async_w:
            main.wait_stream(side)
        return ws["loss"]


def wide_mlp(dtype: torch.dtype = torch.bfloat16, width: int = 8192, depth: int = 8) -> MLP:
    """BASELINE.json config 2."""
    return MLP(width, width, width, depth, dtype=dtype)
