"""Attention head-to-head: dtb_attention_fwd/bwd vs torch SDPA on the same box, same shapes (SURVEY W11 names SDPA as
the bar; K16 left "which SDPA backend torch picks on B200" unverified -- this records it).

    python bench/attention_vs_sdpa.py [--out gpurun_out/attention_vs_sdpa.jsonl]

Shapes: GPT-2 small (B8 H12 T1024 hd64) and GPT-2 XL (B4 H25 T1024 hd64), causal, bf16.  CUDA events, 5 warm-up
calls, L2 flushed (256 MB write) before every timed call, median of 20.  Backward for SDPA = autograd backward of
the forward output (the forward is excluded from its timing by re-using a retained graph)."""
from __future__ import annotations

import argparse
import json
import statistics
import sys
from pathlib import Path

import torch
import torch.nn.functional as F

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from distributed_training_b200.ops import native  # noqa: E402

FLUSH = None


def timed(fn, iters=20, warm=5):
    global FLUSH
    if FLUSH is None:
        FLUSH = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(iters):
        FLUSH.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        ts.append(s.elapsed_time(e) * 1e3)
    return statistics.median(ts)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="")
    ns = ap.parse_args()
    from torch.nn.attention import SDPBackend, sdpa_kernel

    rows = []
    for B, H, T in ((8, 12, 1024), (4, 25, 1024)):
        D = 64
        C = H * D
        torch.manual_seed(0)
        qkv = torch.randn(B, T, 3 * C, device="cuda").to(torch.bfloat16)
        dout = torch.randn(B, T, C, device="cuda").to(torch.bfloat16)
        out = torch.empty(B, T, C, device="cuda", dtype=torch.bfloat16)
        lse = torch.empty(B, H, T, device="cuda")
        dqkv = torch.empty_like(qkv)
        scratch = torch.empty(native.attention_bwd_scratch_floats(B, T, H), device="cuda")
        flops_fwd = 4.0 * B * H * T * T * D / 2
        row = {"shape": f"B{B} H{H} T{T} hd64 causal", "flops_fwd": flops_fwd}
        t = timed(lambda: native.attention_fwd(qkv, out, lse, H))
        row["ours_fwd_us"], row["ours_fwd_tflops"] = round(t, 1), round(flops_fwd / t / 1e6, 1)
        t = timed(lambda: native.attention_bwd(qkv, out, dout, lse, dqkv, scratch, H))
        row["ours_bwd_us"], row["ours_bwd_tflops"] = round(t, 1), round(2.5 * flops_fwd / t / 1e6, 1)

        # torch SDPA on the [B, H, T, D] views of the same packed tensor (what baseline/models.TorchGPT2 does)
        q, k, v = (x.detach().requires_grad_(True) for x in
                   qkv.view(B, T, 3, H, D).permute(2, 0, 3, 1, 4).unbind(0))
        g = dout.view(B, T, H, D).transpose(1, 2)
        backends = {"default": None, "flash": SDPBackend.FLASH_ATTENTION, "efficient": SDPBackend.EFFICIENT_ATTENTION,
                    "cudnn": SDPBackend.CUDNN_ATTENTION}
        for name, be in backends.items():
            try:
                def fwd():
                    if be is None:
                        return F.scaled_dot_product_attention(q, k, v, is_causal=True)
                    with sdpa_kernel(be):
                        return F.scaled_dot_product_attention(q, k, v, is_causal=True)

                y = fwd()
                tf = timed(lambda: fwd())
                if be is None:
                    y = F.scaled_dot_product_attention(q, k, v, is_causal=True)
                else:
                    with sdpa_kernel(be):
                        y = F.scaled_dot_product_attention(q, k, v, is_causal=True)
                tb = timed(lambda: torch.autograd.grad(y, (q, k, v), g, retain_graph=True))
                row[f"sdpa_{name}_fwd_us"], row[f"sdpa_{name}_bwd_us"] = round(tf, 1), round(tb, 1)
                if be is None:      # which kernel did torch pick?  (name of the grad_fn identifies the backend)
                    row["sdpa_default_grad_fn"] = type(y.grad_fn).__name__
            except Exception as exc:  # noqa: BLE001
                row[f"sdpa_{name}_error"] = str(exc).splitlines()[0][:120]
        rows.append(row)
        print(json.dumps(row), flush=True)
    if ns.out:
        Path(ns.out).parent.mkdir(parents=True, exist_ok=True)
        with open(ns.out, "a") as fh:
            for r in rows:
                fh.write(json.dumps(r) + "\n")


if __name__ == "__main__":
    main()
