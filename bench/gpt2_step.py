"""GPT-2 training-step time on one GPU: native hand-scheduled step vs eager PyTorch (bf16, SDPA, fused AdamW).

    python bench/gpt2_step.py [--model small|xl] [--batch 8] [--steps 10]"""
import argparse
import json
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from distributed_training_b200.models.gpt2 import GPT2, GPT2Config  # noqa: E402
from distributed_training_b200.ops import native  # noqa: E402
from distributed_training_b200.optim import FusedAdamW  # noqa: E402


def timed(fn, steps, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(steps):
        fn()
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) / steps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="small")
    ap.add_argument("--batch", type=int, default=8)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--layers", type=int, default=0)
    ap.add_argument("--skip-torch", action="store_true")
    ns = ap.parse_args()
    cfg = GPT2Config.small() if ns.model == "small" else GPT2Config.xl()
    if ns.layers:
        cfg.n_layer = ns.layers
    torch.manual_seed(0)
    model = GPT2(cfg, dtype=torch.bfloat16).cuda()
    tok = torch.randint(0, cfg.vocab_size, (ns.batch, 1024), device="cuda")
    tgt = torch.randint(0, cfg.vocab_size, (ns.batch, 1024), device="cuda")
    opt = FusedAdamW(list(model.parameters()), lr=1e-4, master_weights=False)

    def native_step():
        model.forward_backward(tok, tgt)
        opt.step()

    native.reset_launch_count()
    ms_native = timed(native_step, ns.steps)
    launches = native.launch_count() / (ns.steps + 3)
    out = {"model": ns.model, "layers": cfg.n_layer, "batch": ns.batch, "native_ms": ms_native,
           "native_tokens_per_s": ns.batch * 1024 / ms_native * 1e3, "native_launches_per_step": launches,
           "loss": float(model._ws[next(iter(model._ws))]["loss"])}
    if not ns.skip_torch:
        for p in model.parameters():
            p.grad = None
        topt = torch.optim.AdamW(model.parameters(), lr=1e-4, fused=True)

        def torch_step():
            topt.zero_grad(set_to_none=True)
            model.loss(tok, tgt).backward()
            topt.step()

        ms_torch = timed(torch_step, ns.steps)
        out.update({"torch_ms": ms_torch, "torch_tokens_per_s": ns.batch * 1024 / ms_torch * 1e3,
                    "speedup": ms_torch / ms_native})
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
