"""BASELINE.json configs 2-4 on N GPUs: this framework vs the same torch build the reference runs on
(torch DDP / FSDP1 + NCCL + cuBLAS + SDPA, eager, bf16), device-timed, max over ranks.

    python -m torch.distributed.run --nproc-per-node N bench/configs.py --config gpt2_small --strategy ddp --impl ours
"""
import argparse
import json
import os
import sys
from pathlib import Path

import torch
import torch.distributed as dist

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

CONFIGS = {"wide_mlp": dict(batch=4096), "gpt2_small": dict(batch=8, n_embd=768, n_layer=12, n_head=12),
           "gpt2_xl": dict(batch=4, n_embd=1600, n_layer=48, n_head=25)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="gpt2_small", choices=list(CONFIGS))
    ap.add_argument("--strategy", default="ddp", choices=["ddp", "fsdp"])
    ap.add_argument("--impl", default="ours", choices=["ours", "torch"])
    ap.add_argument("--optimizer", default=None)
    ap.add_argument("--batch", type=int, default=0)
    ap.add_argument("--layers", type=int, default=0)
    ap.add_argument("--ref-loss", action="store_true",
                    help="torch arm: F.cross_entropy(output[B,V,T], targets) exactly as the reference Trainer calls it "
                         "(default: flat fp32 logits, the layout the product model uses)")
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--timeline", default="", help="ours: write a per-kernel device timeline of one step to this file")
    ns = ap.parse_args()
    c = dict(CONFIGS[ns.config])
    if ns.batch:
        c["batch"] = ns.batch
    if ns.layers:
        c["n_layer"] = ns.layers
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29577")
    os.environ.setdefault("RANK", "0")
    os.environ.setdefault("WORLD_SIZE", "1")
    os.environ.setdefault("LOCAL_RANK", "0")
    torch.cuda.set_device(local)
    opt_name = ns.optimizer or ("sgd" if ns.config == "wide_mlp" else "adamw")
    from distributed_training_b200.utils.timing import time_steps_ms

    torch.manual_seed(0)
    is_mlp = ns.config == "wide_mlp"
    if is_mlp:
        x = torch.randn(c["batch"], 8192, device="cuda").to(torch.bfloat16)
        y = torch.randint(0, 8192, (c["batch"],), device="cuda")
    else:
        x = torch.randint(0, 50257, (c["batch"], 1024), device="cuda")
        y = torch.randint(0, 50257, (c["batch"], 1024), device="cuda")

    if ns.impl == "ours":
        dist.init_process_group("gloo")
        from distributed_training_b200.models.gpt2 import GPT2, GPT2Config
        from distributed_training_b200.models.mlp import MLP
        from distributed_training_b200.ops import native
        from distributed_training_b200.parallel.ddp import DDPStrategy
        from distributed_training_b200.parallel.sharded import FSDPStrategy
        from distributed_training_b200.parallel.strategy import OptimizerSpec

        model = (MLP(8192, 8192, 8192, 8, dtype=torch.bfloat16) if is_mlp else
                 GPT2(GPT2Config(n_embd=c["n_embd"], n_layer=c["n_layer"], n_head=c["n_head"]),
                      dtype=torch.bfloat16)).cuda()
        strategy = DDPStrategy(local, True) if ns.strategy == "ddp" else FSDPStrategy(True)
        wrapped = strategy.prepare_model(model, local)
        spec = OptimizerSpec(opt_name, lr=1e-4, weight_decay=0.0 if opt_name == "sgd" else 0.1)
        opt = strategy.build_optimizer(wrapped, spec)
        if hasattr(wrapped, "attach_optimizer"):
            wrapped.attach_optimizer(opt)
        assert wrapped.native_step

        def step(i):
            wrapped.forward_backward(x, y)
            opt.step()

        for i in range(ns.warmup):
            step(i)
        native.reset_launch_count()
        total = time_steps_ms(step, ns.steps)
        extra = {"launches_per_step": native.launch_count() / ns.steps,
                 "exposed_comm_ms": wrapped.exposed_comm_ms() if hasattr(wrapped, "exposed_comm_ms") else None}
        # host time to ISSUE a step (no sync inside): when it approaches the device time the step is launch-bound
        import time
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for i in range(3):
            step(i)
        extra["host_issue_ms_per_step"] = (time.perf_counter() - t0) / 3 * 1e3
        torch.cuda.synchronize()
        if ns.timeline:
            from collections import defaultdict
            from distributed_training_b200.utils.tracing import KernelTimeline
            with KernelTimeline() as tl:
                for i in range(2):
                    torch.cuda.synchronize()
                    mark = len(tl._records)
                    step(i)
                rows = tl.rows()[mark:]
            base = rows[0]["start_ms"]
            by = defaultdict(lambda: [0, 0.0])
            for r in rows:
                r["start_ms"] = round(r["start_ms"] - base, 4)
                key = f"{r['name']} {r['detail']} s{r['stream']}"
                by[key][0] += 1
                by[key][1] += r["ms"]
            busy = {}
            for r in rows:
                busy[r["stream"]] = busy.get(r["stream"], 0.0) + r["ms"]
            if rank == 0:
                Path(ns.timeline).parent.mkdir(parents=True, exist_ok=True)
                Path(ns.timeline).write_text(json.dumps({
                    "span_ms": max(r["start_ms"] + r["ms"] for r in rows), "busy_ms_by_stream": busy,
                    "by_kernel": {k: {"n": v[0], "ms_total": round(v[1], 3)} for k, v in
                                  sorted(by.items(), key=lambda kv: -kv[1][1])}, "rows": rows}, indent=1))
        strategy.comm.close()
    else:
        dist.init_process_group("cpu:gloo,cuda:nccl")
        from baseline.models import TorchGPT2, wide_mlp_torch
        from torch.distributed.fsdp import FullyShardedDataParallel as FSDP
        from torch.nn.parallel import DistributedDataParallel as DDP

        model = (wide_mlp_torch() if is_mlp else
                 TorchGPT2(c["n_embd"], c["n_layer"], c["n_head"])).cuda()
        # exactly what the reference's strategies do (ddp_strategy.py:18, fsdp_strategy.py:20-26)
        wrapped = DDP(model, device_ids=[local]) if ns.strategy == "ddp" else FSDP(model, device_id=local)
        params = wrapped.parameters()
        opt = (torch.optim.SGD(params, lr=1e-4) if opt_name == "sgd"
               else torch.optim.AdamW(params, lr=1e-4, weight_decay=0.1))

        def step(i):
            opt.zero_grad()
            out = wrapped(x)
            if is_mlp or ns.ref_loss:
                loss = torch.nn.functional.cross_entropy(out, y)
            else:       # [B, V, T] -> [B*T, V] fp32
                flat = out.transpose(1, 2).float().reshape(-1, out.shape[1])
                loss = torch.nn.functional.cross_entropy(flat, y.view(-1))
            loss.backward()
            opt.step()

        for i in range(ns.warmup):
            step(i)
        total = time_steps_ms(step, ns.steps)
        extra = {}
    ms = total / ns.steps
    if rank == 0:
        unit = c["batch"] * world
        print(json.dumps({"config": ns.config, "strategy": ns.strategy, "impl": ns.impl, "optimizer": opt_name,
                          "n_gpus": world, "per_gpu_batch": c["batch"], "ms_per_step": ms,
                          "samples_per_s": unit / ms * 1e3,
                          "tokens_per_s": None if is_mlp else unit * 1024 / ms * 1e3,
                          "peak_mem_gb": torch.cuda.max_memory_allocated() / 2**30, **extra}), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
