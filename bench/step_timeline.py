"""Per-kernel device timeline of the headline training step (wide MLP, DDP), to see which kernels stretch when
the gradient reduction runs beside the GEMMs.  Launch like bench.py (plain python for 1 GPU, torchrun for N).

    python bench/step_timeline.py --out gpurun_out/timeline_n1.json
    torchrun --nproc-per-node 2 ... bench/step_timeline.py --out gpurun_out/timeline_n2.json [--no-sync]
"""
from __future__ import annotations

import argparse
import contextlib
import json
import os
import sys
from collections import defaultdict
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--batch", type=int, default=4096)
    ap.add_argument("--width", type=int, default=8192)
    ap.add_argument("--depth", type=int, default=8)
    ap.add_argument("--comm-ctas", type=int, default=0)
    ap.add_argument("--comm-ce", type=int, default=-1, help="1/0: copy-engine path for large all-reduces")
    ap.add_argument("--comm-exclusive", type=int, default=-1)
    ap.add_argument("--no-sync", action="store_true", help="skip the gradient reduction (compute-only step)")
    ap.add_argument("--out", default="gpurun_out/timeline.json")
    args = ap.parse_args()

    import torch
    import torch.distributed as dist

    from distributed_training_b200.models.mlp import MLP
    from distributed_training_b200.ops import native
    from distributed_training_b200.parallel.ddp import DDPStrategy
    from distributed_training_b200.parallel.strategy import OptimizerSpec
    from distributed_training_b200.trainer import DistributedEnvironment, Trainer, TrainingConfig
    from distributed_training_b200.utils.clocks import ClockSampler
    from distributed_training_b200.utils.timing import max_over_ranks, time_steps_ms
    from distributed_training_b200.utils.tracing import KernelTimeline

    for key, val in (("MASTER_ADDR", "127.0.0.1"), ("MASTER_PORT", "29533"), ("RANK", "0"), ("LOCAL_RANK", "0"),
                     ("WORLD_SIZE", "1")):
        os.environ.setdefault(key, val)
    rank, local_rank = int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0))
    native.require()
    env = DistributedEnvironment("cuda")
    env.setup()
    device = env.torch_device
    torch.manual_seed(0)
    model = MLP(args.width, args.width, args.width, args.depth, dtype=torch.bfloat16)
    cfg = TrainingConfig(max_epochs=1, save_every=10 ** 9, batch_size=args.batch, learning_rate=1e-4,
                         snapshot_path=Path("/nonexistent/x.pt"), device="cuda", parallel_strategy="ddp",
                         loss="cross_entropy", optimizer="sgd")
    strategy = DDPStrategy(local_rank, True)
    trainer = Trainer(model, [], OptimizerSpec("sgd", lr=1e-4), cfg, env, strategy=strategy)
    if args.comm_ctas and strategy.comm is not None:
        strategy.comm.comm_ctas = args.comm_ctas
    if args.comm_ce >= 0 and strategy.comm is not None and hasattr(strategy.comm, "copy_engine"):
        strategy.comm.copy_engine = bool(args.comm_ce)
    if args.comm_exclusive >= 0 and strategy.comm is not None and hasattr(strategy.comm, "exclusive_sms"):
        strategy.comm.exclusive_sms = args.comm_exclusive
    x = torch.randn(args.batch, args.width, device=device).to(torch.bfloat16)
    y = torch.randint(0, args.width, (args.batch,), device=device)

    def ctx():
        if args.no_sync and hasattr(trainer.model, "no_sync"):
            return trainer.model.no_sync()
        return contextlib.nullcontext()

    def step(_i):
        with ctx():
            trainer._run_batch(x, y)

    with ClockSampler(local_rank) as sampler:
        for i in range(args.warmup):
            step(i)
        torch.cuda.synchronize()
        sampler.mark_start()
        total_ms = time_steps_ms(step, args.steps)
        sampler.mark_stop()
    clocks = sampler.summary()
    with KernelTimeline() as tl:
        for i in range(3):
            torch.cuda.synchronize()
            mark = len(tl._records)
            step(i)
        rows = tl.rows()[mark:]
    base = rows[0]["start_ms"] if rows else 0.0
    for r in rows:
        r["start_ms"] = round(r["start_ms"] - base, 4)
    by = defaultdict(lambda: [0, 0.0])
    for r in rows:
        key = f"{r['name']} {r['detail']} s{r['stream']}"
        by[key][0] += 1
        by[key][1] += r["ms"]
    span = max(r["start_ms"] + r["ms"] for r in rows) if rows else 0.0
    summary = {"world": env.world_size, "no_sync": args.no_sync, "ms_per_step": max_over_ranks(total_ms / args.steps),
               "traced_step_span_ms": round(span, 3), "clocks": clocks,
               "by_kernel": {k: {"n": v[0], "ms_total": round(v[1], 3), "ms_avg": round(v[1] / v[0], 4)}
                             for k, v in sorted(by.items(), key=lambda kv: -kv[1][1])}}
    if rank == 0:
        Path(args.out).parent.mkdir(parents=True, exist_ok=True)
        Path(args.out).write_text(json.dumps({"summary": summary, "rows": rows}, indent=1))
        print(json.dumps(summary))
    if strategy.comm is not None:
        strategy.comm.close()
    if dist.is_initialized():
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
