#!/usr/bin/env python
"""Headline benchmark: samples/sec of the wide-MLP (8 x 8192, bf16) DDP training step.

    python bench.py --gpus N --steps K --warmup W [--impl reference] [--batch B]

N > 1 is launched by ``python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...``
(one rank per GPU; RANK / LOCAL_RANK / WORLD_SIZE / MASTER_* from the environment).

Own arm      : distributed_training_b200 ``Trainer`` + ``DDPStrategy`` (peer-memory all-reduce, tcgen05 GEMMs).
Reference arm: the UNMODIFIED reference ``Trainer`` / ``DDPStrategy`` from ``baseline/_ref`` (torch DDP + NCCL
               + cuBLAS) on a plain-torch model of the same architecture.

Both arms time exactly K steps after W warm-up steps with CUDA events, a barrier + synchronize on both
sides, max over ranks, and print ONE JSON line on rank 0.  Weak scaling: the per-GPU batch is fixed.
Data is synthetic, weights are random-init.  The step's working set (1 GiB of bf16 weights + activations)
is far larger than the 126 MB L2, so no explicit flush is needed between iterations.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

WIDTH, DEPTH = 8192, 8
METRIC = "samples/sec"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--batch", type=int, default=4096, help="per-GPU batch (weak scaling)")
    ap.add_argument("--lr", type=float, default=1e-4)
    ap.add_argument("--width", type=int, default=WIDTH)
    ap.add_argument("--depth", type=int, default=DEPTH)
    ap.add_argument("--comm-ctas", type=int, default=0)
    ap.add_argument("--comm-ce", type=int, default=-1, help="1/0: copy-engine path for large all-reduces")
    ap.add_argument("--loader-workers", type=int, default=2, help="ShardedLoader gather threads in the e2e loop")
    ap.add_argument("--comm-exclusive", type=int, default=-1,
                    help="whole-SM comm CTAs for large all-reduces (0 = co-resident mode, -1 = library default)")
    return ap.parse_args()


def dist_env(args):
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world == 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29533")
        os.environ.setdefault("RANK", "0")
        os.environ.setdefault("LOCAL_RANK", "0")
        os.environ.setdefault("WORLD_SIZE", "1")
    if world != args.gpus:
        if rank == 0:
            print(f"bench.py: --gpus {args.gpus} but WORLD_SIZE={world}; launch with torch.distributed.run",
                  file=sys.stderr)
        args.gpus = world
    return rank, local_rank, world


def device_pool(batch, width, device, n=4, seed=0):
    """``n`` device-resident batches for the device-timed loop; seeded per rank so that ranks reduce DIFFERENT
    gradients (the loss after K steps then depends on N)."""
    import torch

    g = torch.Generator().manual_seed(seed)
    pool = []
    for _ in range(n):
        x = torch.randn(batch, width, generator=g).to(torch.bfloat16).to(device)
        y = torch.randint(0, width, (batch,), generator=g, dtype=torch.int64).to(device)
        pool.append((x, y))
    return pool


def result_line(args, world, ms_per_step, e2e_ms_per_step, clocks, launches, impl, h2d, d2h, extra=None):
    global_batch = args.batch * world
    value = global_batch / (ms_per_step / 1e3)
    line = {
        "metric": METRIC, "value": value, "unit": "samples/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic", "impl": impl,
        "config": {"model": f"wide MLP {args.depth}x{args.width} bf16 (Linear+GELU(tanh), CE over {args.width} classes)",
                   "global_batch": global_batch, "per_gpu_batch": args.batch, "seq_len": None,
                   "parallelism": f"dp{world}", "optimizer": "SGD (bf16 params, no momentum)",
                   "l2": "working set (1 GiB weights + activations) >> 126 MB L2; no explicit flush"},
        "clocks": clocks,
        "e2e": {"value": global_batch / (e2e_ms_per_step / 1e3) if e2e_ms_per_step else None, "unit": "samples/s",
                "ms_per_step": e2e_ms_per_step, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
        "gpu_launches": launches,
    }
    if extra:
        line.update(extra)
    return line


# ---------------------------------------------------------------------------------------------
def run_ours(args):
    import torch
    import torch.distributed as dist

    from distributed_training_b200.models.mlp import MLP
    from distributed_training_b200.ops import native
    from distributed_training_b200.parallel.ddp import DDPStrategy
    from distributed_training_b200.parallel.strategy import OptimizerSpec
    from distributed_training_b200.trainer import DistributedEnvironment, Trainer, TrainingConfig
    from distributed_training_b200.utils.clocks import ClockSampler
    from distributed_training_b200.utils.timing import barrier_sync, max_over_ranks, time_steps_ms

    rank, local_rank, world = dist_env(args)
    native.require()
    env = DistributedEnvironment("cuda")
    env.setup()
    device = env.torch_device
    torch.manual_seed(0)
    model = MLP(args.width, args.width, args.width, args.depth, dtype=torch.bfloat16)
    cfg = TrainingConfig(max_epochs=1, save_every=10 ** 9, batch_size=args.batch, learning_rate=args.lr,
                         snapshot_path=Path("/nonexistent/bench_snapshot.pt"), device="cuda",
                         parallel_strategy="ddp", loss="cross_entropy", optimizer="sgd")
    strategy = DDPStrategy(local_rank, True)
    trainer = Trainer(model, [], OptimizerSpec("sgd", lr=args.lr), cfg, env, strategy=strategy)
    if args.comm_ctas and strategy.comm is not None:
        strategy.comm.comm_ctas = args.comm_ctas
    if args.comm_ce >= 0 and strategy.comm is not None and hasattr(strategy.comm, "copy_engine"):
        strategy.comm.copy_engine = bool(args.comm_ce)
    if args.comm_exclusive >= 0 and strategy.comm is not None and hasattr(strategy.comm, "exclusive_sms"):
        strategy.comm.exclusive_sms = args.comm_exclusive
    assert trainer.model.native_step, "the native tcgen05 path must be the one that runs"

    dev_pool = device_pool(args.batch, args.width, device, seed=1000 + rank)

    def step_dev(i):
        x, y = dev_pool[i % len(dev_pool)]
        trainer._run_batch(x, y)

    with ClockSampler(local_rank) as sampler:
        for i in range(args.warmup):
            step_dev(i)
        torch.cuda.synchronize()
        native.reset_launch_count()
        sampler.mark_start()
        total_ms = time_steps_ms(step_dev, args.steps)
        sampler.mark_stop()
    launches = native.launch_count()
    clocks = sampler.summary()
    exposed = trainer.model.exposed_comm_ms() if hasattr(trainer.model, "exposed_comm_ms") else 0.0
    exposed = max_over_ranks(exposed)
    loss_val = float(trainer.last_loss)
    del dev_pool

    # End to end = the body of Trainer._run_epoch over the framework's own data path, the same role the reference
    # arm's loop plays over the reference's DataLoader: a map-style dataset (sample i = row i % 8192 of a pool, as in
    # the reference arm), rank-sharded by ShardedSampler, batches assembled by ShardedLoader (index_select into pinned
    # staging by worker threads -> H2D on a copy stream), then _run_batch and a device->host read of the loss.
    from distributed_training_b200.data import PooledWideDataset
    from distributed_training_b200.trainer import prepare_dataloader

    e2e_warm = 3
    e2e_steps = args.steps + e2e_warm
    dataset = PooledWideDataset(e2e_steps * args.batch * world, args.width, args.width, torch.bfloat16)
    loader = prepare_dataloader(dataset, args.batch, device=device, seed=0, num_workers=args.loader_workers)
    trainer.train_data = loader
    loader.sampler.set_epoch(0)
    timed, h2d = 0, 0
    start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for i, (source, targets) in enumerate(trainer.train_data):
        if i == e2e_warm:
            barrier_sync()
            start.record()
        h2d = source.numel() * source.element_size() + targets.numel() * targets.element_size()
        loss = trainer._run_batch(source, targets)
        loss.item()
        timed += int(i >= e2e_warm)
    stop.record()
    torch.cuda.synchronize()
    e2e_ms = max_over_ranks(start.elapsed_time(stop))
    barrier_sync()
    trainer.check_comm_errors()

    if rank == 0:
        comm = strategy.comm
        line = result_line(args, world, total_ms / args.steps, e2e_ms / max(timed, 1), clocks, launches, "ours",
                           h2d, 4,
                           extra={"exposed_comm_ms_per_step": exposed, "final_loss": loss_val,
                                  "comm": {"heap": getattr(comm, "backend", None),
                                           "nvls": bool(getattr(comm, "nvls_active", False)),
                                           "exclusive_sms": getattr(comm, "exclusive_sms", 0)},
                                  "e2e_path": "PooledWideDataset -> ShardedSampler -> ShardedLoader(pinned, "
                                              f"{args.loader_workers} worker threads) -> Trainer._run_batch -> loss.item()"})
        print(json.dumps(line), flush=True)
    if strategy.comm is not None:
        strategy.comm.close()
    if dist.is_initialized():
        dist.destroy_process_group()


# ---------------------------------------------------------------------------------------------
def run_reference(args):
    from baseline.ref_loader import load_reference, reference_available

    why = reference_available()
    if why:
        print(json.dumps({"impl": "reference", "unavailable": why}), flush=True)
        return
    rank, local_rank, world = dist_env(args)
    import torch
    import torch.distributed as dist

    try:
        ref = load_reference()
    except Exception as exc:  # noqa: BLE001
        if rank == 0:
            print(json.dumps({"impl": "reference", "unavailable": f"import failed: {exc}"[:200]}), flush=True)
        return
    from baseline.models import PooledClassificationDataset, wide_mlp_torch
    from distributed_training_b200.utils.clocks import ClockSampler
    from distributed_training_b200.utils.timing import barrier_sync, max_over_ranks, time_steps_ms

    env = ref.DistributedEnvironment("cuda")
    env.setup()                                           # init_process_group("nccl") + set_device
    torch.manual_seed(0)
    model = wide_mlp_torch(args.width, args.depth, torch.bfloat16)
    optimizer = torch.optim.SGD(model.parameters(), lr=args.lr)
    cfg = ref.TrainingConfig(max_epochs=1, save_every=10 ** 9, batch_size=args.batch, learning_rate=args.lr,
                             snapshot_path=Path("/nonexistent/bench_snapshot.pt"), device="cuda",
                             parallel_strategy="ddp")
    e2e_steps = args.steps + 2
    dataset = PooledClassificationDataset(e2e_steps * args.batch * world, args.width, args.width, torch.bfloat16)
    loader = ref.prepare_dataloader(dataset, args.batch)   # DataLoader(pin_memory=True, DistributedSampler)
    trainer = ref.Trainer(model, loader, optimizer, cfg, env)
    device = torch.device("cuda", local_rank)

    g = torch.Generator().manual_seed(0)
    dev_pool = [(torch.randn(args.batch, args.width, generator=g).to(torch.bfloat16).to(device),
                 torch.randint(0, args.width, (args.batch,), generator=g).to(device)) for _ in range(4)]

    def step_dev(i):
        x, y = dev_pool[i % len(dev_pool)]
        trainer._run_batch(x, y)                           # the reference's own step

    with ClockSampler(local_rank) as sampler:
        for i in range(args.warmup):
            step_dev(i)
        torch.cuda.synchronize()
        sampler.mark_start()
        total_ms = time_steps_ms(step_dev, args.steps)
        sampler.mark_stop()
    clocks = sampler.summary()

    # end to end = the body of the reference's Trainer._run_epoch on its own DataLoader, plus a 4-byte
    # read-back of the step's result (the reference never exposes its loss, so a parameter element stands in)
    probe = next(trainer.model.parameters())
    loader.sampler.set_epoch(0)
    it = iter(loader)
    timed = 0
    e2e_ms = None
    bytes_in = 0
    for i in range(e2e_steps):
        source, targets = next(it)
        if i == 2:
            barrier_sync()
            start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            start.record()
        bytes_in = source.numel() * source.element_size() + targets.numel() * targets.element_size()
        source = source.to(local_rank)
        targets = targets.to(local_rank)
        trainer._run_batch(source, targets)
        probe.view(-1)[0].item()
        timed += int(i >= 2)
    stop.record()
    torch.cuda.synchronize()
    e2e_ms = max_over_ranks(start.elapsed_time(stop))
    barrier_sync()

    if rank == 0:
        line = result_line(args, world, total_ms / args.steps, e2e_ms / max(timed, 1), clocks, 0, "reference",
                           bytes_in, probe.element_size(),
                           extra={"gpu_launches_note": "reference arm: torch/cuBLAS/NCCL kernels only"})
        print(json.dumps(line), flush=True)
    if dist.is_initialized():
        dist.destroy_process_group()


if __name__ == "__main__":
    ns = parse_args()
    if ns.impl == "reference":
        run_reference(ns)
    else:
        run_ours(ns)
