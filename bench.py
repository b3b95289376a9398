#!/usr/bin/env python
"""Headline benchmark: samples/sec of a BASELINE.json training config (default: wide MLP 8 x 8192 bf16, DDP).

    python bench.py --gpus N --steps K --warmup W [--impl reference] [--config wide_mlp|gpt2_small|gpt2_xl_fsdp]

N > 1 is launched by ``python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...``
(one rank per GPU; RANK / LOCAL_RANK / WORLD_SIZE / MASTER_* from the environment).

Own arm      : distributed_training_b200 ``Trainer`` + ``DDPStrategy`` / ``FSDPStrategy`` (peer-memory / NVLS
               collectives, tcgen05 GEMMs and attention, fused optimizer paths).
Reference arm: the UNMODIFIED reference ``Trainer`` with its own ``DDPStrategy`` / ``FSDPStrategy`` from
               ``baseline/_ref`` (torch DDP / FSDP1 + NCCL + cuBLAS + SDPA) on a plain-torch model of the same
               architecture (``baseline/models.py``).

Both arms time exactly K steps after W warm-up steps with CUDA events, a barrier + synchronize on both
sides, max over ranks, and print ONE JSON line on rank 0.  Weak scaling: the per-GPU batch is fixed.
Data is synthetic, weights are random-init.  Every config's working set (>= 0.25 GB of weights plus
activations) is far larger than the 126 MB L2, so no explicit flush is needed between iterations.

The ``e2e`` number walks each framework's own data path every step: map-style dataset -> rank-sharding sampler ->
the framework's loader (ours: ``ShardedLoader``, pinned staging + worker threads + copy stream; reference:
``DataLoader(pin_memory=True)`` + blocking ``.to``) -> the trainer's ``_run_batch`` -> a device->host read of the
step's result.
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
VOCAB = 50257

CONFIGS = {
    # BASELINE.json config 2 (the headline)
    "wide_mlp": dict(kind="mlp", batch=4096, strategy="ddp", optimizer="sgd", seq_len=None),
    # config 3
    "gpt2_small": dict(kind="gpt2", n_embd=768, n_layer=12, n_head=12, batch=8, seq_len=1024, strategy="ddp",
                       optimizer="adamw"),
    # config 4
    "gpt2_xl_fsdp": dict(kind="gpt2", n_embd=1600, n_layer=48, n_head=25, batch=4, seq_len=1024, strategy="fsdp",
                         optimizer="adamw"),
}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--config", choices=list(CONFIGS), default="wide_mlp")
    ap.add_argument("--batch", type=int, default=0, help="per-GPU batch (weak scaling); 0 = the config's")
    ap.add_argument("--layers", type=int, default=0, help="gpt2 configs: override n_layer (smoke runs only)")
    ap.add_argument("--lr", type=float, default=0.0,
                    help="0 = the config's (SGD 0.05: large enough that bf16 weights move every step, so the loss "
                         "after K steps depends on the gradients every rank contributed; AdamW 1e-4)")
    ap.add_argument("--width", type=int, default=WIDTH)
    ap.add_argument("--depth", type=int, default=DEPTH)
    ap.add_argument("--comm-ctas", type=int, default=0)
    ap.add_argument("--comm-ce", type=int, default=-1, help="1/0: copy-engine path for large all-reduces")
    ap.add_argument("--loader-workers", type=int, default=4, help="ShardedLoader gather threads in the e2e loop")
    ap.add_argument("--comm-exclusive", type=int, default=-1,
                    help="whole-SM comm CTAs for large all-reduces (0 = co-resident mode, -1 = library default)")
    args = ap.parse_args()
    args.cfg = dict(CONFIGS[args.config])
    if args.batch:
        args.cfg["batch"] = args.batch
    args.batch = args.cfg["batch"]
    if not args.lr:
        args.lr = 0.05 if args.cfg["optimizer"] == "sgd" else 1e-4
    if args.layers and args.cfg["kind"] == "gpt2":
        args.cfg["n_layer"] = args.layers
    return args


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


def device_pool(args, device, n=4, seed=0):
    """``n`` device-resident batches for the device-timed loop; seeded per rank so that ranks reduce DIFFERENT
    gradients (the loss after K steps then depends on N)."""
    import torch

    g = torch.Generator().manual_seed(seed)
    pool = []
    for _ in range(n):
        if args.cfg["kind"] == "mlp":
            x = torch.randn(args.batch, args.width, generator=g).to(torch.bfloat16).to(device)
            y = torch.randint(0, args.width, (args.batch,), generator=g, dtype=torch.int64).to(device)
        else:
            t = torch.randint(0, VOCAB, (args.batch, args.cfg["seq_len"] + 1), generator=g, dtype=torch.int64)
            x, y = t[:, :-1].contiguous().to(device), t[:, 1:].contiguous().to(device)
        pool.append((x, y))
    return pool


def model_text(args):
    c = args.cfg
    if c["kind"] == "mlp":
        return f"wide MLP {args.depth}x{args.width} bf16 (Linear+GELU(tanh), CE over {args.width} classes)"
    return (f"GPT-2 n_embd={c['n_embd']} n_layer={c['n_layer']} n_head={c['n_head']} bf16, vocab 50257 (padded 50304), "
            f"tied LM head, next-token CE")


def result_line(args, world, ms_per_step, e2e_ms_per_step, clocks, launches, impl, h2d, d2h, extra=None):
    c = args.cfg
    global_batch = args.batch * world
    value = global_batch / (ms_per_step / 1e3)
    par = ("dp" if c["strategy"] == "ddp" else "fsdp") + str(world)
    opt = "SGD (bf16 params, no momentum)" if c["optimizer"] == "sgd" else "AdamW (fp32 master + moments)"
    line = {
        "metric": METRIC, "value": value, "unit": "samples/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic", "impl": impl,
        "config": {"name": args.config, "model": model_text(args),
                   "global_batch": global_batch, "per_gpu_batch": args.batch, "seq_len": c["seq_len"],
                   "parallelism": par, "optimizer": opt,
                   "l2": "working set (weights + activations) >> 126 MB L2; no explicit flush"},
        "clocks": clocks,
        "e2e": {"value": global_batch / (e2e_ms_per_step / 1e3) if e2e_ms_per_step else None, "unit": "samples/s",
                "ms_per_step": e2e_ms_per_step, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
        "gpu_launches": launches,
    }
    if c["seq_len"]:
        line["tokens_per_s"] = value * c["seq_len"]
    if extra:
        line.update(extra)
    return line


def heap_bytes(comm) -> int:
    return sum(getattr(s, "total_bytes", 0) or s.nbytes for s in getattr(comm, "segments", []))


# ---------------------------------------------------------------------------------------------
def run_ours(args):
    import torch
    import torch.distributed as dist

    from distributed_training_b200.data import PooledWideDataset, TokenDataset
    from distributed_training_b200.ops import native
    from distributed_training_b200.parallel.ddp import DDPStrategy
    from distributed_training_b200.parallel.sharded import FSDPStrategy
    from distributed_training_b200.parallel.strategy import OptimizerSpec
    from distributed_training_b200.trainer import DistributedEnvironment, Trainer, TrainingConfig, prepare_dataloader
    from distributed_training_b200.utils.clocks import ClockSampler
    from distributed_training_b200.utils.timing import barrier_sync, max_over_ranks, time_steps_ms

    rank, local_rank, world = dist_env(args)
    c = args.cfg
    native.require()
    env = DistributedEnvironment("cuda")
    env.setup()
    device = env.torch_device
    torch.manual_seed(0)
    if c["kind"] == "mlp":
        from distributed_training_b200.models.mlp import MLP

        model = MLP(args.width, args.width, args.width, args.depth, dtype=torch.bfloat16)
    else:
        from distributed_training_b200.models.gpt2 import GPT2, GPT2Config

        model = GPT2(GPT2Config(n_embd=c["n_embd"], n_layer=c["n_layer"], n_head=c["n_head"]), dtype=torch.bfloat16)
    wd = 0.0 if c["optimizer"] == "sgd" else 0.1
    cfg = TrainingConfig(max_epochs=1, save_every=10 ** 9, batch_size=args.batch, learning_rate=args.lr,
                         snapshot_path=Path("/nonexistent/bench_snapshot.pt"), device="cuda",
                         parallel_strategy=c["strategy"], loss="cross_entropy", optimizer=c["optimizer"],
                         weight_decay=wd)
    strategy = DDPStrategy(local_rank, True) if c["strategy"] == "ddp" else FSDPStrategy(True)
    trainer = Trainer(model, [], OptimizerSpec(c["optimizer"], lr=args.lr, weight_decay=wd), cfg, env,
                      strategy=strategy)
    comm = strategy.comm
    if args.comm_ctas and comm is not None:
        comm.comm_ctas = args.comm_ctas
    if args.comm_ce >= 0 and comm is not None and hasattr(comm, "copy_engine"):
        comm.copy_engine = bool(args.comm_ce)
    if args.comm_exclusive >= 0 and comm is not None and hasattr(comm, "exclusive_sms"):
        comm.exclusive_sms = args.comm_exclusive
    assert trainer.model.native_step, "the native tcgen05 path must be the one that runs"
    mem_trace = os.environ.get("DTB_MEM_TRACE", "0") == "1" and rank == 0

    def mem(tag):
        if mem_trace:
            print(f"[mem] {tag}: allocated {torch.cuda.memory_allocated() / 2**30:.2f} GB, peak "
                  f"{torch.cuda.max_memory_allocated() / 2**30:.2f} GB, heap {heap_bytes(comm) / 2**30:.2f} GB",
                  file=sys.stderr, flush=True)

    mem("after trainer construction")
    dev_pool = device_pool(args, device, seed=1000 + rank)

    def step_dev(i):
        x, y = dev_pool[i % len(dev_pool)]
        trainer._run_batch(x, y)

    with ClockSampler(local_rank) as sampler:
        for i in range(args.warmup):
            step_dev(i)
            if i == 0:
                torch.cuda.synchronize()
                mem("after first step")
                if mem_trace:      # the largest live blocks, by size
                    sizes = {}
                    for seg in torch.cuda.memory_snapshot():
                        for blk in seg["blocks"]:
                            if blk["state"] == "active_allocated":
                                sizes[blk["size"]] = sizes.get(blk["size"], 0) + 1
                    top = sorted(sizes.items(), key=lambda kv: -kv[0] * kv[1])[:12]
                    print("[mem] live blocks (MiB x count): " + ", ".join(f"{k / 2**20:.1f}x{v}" for k, v in top),
                          file=sys.stderr, flush=True)
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
    # host time to ISSUE a step (no sync inside): when it approaches the device time the step is launch-bound
    import time
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(3):
        step_dev(i)
    host_issue_ms = (time.perf_counter() - t0) / 3 * 1e3
    torch.cuda.synchronize()
    del dev_pool

    # End to end = the body of Trainer._run_epoch over the framework's own data path, the same role the reference
    # arm's loop plays over the reference's DataLoader: a map-style dataset (wide MLP: sample i = row i % 8192 of a
    # pool, as in the reference arm), rank-sharded by ShardedSampler, batches assembled by ShardedLoader
    # (index_select into pinned staging by worker threads -> H2D on a copy stream), then _run_batch and a
    # device->host read of the loss.
    e2e_warm = 3
    e2e_steps = args.steps + e2e_warm
    if c["kind"] == "mlp":
        dataset = PooledWideDataset(e2e_steps * args.batch * world, args.width, args.width, torch.bfloat16)
    else:
        dataset = TokenDataset(e2e_steps * args.batch * world, c["seq_len"], VOCAB)
    loader = prepare_dataloader(dataset, args.batch, device=device, seed=0, num_workers=args.loader_workers)
    trainer.train_data = loader
    loader.sampler.set_epoch(0)
    timed, h2d, waited = 0, 0, 0.0
    start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    it = iter(trainer.train_data)
    for i in range(e2e_steps):
        if i == e2e_warm:
            barrier_sync()
            start.record()
        t0 = time.perf_counter()
        source, targets = next(it)                   # host time spent waiting for the loader (diagnostic)
        waited += (time.perf_counter() - t0) if i >= e2e_warm else 0.0
        h2d = source.numel() * source.element_size() + targets.numel() * targets.element_size()
        loss = trainer._run_batch(source, targets)
        loss.item()
        timed += int(i >= e2e_warm)
    stop.record()
    loader_wait_ms = waited / max(timed, 1) * 1e3
    torch.cuda.synchronize()
    e2e_ms = max_over_ranks(start.elapsed_time(stop))
    barrier_sync()
    trainer.check_comm_errors()
    peak = torch.cuda.max_memory_allocated() / 2 ** 30
    heap = heap_bytes(comm) / 2 ** 30

    if rank == 0:
        line = result_line(args, world, total_ms / args.steps, e2e_ms / max(timed, 1), clocks, launches, "ours",
                           h2d, 4,
                           extra={"exposed_comm_ms_per_step": exposed, "final_loss": loss_val,
                                  "host_issue_ms_per_step": host_issue_ms,
                                  "peak_mem_gb": peak + heap, "peak_mem_note": f"torch allocator peak {peak:.2f} GB + "
                                  f"symmetric heap {heap:.2f} GB",
                                  "comm": {"heap": getattr(comm, "backend", None),
                                           "nvls": bool(getattr(comm, "nvls_active", False)),
                                           "exclusive_sms": getattr(comm, "exclusive_sms", 0)},
                                  "e2e_path": "dataset -> ShardedSampler -> ShardedLoader("
                                              + ("rows pulled from the pinned dataset by dtb_gather_rows on the copy "
                                                 "stream" if loader.device_gather else
                                                 f"pinned staging, {args.loader_workers} worker threads")
                                              + ") -> Trainer._run_batch -> loss.item()",
                                  "loader_wait_ms_per_step": loader_wait_ms, "host_cpus": os.cpu_count()})
        print(json.dumps(line), flush=True)
    if comm is not None:
        comm.close()
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
    from baseline.models import PooledClassificationDataset, TokenPairDataset, TorchGPT2, wide_mlp_torch
    from distributed_training_b200.utils.clocks import ClockSampler
    from distributed_training_b200.utils.timing import barrier_sync, max_over_ranks, time_steps_ms

    c = args.cfg
    env = ref.DistributedEnvironment("cuda")
    env.setup()                                           # init_process_group("nccl") + set_device
    torch.manual_seed(0)
    if c["kind"] == "mlp":
        model = wide_mlp_torch(args.width, args.depth, torch.bfloat16)
    else:
        model = TorchGPT2(c["n_embd"], c["n_layer"], c["n_head"])
    # the reference builds its optimizer on the unwrapped model (src/distributed_trainer.py:199-201)
    optimizer = (torch.optim.SGD(model.parameters(), lr=args.lr) if c["optimizer"] == "sgd" else
                 torch.optim.AdamW(model.parameters(), lr=args.lr, weight_decay=0.1))
    cfg = ref.TrainingConfig(max_epochs=1, save_every=10 ** 9, batch_size=args.batch, learning_rate=args.lr,
                             snapshot_path=Path("/nonexistent/bench_snapshot.pt"), device="cuda",
                             parallel_strategy=c["strategy"])
    e2e_steps = args.steps + 2
    if c["kind"] == "mlp":
        dataset = PooledClassificationDataset(e2e_steps * args.batch * world, args.width, args.width, torch.bfloat16)
    else:
        dataset = TokenPairDataset(e2e_steps * args.batch * world, c["seq_len"], VOCAB)
    loader = ref.prepare_dataloader(dataset, args.batch)   # DataLoader(pin_memory=True, DistributedSampler)
    trainer = ref.Trainer(model, loader, optimizer, cfg, env)   # wraps with the reference's DDPStrategy / FSDPStrategy
    device = torch.device("cuda", local_rank)

    dev_pool = device_pool(args, device, seed=1000 + rank)

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
    del dev_pool

    # end to end = the body of the reference's Trainer._run_epoch on its own DataLoader, plus a small
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
                           extra={"gpu_launches_note": "reference arm: torch/cuBLAS/NCCL kernels only",
                                  "peak_mem_gb": torch.cuda.max_memory_allocated() / 2 ** 30,
                                  "peak_mem_note": "torch allocator peak (NCCL's own buffers not counted)"})
        print(json.dumps(line), flush=True)
    if dist.is_initialized():
        dist.destroy_process_group()


if __name__ == "__main__":
    ns = parse_args()
    if ns.impl == "reference":
        run_reference(ns)
    else:
        run_ours(ns)
