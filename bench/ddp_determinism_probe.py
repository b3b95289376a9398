"""Run-to-run reproducibility of the native GPT-2 DDP step: per-step losses and a parameter checksum.

    torchrun --nproc-per-node 2 bench/ddp_determinism_probe.py [--layers 4] [--steps 12]

Two launches with the same arguments should print the same numbers up to the fp32-atomic noise of the attention /
embedding backward kernels; a difference that appears abruptly at one step points at an ordering bug instead.
"""
from __future__ import annotations

import argparse
import os
import sys
from pathlib import Path

import torch
import torch.distributed as dist

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--layers", type=int, default=4)
    ap.add_argument("--steps", type=int, default=12)
    ap.add_argument("--lr", type=float, default=1e-4)
    ns = ap.parse_args()
    from distributed_training_b200.models.gpt2 import GPT2, GPT2Config
    from distributed_training_b200.parallel.ddp import DDPStrategy
    from distributed_training_b200.parallel.strategy import OptimizerSpec

    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
    if world > 1:
        dist.init_process_group("gloo")
    strategy = DDPStrategy(rank, True)
    torch.manual_seed(0)
    cfg = GPT2Config(n_layer=ns.layers)
    model = GPT2(cfg, dtype=torch.bfloat16).cuda()
    wrapped = strategy.prepare_model(model, rank)
    opt = strategy.build_optimizer(wrapped, OptimizerSpec("adamw", lr=ns.lr, weight_decay=0.1))
    wrapped.attach_optimizer(opt)
    g = torch.Generator().manual_seed(1234 + rank)
    batches = [torch.randint(0, 50257, (8, 1025), generator=g).cuda() for _ in range(4)]
    losses = []
    for step in range(ns.steps):
        t = batches[step % 4]
        loss = wrapped.forward_backward(t[:, :-1].contiguous(), t[:, 1:].contiguous())
        opt.step()
        losses.append(float(loss))
    torch.cuda.synchronize()
    flat = torch.cat([p.detach().float().view(-1) for p in model.parameters()])
    checksum = float(flat.double().sum()), float(flat.double().abs().sum())
    same = True
    if world > 1:
        mine = torch.tensor(checksum, dtype=torch.float64)
        got = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(got, mine)
        same = all(torch.equal(got[0], x) for x in got)
    if rank == 0:
        print("losses", " ".join(f"{x:.6f}" for x in losses))
        print("checksum", f"{checksum[0]:.6f} {checksum[1]:.6f}", "replicas_identical", same)
    if strategy.comm is not None:
        strategy.comm.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
