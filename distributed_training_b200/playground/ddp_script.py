"""Data parallelism spelled out with bare collectives (the "playground").

Feature parity with /root/reference/src/playground/ddp_script.py: same CLI
flags and defaults (:193-240), ``mp.spawn`` launch (:254-256), per-rank log
files ``logs/ddp_rank_{rank}.log`` with ``Rank N`` in every line (:56-92),
seed 42 (:108), rank-0 parameter broadcast (:120-121), gradient averaging
after ``backward()`` (:149-154), rank-0 loss lines every ``--log-interval``
batches and an epoch average (:170-181).

Differences, all deliberate:

* gradients are averaged through one flat buffer per step (one collective
  with the ``1/world`` folded in) instead of one ``all_reduce`` + one ``/=``
  per parameter; ``--per-param`` restores the reference's message pattern;
* the two ``torch.norm`` host syncs per parameter per step are
  opt-in via ``--log-grad-norms``;
* on B200 the collectives are the peer-memory kernels (``SymmComm``); gloo is
  used on CPU and for bootstrap only.
"""
from __future__ import annotations

import argparse
import logging
import os
import sys
from pathlib import Path
from typing import List

import torch
import torch.distributed as dist
import torch.multiprocessing as mp
from torch import nn

from ..data import ShardedLoader, ShardedSampler, TensorPairDataset
from ..parallel.comm import Communicator, make_communicator


class SimpleModel(nn.Module):
    """One linear unit, 10 features in, 1 out."""

    def __init__(self, in_features: int = 10, out_features: int = 1) -> None:
        super().__init__()
        self.linear = nn.Linear(in_features, out_features)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return self.linear(x)


class DummyDataset(TensorPairDataset):
    """Standard-normal features and targets, identical on every rank (seeded by the caller)."""

    def __init__(self, size: int = 1000, in_features: int = 10) -> None:
        self.size = size
        self.inputs = torch.randn(size, in_features)
        self.targets = torch.randn(size, 1)

    @property
    def data(self) -> torch.Tensor:
        return self.inputs

    @property
    def labels(self) -> torch.Tensor:
        return self.targets


def setup(rank: int, world_size: int, backend: str = "nccl", port: int = 12355) -> None:
    """Join the bootstrap process group.  ``backend`` is accepted for CLI parity; the group itself
    is always gloo because gradient traffic does not go through it on GPU."""
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", str(port))
    dist.init_process_group(backend="gloo", rank=rank, world_size=world_size)


def cleanup() -> None:
    if dist.is_initialized():
        dist.destroy_process_group()


def setup_logging(rank: int, log_dir: str = "logs") -> logging.Logger:
    """Per-rank logger: ``<log_dir>/ddp_rank_<rank>.log`` + stdout."""
    Path(log_dir).mkdir(parents=True, exist_ok=True)
    log = logging.getLogger(f"ddp_trainer_{rank}")
    log.setLevel(logging.INFO)
    log.propagate = False
    fmt = logging.Formatter("%(asctime)s - Rank %(rank)s - %(levelname)s - %(message)s", datefmt="%Y-%m-%d %H:%M:%S")
    if not log.handlers:
        for handler in (logging.FileHandler(str(Path(log_dir) / f"ddp_rank_{rank}.log")),
                        logging.StreamHandler(sys.stdout)):
            handler.setLevel(logging.INFO)
            handler.setFormatter(fmt)
            log.addHandler(handler)
    return log


class FlatGradients:
    """All gradients of a model as views of one communication buffer."""

    def __init__(self, params: List[nn.Parameter], comm: Communicator) -> None:
        self.params = params
        offsets, total = [], 0
        for p in params:
            offsets.append(total)
            total += (p.numel() + 31) // 32 * 32
        self.flat = comm.alloc(total, params[0].dtype)
        self.views = [self.flat[o: o + p.numel()].view_as(p) for o, p in zip(offsets, params)]
        for p, v in zip(params, self.views):
            p.grad = v

    def reattach(self) -> None:
        for p, v in zip(self.params, self.views):
            if p.grad is None:
                v.zero_()
            elif p.grad.data_ptr() != v.data_ptr():
                v.copy_(p.grad)
            p.grad = v


def average_gradients(model: nn.Module, comm: Communicator, flat: FlatGradients, per_param: bool) -> None:
    if per_param:   # the reference's message pattern: one collective per parameter, on that parameter's slice of
        flat.reattach()    # the flat buffer (allocated once; nothing is taken from the symmetric heap per step)
        for view in flat.views:
            comm.all_reduce_avg_(view.view(-1)).wait()
        return
    flat.reattach()
    comm.all_reduce_avg_(flat.flat).wait()


def train(rank: int, world_size: int, args: argparse.Namespace) -> None:
    """Body of one rank."""
    logger = setup_logging(rank, args.log_dir)
    extra = {"rank": rank}
    torch.manual_seed(42)
    setup(rank, world_size, args.backend, getattr(args, "master_port", 12355))
    try:
        use_cuda = torch.cuda.is_available() and args.backend != "gloo"
        device = torch.device("cuda", rank % max(torch.cuda.device_count(), 1)) if use_cuda else torch.device("cpu")
        if use_cuda:
            torch.cuda.set_device(device)
        model = SimpleModel().to(device)
        comm = make_communicator(device)
        comm.broadcast_([p.data for p in model.parameters()], src=0)

        dataset = DummyDataset()
        sampler = ShardedSampler(dataset, num_replicas=world_size, rank=rank)
        loader = ShardedLoader(dataset, args.batch_size, sampler=sampler, device=device,
                               num_workers=args.num_workers)
        optimizer = torch.optim.SGD(model.parameters(), lr=args.learning_rate)
        criterion = nn.MSELoss()
        flat = FlatGradients(list(model.parameters()), comm)

        model.train()
        for epoch in range(args.epochs):
            sampler.set_epoch(epoch)
            running = torch.zeros((), device=device)
            for batch_idx, (data, target) in enumerate(loader):
                optimizer.zero_grad(set_to_none=False)
                loss = criterion(model(data), target)
                loss.backward()
                average_gradients(model, comm, flat, args.per_param)
                if args.log_grad_norms:
                    for name, param in model.named_parameters():
                        logger.info(f"Epoch {epoch} - Parameter: {name}, Gradient: {torch.norm(param.grad)}",
                                    extra=extra)
                        logger.info(f"Epoch {epoch} - Parameter: {name}, Weight: {torch.norm(param.data)}",
                                    extra=extra)
                optimizer.step()
                running += loss.detach()
                if batch_idx % args.log_interval == 0 and rank == 0:
                    logger.info(f"Epoch: {epoch}, Batch: {batch_idx}, Loss: {loss.item():.6f}", extra=extra)
            if rank == 0:
                logger.info(f"Epoch: {epoch}, Average Loss: {running.item() / max(len(loader), 1):.6f}", extra=extra)
        if getattr(args, "dump_state", None):
            torch.save({k: v.cpu() for k, v in model.state_dict().items()},
                       os.path.join(args.dump_state, f"state_rank{rank}.pt"))
        comm.close()
    finally:
        cleanup()


def get_argument_parser() -> argparse.ArgumentParser:
    parser = argparse.ArgumentParser(description="From-scratch distributed data parallel training")
    parser.add_argument("--batch-size", type=int, default=32, help="per-rank batch size (default: 32)")
    parser.add_argument("--epochs", type=int, default=10, help="number of epochs (default: 10)")
    parser.add_argument("--learning-rate", type=float, default=0.01, help="SGD learning rate (default: 0.01)")
    parser.add_argument("--num-workers", type=int, default=4,
                        help="accepted for CLI parity; the sharded loader gathers whole batches in-process")
    parser.add_argument("--backend", type=str, default="nccl",
                        help="'nccl' = GPU path (peer-memory kernels), 'gloo' = CPU (default: nccl, gloo without CUDA)")
    parser.add_argument("--log-interval", type=int, default=10, help="batches between rank-0 loss lines")
    parser.add_argument("--world-size", type=int,
                        default=torch.cuda.device_count() if torch.cuda.is_available() else 4,
                        help="number of processes (default: #GPUs, or 4 on CPU)")
    parser.add_argument("--log-dir", type=str, default="logs", help="directory for per-rank logs")
    parser.add_argument("--master-port", type=int, default=12355)
    parser.add_argument("--per-param", action="store_true", help="one all-reduce per parameter, as the reference does")
    parser.add_argument("--log-grad-norms", action="store_true", help="log per-parameter grad/weight norms every step")
    parser.add_argument("--dump-state", type=str, default=None, help="directory to save each rank's final weights")
    return parser


def main(argv=None) -> None:
    args = get_argument_parser().parse_args(argv)
    if torch.cuda.is_available():
        logging.info(f"Training on {args.world_size} GPUs")
    else:
        logging.info(f"Training on CPU with {args.world_size} processes")
    mp.spawn(train, args=(args.world_size, args), nprocs=args.world_size, join=True)


if __name__ == "__main__":
    main()
