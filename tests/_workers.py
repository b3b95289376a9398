"""Per-rank bodies for the multi-process (gloo) tests; must be importable by spawned children."""
import os
import sys
from pathlib import Path

import torch
import torch.distributed as dist
from torch import nn

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))


def init(rank, world, port):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    dist.init_process_group("gloo", rank=rank, world_size=world)


def make_mlp(seed):
    torch.manual_seed(seed)
    return nn.Sequential(nn.Linear(20, 64), nn.Tanh(), nn.Linear(64, 32), nn.Tanh(), nn.Linear(32, 1))


def ddp_parity(rank, world, port, out_dir):
    """Our ReplicatedModule vs torch DDP from the same (rank-dependent) init on the same data."""
    from distributed_training_b200.parallel.comm import TorchDistComm
    from distributed_training_b200.parallel.ddp import ReplicatedModule
    from torch.nn.parallel import DistributedDataParallel as DDP

    init(rank, world, port)
    try:
        ours = ReplicatedModule(make_mlp(100 + rank), TorchDistComm(torch.device("cpu")), bucket_cap_mb=0.004,
                                first_bucket_mb=0.002)
        ref = DDP(make_mlp(100 + rank))
        for (n1, p1), (n2, p2) in zip(ours.state_dict().items(), ref.state_dict().items()):
            assert n1 == n2 and torch.equal(p1, p2), f"init broadcast mismatch at {n1}"
        assert len(ours.buckets) > 2
        o1 = torch.optim.SGD(ours.parameters(), lr=0.05)
        o2 = torch.optim.SGD(ref.parameters(), lr=0.05)
        g = torch.Generator().manual_seed(7 + rank)
        for step in range(6):
            x, y = torch.randn(16, 20, generator=g), torch.randn(16, 1, generator=g)
            for model, opt in ((ours, o1), (ref, o2)):
                opt.zero_grad()
                nn.functional.mse_loss(model(x), y).backward()
                opt.step()
            for (n1, p1), (_, p2) in zip(ours.state_dict().items(), ref.state_dict().items()):
                assert torch.allclose(p1, p2, atol=1e-6, rtol=1e-5), f"step {step} param {n1} diverged"
        # replicas bit-identical across ranks
        flat = torch.cat([p.detach().view(-1) for p in ours.parameters()])
        gathered = [torch.empty_like(flat) for _ in range(world)]
        dist.all_gather(gathered, flat)
        assert all(torch.equal(gathered[0], t) for t in gathered)
        # no_sync accumulates locally, next synced step averages the sum
        with ours.no_sync():
            nn.functional.mse_loss(ours(torch.ones(4, 20)), torch.zeros(4, 1)).backward()
        torch.save({"ok": True}, os.path.join(out_dir, f"ok{rank}.pt"))
    finally:
        dist.destroy_process_group()


def trainer_run(rank, world, port, run_dir, overrides, result_name):
    """Full Hydra-configured run through src.distributed_trainer (config 1 of BASELINE.json)."""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    os.chdir(run_dir)
    from distributed_training_b200 import hydra_compat as hc
    from src import distributed_trainer as dt

    conf = ROOT / "conf"
    captured = {}

    def grab(trainer):
        captured["epochs_run"] = trainer.epochs_run
        captured["loss"] = float(trainer.last_loss) if trainer.last_loss is not None else None
        captured["state"] = {k: v.detach().cpu().clone() for k, v in trainer.model.state_dict().items()}

    def task(cfg):
        dt.run(cfg, before_teardown=grab)

    hc.run_job(task, conf, "config", list(overrides))
    torch.save(captured, os.path.join(run_dir, f"{result_name}_rank{rank}.pt"))


def sharded_parity(rank, world, port, out_dir):
    """ShardedModule + ShardedOptimizer vs a single-process reference on the global batch."""
    from distributed_training_b200.parallel.comm import TorchDistComm
    from distributed_training_b200.parallel.sharded import FSDPStrategy
    from distributed_training_b200.parallel.strategy import OptimizerSpec

    init(rank, world, port)
    try:
        strategy = FSDPStrategy(is_gpu=False, comm=TorchDistComm(torch.device("cpu")))
        model = strategy.prepare_model(make_mlp(200 + rank), "cpu")       # rank-dependent init: must be broadcast
        assert len(model.units) == 3
        opt = strategy.build_optimizer(model, OptimizerSpec("adamw", lr=0.01, weight_decay=0.1))
        ref = make_mlp(200)                                                # rank 0's init
        ref_opt = torch.optim.AdamW(ref.parameters(), lr=0.01, weight_decay=0.1)
        g = torch.Generator().manual_seed(3)
        for step in range(5):
            xs, ys = torch.randn(world * 8, 20, generator=g), torch.randn(world * 8, 1, generator=g)
            opt.zero_grad()
            nn.functional.mse_loss(model(xs[rank * 8:(rank + 1) * 8]), ys[rank * 8:(rank + 1) * 8]).backward()
            opt.step()
            ref_opt.zero_grad()
            nn.functional.mse_loss(ref(xs), ys).backward()
            ref_opt.step()
        full = model.state_dict()                                          # collective, unprefixed keys
        assert list(full.keys()) == list(ref.state_dict().keys())
        for k, v in ref.state_dict().items():
            assert torch.allclose(full[k], v, atol=2e-5, rtol=1e-4), f"{k} differs: {(full[k]-v).abs().max()}"
        # checkpoint through the strategy, reload into a fresh sharded model
        path = os.path.join(out_dir, "fsdp_snapshot.pt")
        strategy.save_checkpoint(model, {"MODEL_STATE": full, "EPOCHS_RUN": 3}, path)
        dist.barrier()
        strategy2 = FSDPStrategy(is_gpu=False, comm=TorchDistComm(torch.device("cpu")))
        model2 = strategy2.prepare_model(make_mlp(999), "cpu")
        model2, epoch = strategy2.load_checkpoint(model2, torch.load(path, weights_only=False), "cpu")
        assert epoch == 3
        x = torch.randn(4, 20, generator=torch.Generator().manual_seed(1))
        assert torch.allclose(model2(x), model(x), atol=1e-6)
        # the same file loads into a bare (unwrapped) model: FULL_STATE_DICT layout
        bare = make_mlp(5)
        bare.load_state_dict(torch.load(path, weights_only=False)["MODEL_STATE"])
        assert torch.allclose(bare(x), ref(x), atol=2e-5)
        torch.save({"ok": True}, os.path.join(out_dir, f"ok{rank}.pt"))
    finally:
        dist.destroy_process_group()


def sharded_optimizer_resume(rank, world, port, out_dir):
    """fsdp + save_optimizer: a run that stops after 3 steps, is reloaded from the snapshot (model + optimizer state
    written by rank 0) and continues must end, ON EVERY RANK, with the same master weights and moments as a run that
    never stopped.  (Round 1 saved rank 0's shard only and loaded it into every rank.)"""
    from distributed_training_b200.parallel.comm import TorchDistComm
    from distributed_training_b200.parallel.sharded import FSDPStrategy
    from distributed_training_b200.parallel.strategy import OptimizerSpec, atomic_torch_save

    init(rank, world, port)
    try:
        def build(seed):
            strategy = FSDPStrategy(is_gpu=False, comm=TorchDistComm(torch.device("cpu")))
            model = strategy.prepare_model(make_mlp(seed + rank), "cpu")
            return strategy, model, strategy.build_optimizer(model, OptimizerSpec("adamw", lr=0.01, weight_decay=0.1))

        def steps(model, opt, lo, hi):
            for step in range(lo, hi):
                g = torch.Generator().manual_seed(50 + step)
                xs, ys = torch.randn(world * 8, 20, generator=g), torch.randn(world * 8, 1, generator=g)
                opt.zero_grad()
                nn.functional.mse_loss(model(xs[rank * 8:(rank + 1) * 8]), ys[rank * 8:(rank + 1) * 8]).backward()
                opt.step()

        strategy, model, opt = build(400)
        steps(model, opt, 0, 3)
        snap = {"MODEL_STATE": model.state_dict(), "EPOCHS_RUN": 0, "OPTIMIZER_STATE": opt.state_dict()}  # collective
        path = os.path.join(out_dir, "snap_opt.pt")
        if rank == 0:
            assert snap["OPTIMIZER_STATE"]["format"] == "full" and snap["OPTIMIZER_STATE"]["step"] == 3
            for unit in model.units:       # rank 0 holds every rank's shard, not just its own
                assert snap["OPTIMIZER_STATE"]["state"][unit.name]["exp_avg"].numel() == unit.full_numel
            atomic_torch_save(snap, path)
        dist.barrier()
        steps(model, opt, 3, 5)                                    # the run that never stopped
        want = {u.name: {k: v.clone() for k, v in opt.state[u.name].items()} for u in model.units}

        strategy2, model2, opt2 = build(900)                       # "relaunch": different init, fresh optimizer
        loaded = torch.load(path, weights_only=False)
        strategy2.load_checkpoint(model2, loaded, "cpu")
        opt2.load_state_dict(loaded["OPTIMIZER_STATE"])
        assert opt2.step_count == 3
        steps(model2, opt2, 3, 5)
        for u in model2.units:
            for k, v in opt2.state[u.name].items():
                assert torch.equal(v, want[u.name][k]), (rank, u.name, k, (v - want[u.name][k]).abs().max())
        torch.save({"ok": True}, os.path.join(out_dir, f"ok{rank}.pt"))
    finally:
        dist.destroy_process_group()


def mismatched_replicas(rank, world, port, out_dir):
    """Ranks wrap different architectures: the wrapper must refuse on EVERY rank (torch's shape verification)."""
    from distributed_training_b200.parallel.comm import TorchDistComm
    from distributed_training_b200.parallel.ddp import ReplicatedModule

    init(rank, world, port)
    try:
        model = nn.Linear(20, 1 + rank)
        try:
            ReplicatedModule(model, TorchDistComm(torch.device("cpu")))
        except RuntimeError as err:
            assert "disagree" in str(err)
            torch.save({"ok": True}, os.path.join(out_dir, f"ok{rank}.pt"))
    finally:
        dist.destroy_process_group()
