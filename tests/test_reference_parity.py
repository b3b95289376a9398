"""Snapshot interchange with the reference's own classes, both directions.

The reference runs in a subprocess (it is imported as ``src.*`` from baseline/_ref, which would
collide with this repo's ``src`` package in-process)."""
import os
import subprocess
import sys
import textwrap
from pathlib import Path

import pytest
import torch

from conftest import free_port

ROOT = Path(__file__).resolve().parent.parent
pytestmark = pytest.mark.skipif(not (ROOT / "baseline" / "_ref" / "src").exists(),
                                reason="reference not installed (baseline/install_reference.sh)")


def _run_ref(code: str, port: int) -> str:
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK="0", WORLD_SIZE="1", LOCAL_RANK="0",
               PYTHONPATH=str(ROOT))
    prelude = ("import sys; sys.path.insert(0, %r)\nfrom baseline.ref_loader import load_reference\n"
               "ref = load_reference()\n") % str(ROOT)
    out = subprocess.run([sys.executable, "-c", prelude + textwrap.dedent(code)], env=env, capture_output=True,
                         text=True, timeout=300)
    assert out.returncode == 0, out.stderr[-3000:]
    return out.stdout


def test_reference_snapshot_loads_in_ours_and_resumes(tmp_path):
    snap = tmp_path / "ref_snapshot.pt"
    _run_ref(f"""
        import torch
        from pathlib import Path
        cfgT = ref.TrainingConfig(max_epochs=3, save_every=2, batch_size=32, learning_rate=0.001,
                                  snapshot_path=Path({str(snap)!r}), device="cpu")
        env = ref.DistributedEnvironment("cpu"); env.setup()
        ds = ref.MyTrainDataset(256)
        model = torch.nn.Linear(20, 1)
        opt = torch.optim.SGD(model.parameters(), lr=0.001)
        tr = ref.Trainer(model, ref.prepare_dataloader(ds, 32), opt, cfgT, env)
        tr.train(3)
        torch.distributed.destroy_process_group()
    """, free_port())
    ckpt = torch.load(snap, weights_only=False)
    assert set(ckpt) == {"MODEL_STATE", "EPOCHS_RUN"} and ckpt["EPOCHS_RUN"] == 2
    assert list(ckpt["MODEL_STATE"]) == ["module.weight", "module.bias"]

    # ours: load the reference's file and resume at the stored epoch
    from distributed_training_b200.data import MyTrainDataset
    from distributed_training_b200.parallel.comm import SingleProcessComm
    from distributed_training_b200.parallel.ddp import DDPStrategy
    from distributed_training_b200.trainer import (DistributedEnvironment, Trainer, TrainingConfig,
                                                   prepare_dataloader)

    cfg = TrainingConfig(4, 2, 32, 0.001, snap, device="cpu")
    env = DistributedEnvironment("cpu")
    strategy = DDPStrategy(0, False, comm=SingleProcessComm(torch.device("cpu")))
    ds = MyTrainDataset(64)
    trainer = Trainer(torch.nn.Linear(20, 1), prepare_dataloader(ds, 32), None, cfg, env, strategy=strategy)
    assert trainer.epochs_run == 2
    assert torch.equal(trainer.model.module.weight.detach(), ckpt["MODEL_STATE"]["module.weight"])
    trainer.train(4)
    assert torch.load(snap, weights_only=False)["EPOCHS_RUN"] == 2     # epoch 2 re-run and re-saved (Q3 semantics)


def test_our_snapshot_loads_in_reference(tmp_path):
    from distributed_training_b200.data import LinearTeacherDataset
    from distributed_training_b200.parallel.comm import SingleProcessComm
    from distributed_training_b200.parallel.ddp import DDPStrategy
    from distributed_training_b200.trainer import (DistributedEnvironment, Trainer, TrainingConfig,
                                                   prepare_dataloader)

    snap = tmp_path / "our_snapshot.pt"
    cfg = TrainingConfig(3, 2, 16, 0.05, snap, device="cpu")
    strategy = DDPStrategy(0, False, comm=SingleProcessComm(torch.device("cpu")))
    trainer = Trainer(torch.nn.Linear(20, 1), prepare_dataloader(LinearTeacherDataset(64), 16), None, cfg,
                      DistributedEnvironment("cpu"), strategy=strategy)
    trainer.train(3)
    ours = torch.load(snap, weights_only=False)
    out = _run_ref(f"""
        import torch
        from src.dist_strategy.ddp_strategy import DDPStrategy
        env = ref.DistributedEnvironment("cpu"); env.setup()
        strat = DDPStrategy(0, False)
        model = strat.prepare_model(torch.nn.Linear(20, 1), "cpu")
        ckpt = torch.load({str(snap)!r}, map_location="cpu")
        model, epoch = strat.load_checkpoint(model, ckpt, "cpu")
        print("EPOCH", epoch, float(model.module.weight.sum()))
        torch.distributed.destroy_process_group()
    """, free_port())
    tag, epoch, wsum = out.strip().splitlines()[-1].split()
    assert tag == "EPOCH" and int(epoch) == 2
    assert abs(float(wsum) - float(ours["MODEL_STATE"]["module.weight"].sum())) < 1e-6


def test_reference_loss_is_identically_zero_quirk(tmp_path):
    """Q1: the reference's default objective has zero gradient; ours keeps that form selectable."""
    from distributed_training_b200.trainer import Trainer, TrainingConfig, DistributedEnvironment, prepare_dataloader
    from distributed_training_b200.data import MyTrainDataset
    from distributed_training_b200.parallel.comm import SingleProcessComm
    from distributed_training_b200.parallel.ddp import DDPStrategy

    cfg = TrainingConfig(1, 5, 32, 0.1, tmp_path / "s.pt", device="cpu", loss="cross_entropy")
    trainer = Trainer(torch.nn.Linear(20, 1), prepare_dataloader(MyTrainDataset(64), 32), None, cfg,
                      DistributedEnvironment("cpu"),
                      strategy=DDPStrategy(0, False, comm=SingleProcessComm(torch.device("cpu"))))
    before = trainer.model.module.weight.detach().clone()
    trainer.train(1)
    assert float(trainer.last_loss) == 0.0 and torch.equal(before, trainer.model.module.weight.detach())
