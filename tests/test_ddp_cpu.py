"""Distributed plumbing on CPU/gloo, world 2 and 4 (SURVEY §4.3 tier 2; BASELINE.json config 1)."""
import pytest
import torch
import torch.multiprocessing as mp

import _workers as W
from conftest import free_port


@pytest.mark.parametrize("world", [2, 4])
def test_replicated_module_matches_torch_ddp(world, tmp_path):
    mp.spawn(W.ddp_parity, args=(world, free_port(), str(tmp_path)), nprocs=world, join=True)
    assert all((tmp_path / f"ok{r}.pt").exists() for r in range(world))


def test_mismatched_replicas_are_rejected(tmp_path):
    mp.spawn(W.mismatched_replicas, args=(2, free_port(), str(tmp_path)), nprocs=2, join=True)
    assert all((tmp_path / f"ok{r}.pt").exists() for r in range(2))


def test_sharded_strategy_matches_single_process(tmp_path):
    mp.spawn(W.sharded_parity, args=(2, free_port(), str(tmp_path)), nprocs=2, join=True)
    assert all((tmp_path / f"ok{r}.pt").exists() for r in range(2))


def test_sharded_optimizer_state_resumes_on_every_rank(tmp_path):
    mp.spawn(W.sharded_optimizer_resume, args=(2, free_port(), str(tmp_path)), nprocs=2, join=True)
    assert all((tmp_path / f"ok{r}.pt").exists() for r in range(2))


def test_fused_optimizer_state_dict_keeps_the_step_count():
    from distributed_training_b200.optim import FusedAdamW

    p = torch.nn.Parameter(torch.ones(4))
    opt = FusedAdamW([p], lr=0.1, master_weights=False)
    for _ in range(3):
        p.grad = torch.ones(4)
        opt.step()
    sd = opt.state_dict()
    assert sd["fused_step_count"] == 3
    q = torch.nn.Parameter(torch.ones(4))
    opt2 = FusedAdamW([q], lr=0.1, master_weights=False)
    opt2.load_state_dict(sd)
    assert opt2._step_count == 3 and torch.equal(opt2.state[q]["exp_avg"], opt.state[p]["exp_avg"])


def _run_trainer(tmp_path, world, overrides, name):
    mp.spawn(W.trainer_run, args=(world, free_port(), str(tmp_path), overrides, name), nprocs=world, join=True)
    return [torch.load(tmp_path / f"{name}_rank{r}.pt", weights_only=False) for r in range(world)]


def test_trainer_end_to_end_save_cadence_and_resume(tmp_path):
    snap = tmp_path / "snap.pt"
    common = [f"train.snapshot_path={snap}", "train.dataset_size=256", "train.batch_size=16", "train.dataset=teacher",
              "train.learning_rate=0.05", "train.device=cpu", "logging.log_every=4", "model=mlp_toy"]
    first = _run_trainer(tmp_path, 2, common + ["train.total_epochs=5", f"hydra.run.dir={tmp_path}/run1"], "first")
    ckpt = torch.load(snap, weights_only=False)
    assert set(ckpt) == {"MODEL_STATE", "EPOCHS_RUN"}
    assert ckpt["EPOCHS_RUN"] == 4                                  # epochs 0,2,4 saved; last wins
    assert all(k.startswith("module.") for k in ckpt["MODEL_STATE"])
    # replicas identical
    for k, v in first[0]["state"].items():
        assert torch.equal(v, first[1]["state"][k])
    # log + metrics + hydra dump exist
    run1 = tmp_path / "run1"
    assert (run1 / "train.log").exists() and (run1 / ".hydra" / "config.yaml").exists()
    lines = (run1 / "metrics.jsonl").read_text().strip().splitlines()
    assert len(lines) >= 2
    import json
    losses = [json.loads(l)["loss"] for l in lines]
    assert losses[-1] < losses[0]                                   # it actually learns (Q1 fixed via MSE)
    # resume: starts at the stored epoch (Q2 fixed, Q3 semantics kept) and continues to 7
    second = _run_trainer(tmp_path, 2, common + ["train.total_epochs=7", f"hydra.run.dir={tmp_path}/run2"], "second")
    log2 = (tmp_path / "run2" / "train.log").read_text()
    assert "Loading snapshot" in log2
    assert "Epoch 4 |" in log2 and "Epoch 3 |" not in log2
    assert torch.load(snap, weights_only=False)["EPOCHS_RUN"] == 6


def test_trainer_fsdp_strategy_on_cpu(tmp_path):
    snap = tmp_path / "snap_fsdp.pt"
    overrides = [f"train.snapshot_path={snap}", "train.dataset_size=128", "train.batch_size=16",
                 "train.dataset=teacher", "train.total_epochs=2", "train.device=cpu",
                 "train.parallel_strategy=fsdp", "model=mlp_toy",
                 "train.optimizer=adamw", "train.learning_rate=0.01", f"hydra.run.dir={tmp_path}/runf"]
    res = _run_trainer(tmp_path, 2, overrides, "fsdp")
    ckpt = torch.load(snap, weights_only=False)
    assert ckpt["EPOCHS_RUN"] == 0 and not any(k.startswith("module.") for k in ckpt["MODEL_STATE"])
    for k, v in res[0]["state"].items():
        assert torch.equal(v, res[1]["state"][k])


def test_unknown_strategy_raises():
    from distributed_training_b200.trainer import DistributedEnvironment, Trainer, TrainingConfig

    cfg = TrainingConfig(1, 1, 4, 0.1, "x.pt", device="cpu", parallel_strategy="zero9")
    with pytest.raises(ValueError, match="parallel_strategy"):
        Trainer(torch.nn.Linear(2, 1), [], None, cfg, DistributedEnvironment("cpu"))


def test_playground_runs_and_replicas_agree(tmp_path):
    from distributed_training_b200.playground import ddp_script

    args = ddp_script.get_argument_parser().parse_args(
        ["--world-size", "2", "--epochs", "2", "--backend", "gloo", "--log-dir", str(tmp_path / "logs"),
         "--master-port", str(free_port()), "--dump-state", str(tmp_path)])
    mp.spawn(ddp_script.train, args=(2, args), nprocs=2, join=True)
    s0 = torch.load(tmp_path / "state_rank0.pt")
    s1 = torch.load(tmp_path / "state_rank1.pt")
    assert all(torch.equal(s0[k], s1[k]) for k in s0)
    log = (tmp_path / "logs" / "ddp_rank_0.log").read_text()
    assert "Rank 0" in log and "Average Loss" in log
    import re
    avgs = [float(m) for m in re.findall(r"Average Loss: ([0-9.]+)", log)]
    assert len(avgs) == 2 and avgs[1] < avgs[0]
