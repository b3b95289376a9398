"""Peer-memory collectives and the native DDP step on >= 2 GPUs (SURVEY §4.3 collective tier)."""
import pytest
import torch
import torch.multiprocessing as mp

import _gpu_workers as G
from conftest import free_port

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]


def _world():
    n = torch.cuda.device_count()
    return 8 if n >= 8 else (4 if n >= 4 else 2)


@pytest.mark.parametrize("nvls", ["0", "1"])
def test_collectives_exact_and_bitwise_identical(tmp_path, nvls):
    """nvls=0: peer-load/store kernels (bit-exact vs a fixed-order fp32 sum); nvls=1: multimem kernels on the
    multicast mapping (identical on every rank, equal to the oracle within one rounding)."""
    w = _world()
    mp.spawn(G.collectives, args=(w, free_port(), str(tmp_path), nvls), nprocs=w, join=True)
    assert all((tmp_path / f"ok{r}.pt").exists() for r in range(w))
    res = torch.load(tmp_path / "ok0.pt", weights_only=False)
    if res.get("skipped"):
        pytest.skip(res["skipped"])
    assert res["nvls"] == (nvls == "1")


@pytest.mark.timeout(600)
def test_collectives_stress_sizes_back_to_back_and_beside_gemm(tmp_path):
    """1 KiB - 1 GiB against an NCCL fp32 oracle, queued back to back, and beside a GEMM burst with launch skew."""
    import os

    w = _world()
    max_mb = int(os.environ.get("DTB_STRESS_MAX_MB", "1024"))
    mp.spawn(G.collectives_stress, args=(w, free_port(), str(tmp_path), max_mb), nprocs=w, join=True)
    res = [torch.load(tmp_path / f"ok{r}.pt", weights_only=False) for r in range(w)]
    assert all(r["ok"] and r["checked"] >= 30 for r in res), res


@pytest.mark.timeout(120)
def test_watchdog_turns_a_missing_peer_into_an_error(tmp_path):
    w = 2
    mp.spawn(G.watchdog, args=(w, free_port(), str(tmp_path)), nprocs=w, join=True)
    assert torch.load(tmp_path / "ok0.pt", weights_only=False)["raised"] is True


def test_native_ddp_step_matches_single_process(tmp_path):
    w = _world()
    mp.spawn(G.ddp_native, args=(w, free_port(), str(tmp_path)), nprocs=w, join=True)
    assert all((tmp_path / f"ok{r}.pt").exists() for r in range(w))


def test_native_gpt2_ddp_step_matches_single_process(tmp_path):
    w = 2
    mp.spawn(G.ddp_native_gpt2, args=(w, free_port(), str(tmp_path)), nprocs=w, join=True)
    assert all((tmp_path / f"ok{r}.pt").exists() for r in range(w))


def test_native_sharded_strategy_matches_single_process(tmp_path):
    w = _world()
    mp.spawn(G.sharded_native, args=(w, free_port(), str(tmp_path)), nprocs=w, join=True)
    assert all((tmp_path / f"ok{r}.pt").exists() for r in range(w))


@pytest.mark.timeout(150)
def test_trainer_minimum_slice_on_gpus(tmp_path):
    w = 2
    for total, tag in ((3, "a"), (5, "b")):      # second launch = a relaunch that resumes from the snapshot
        mp.spawn(G.trainer_gpu, args=(w, free_port(), str(tmp_path), total, tag), nprocs=w, join=True)
    a = [torch.load(tmp_path / f"res_a_rank{r}.pt", weights_only=False) for r in range(w)]
    b = [torch.load(tmp_path / f"res_b_rank{r}.pt", weights_only=False) for r in range(w)]
    assert a[0]["comm"] == "SymmComm" and a[0]["launches"] > 0          # native kernels, not a fallback
    for k in a[0]["state"]:
        assert torch.equal(a[0]["state"][k], a[1]["state"][k]) and torch.equal(b[0]["state"][k], b[1]["state"][k])
    ckpt = torch.load(tmp_path / "snap.pt", weights_only=False)
    assert ckpt["EPOCHS_RUN"] == 4 and list(ckpt["MODEL_STATE"]) == ["module.weight", "module.bias"]
    log_b = (tmp_path / "run_b" / "train.log").read_text()
    assert "Loading snapshot" in log_b and "Epoch 2 |" in log_b and "Epoch 1 |" not in log_b
