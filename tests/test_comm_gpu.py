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


def test_collectives_exact_and_bitwise_identical(tmp_path):
    w = _world()
    mp.spawn(G.collectives, args=(w, free_port(), str(tmp_path)), nprocs=w, join=True)
    assert all((tmp_path / f"ok{r}.pt").exists() for r in range(w))


def test_native_ddp_step_matches_single_process(tmp_path):
    w = _world()
    mp.spawn(G.ddp_native, args=(w, free_port(), str(tmp_path)), nprocs=w, join=True)
    assert all((tmp_path / f"ok{r}.pt").exists() for r in range(w))


def test_native_sharded_strategy_matches_single_process(tmp_path):
    w = _world()
    mp.spawn(G.sharded_native, args=(w, free_port(), str(tmp_path)), nprocs=w, join=True)
    assert all((tmp_path / f"ok{r}.pt").exists() for r in range(w))
