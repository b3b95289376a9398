"""Config layer: composition, overrides, interpolation, run-dir handling."""
import os
from pathlib import Path

import pytest
import yaml

from distributed_training_b200 import hydra_compat as hc

CONF = Path(__file__).resolve().parent.parent / "conf"


def test_defaults_match_reference_keys():
    cfg = hc.compose(CONF, "config")
    assert cfg.model.input_size == 20 and cfg.model.output_size == 1
    t = cfg.train
    assert (t.batch_size, t.total_epochs, t.save_every, t.dataset_size) == (32, 10, 2, 2048)
    assert t.snapshot_path == "snapshot.pt" and t.learning_rate == 0.001
    assert t.device == "auto" and t.parallel_strategy == "ddp"
    assert cfg.logging.level == "INFO"


def test_value_and_group_overrides():
    cfg = hc.compose(CONF, "config", ["train.batch_size=64", "model=mlp_wide", "train.parallel_strategy=fsdp"])
    assert cfg.train.batch_size == 64
    assert cfg.model.hidden_size == 8192 and cfg.model.num_layers == 8
    assert cfg.train.parallel_strategy == "fsdp"
    assert cfg.hydra.runtime.choices["model"] == "mlp_wide"


def test_add_and_delete_and_unknown_key():
    cfg = hc.compose(CONF, "config", ["+train.extra=3", "~train.seed"])
    assert cfg.train.extra == 3 and "seed" not in cfg.train
    with pytest.raises(hc.ConfigError):
        hc.compose(CONF, "config", ["train.no_such_key=1"])
    with pytest.raises(hc.ConfigError):
        hc.compose(CONF, "config", ["model=does_not_exist"])


def test_get_with_default_and_types():
    cfg = hc.compose(CONF, "config", ["train.learning_rate=1e-2", "train.snapshot_path=/tmp/x.pt"])
    assert cfg.train.get("device", "cpu") == "auto"
    assert cfg.train.get("missing", 7) == 7
    assert isinstance(cfg.train.learning_rate, float) and cfg.train.learning_rate == 0.01
    assert cfg.train.snapshot_path == "/tmp/x.pt"


def test_interpolations_resolve_against_run_dir(tmp_path):
    cfg = hc.compose(CONF, "config", [f"hydra.run.dir={tmp_path}/run"])
    assert cfg.logging.file == f"{tmp_path}/run/train.log"
    cfg2 = hc.compose(CONF, "config")
    parts = Path(cfg2.hydra.run.dir).parts
    assert parts[0] == "outputs" and len(parts) == 3     # outputs/<date>/<time>


def test_main_decorator_creates_run_dir_and_dumps(tmp_path, monkeypatch):
    monkeypatch.chdir(tmp_path)
    seen = {}

    def task(cfg):
        seen["cwd"] = os.getcwd()
        seen["bs"] = cfg.train.batch_size
        seen["log"] = cfg.logging.file
        assert "hydra" not in cfg
        return 5

    rc = hc.run_job(task, CONF, "config", ["train.batch_size=8", "hydra.run.dir=myrun"])
    assert rc == 5 and seen["bs"] == 8
    run = tmp_path / "myrun"
    assert Path(seen["cwd"]).resolve() == run.resolve()          # hydra.job.chdir: True
    assert seen["log"] == str(run / "train.log")
    saved = yaml.safe_load((run / ".hydra" / "overrides.yaml").read_text())
    assert saved == ["train.batch_size=8", "hydra.run.dir=myrun"]
    dumped = yaml.safe_load((run / ".hydra" / "config.yaml").read_text())
    assert dumped["train"]["batch_size"] == 8 and "hydra" not in dumped
    assert os.getcwd() == str(tmp_path)                             # restored


def test_stub_modules_satisfy_reference_imports():
    import sys

    saved = {k: sys.modules.get(k) for k in ("hydra", "omegaconf")}
    try:
        assert hc.install_stub_modules(force=True)
        import hydra
        from omegaconf import DictConfig

        assert callable(hydra.main) and DictConfig is hc.ConfigNode
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
