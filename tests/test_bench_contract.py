"""bench.py's command-line contract (what the driver relies on), checked without a GPU."""
import importlib.util
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent


def _load_bench():
    spec = importlib.util.spec_from_file_location("bench_main", ROOT / "bench.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def test_defaults_and_config_table(monkeypatch):
    bench = _load_bench()
    monkeypatch.setattr(sys, "argv", ["bench.py"])
    args = bench.parse_args()
    assert args.gpus == 1 and args.warmup >= 3 and args.steps > 0 and args.impl == "ours"
    assert args.config == "wide_mlp" and args.batch == 4096 and args.cfg["strategy"] == "ddp"
    assert args.lr == 0.05                      # SGD: large enough that bf16 weights move every step
    monkeypatch.setattr(sys, "argv", ["bench.py", "--config", "gpt2_xl_fsdp", "--gpus", "8", "--impl", "reference"])
    args = bench.parse_args()
    assert args.cfg["strategy"] == "fsdp" and args.cfg["n_embd"] == 1600 and args.cfg["n_layer"] == 48
    assert args.batch == 4 and args.lr == 1e-4 and args.impl == "reference"
    assert set(bench.CONFIGS) == {"wide_mlp", "gpt2_small", "gpt2_xl_fsdp"}


def test_result_line_schema(monkeypatch):
    bench = _load_bench()
    monkeypatch.setattr(sys, "argv", ["bench.py", "--config", "gpt2_small"])
    args = bench.parse_args()
    line = bench.result_line(args, 8, 12.5, 12.9, {"sm_mhz": 1500, "sm_max_mhz": 1965, "reasons": []}, 4200, "ours",
                             h2d=131200, d2h=4)
    for key in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
                "vs_baseline", "dtype", "data", "config", "clocks", "e2e", "gpu_launches", "impl"):
        assert key in line, key
    assert line["n_gpus"] == 8 and line["scaling"] == "weak" and line["higher_is_better"] is True
    assert abs(line["value"] - 8 * 8 / 12.5e-3) < 1e-6          # whole-job samples/s
    assert line["config"]["parallelism"] == "dp8" and line["config"]["global_batch"] == 64
    assert line["e2e"]["h2d_bytes_per_step"] == 131200 and line["e2e"]["d2h_bytes_per_step"] == 4
    assert abs(line["tokens_per_s"] - line["value"] * 1024) < 1e-3
