#!/usr/bin/env bash
# Re-measure every number quoted in BASELINE.md / profiles/README.md on an 8 x B200 box.
#
#   scripts/reproduce_baseline.sh [OUT_DIR] [MAX_GPUS]
#
# Every line goes to OUT_DIR (default gpurun_out/repro) as JSON lines; nothing here runs under a profiler.
# Multi-GPU runs are launched like the driver launches bench.py (torch.distributed.run, 127.0.0.1 rendezvous).
set -uo pipefail
cd "$(dirname "${BASH_SOURCE[0]}")/.."
OUT="${1:-gpurun_out/repro}"
MAXG="${2:-$(nvidia-smi -L 2>/dev/null | wc -l)}"
mkdir -p "$OUT"
python -m distributed_training_b200._native >/dev/null

run() {   # run N ARGS...: bench.py on N GPUs, last stdout line appended to $OUT/bench_lines.jsonl
  local n="$1"; shift
  if [ "$n" -gt "$MAXG" ]; then return 0; fi
  if [ "$n" -eq 1 ]; then
    timeout 900 python bench.py --gpus 1 "$@" 2>>"$OUT/stderr.log" | tail -1 >>"$OUT/bench_lines.jsonl"
  else
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node "$n" --master-addr 127.0.0.1 \
      --master-port $((29600 + n)) bench.py --gpus "$n" "$@" 2>>"$OUT/stderr.log" | tail -1 >>"$OUT/bench_lines.jsonl"
  fi
}

# config 2 (headline): wide MLP, both arms, 1 -> 8 GPUs
for n in 1 2 4 8; do
  run "$n" --impl reference --steps 50 --warmup 5
  run "$n" --steps 50 --warmup 5
done
# config 3: GPT-2 small, DDP
for n in 1 2 8; do
  run "$n" --config gpt2_small --impl reference --steps 20 --warmup 5
  run "$n" --config gpt2_small --steps 20 --warmup 5
done
# config 4: GPT-2 XL, sharded
run 8 --config gpt2_xl_fsdp --impl reference --steps 10 --warmup 3
run 8 --config gpt2_xl_fsdp --steps 10 --warmup 3
# config 5: all-reduce sweep vs NCCL + NVLink denominators
for n in 2 4 8; do
  if [ "$n" -le "$MAXG" ]; then
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node "$n" --master-addr 127.0.0.1 \
      --master-port $((29700 + n)) bench/allreduce_sweep.py --p2p --out "$OUT/allreduce_sweep_${n}gpu.jsonl" \
      >/dev/null 2>>"$OUT/stderr.log"
  fi
done
# per-kernel table, attention head-to-head, determinism probe
timeout 900 python bench/kernel_zoo.py >"$OUT/kernel_zoo_times.jsonl" 2>>"$OUT/stderr.log"
timeout 300 python bench/attention_vs_sdpa.py --out "$OUT/attention_vs_sdpa.jsonl" >/dev/null 2>>"$OUT/stderr.log"
if [ "$MAXG" -ge 2 ]; then
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
    --master-port 29810 bench/ddp_determinism_probe.py >"$OUT/gpt2_ddp_determinism_n2.txt" 2>>"$OUT/stderr.log"
fi
echo "wrote $(ls "$OUT" | wc -l) files to $OUT"
