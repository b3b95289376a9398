#!/usr/bin/env bash
# Single-node launcher: the 8xB200-box counterpart of the torchrun line the reference's cloud-init generates
# (/root/reference/infrastructure/nebius/cluster/scripts/cloud-init.tftpl:61-77), pointed at the real entry point
# (the reference's scripts name a file that does not exist).
#
#   scripts/launch_single_node.sh [NUM_GPUS] [hydra overrides...]
#   scripts/launch_single_node.sh 8 model=mlp_wide train=wide train.snapshot_path=/mnt/filesystem/snapshot.pt
#
# Automatic resume: a relaunch with the same snapshot_path continues at the stored epoch, so the script is safe to
# wrap in `torchrun --max-restarts`.
set -euo pipefail
NGPU="${1:-$(nvidia-smi -L 2>/dev/null | wc -l)}"
shift || true
cd "$(dirname "${BASH_SOURCE[0]}")/.."
python -m distributed_training_b200._native >/dev/null          # build libdtb200.so if sources changed
exec python -m torch.distributed.run --nnodes=1 --nproc-per-node "${NGPU}" \
    --master-addr 127.0.0.1 --master-port "${MASTER_PORT:-29500}" --max-restarts "${MAX_RESTARTS:-0}" \
    -m src.distributed_trainer "$@"
