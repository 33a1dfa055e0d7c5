#!/usr/bin/env bash
# Single-box launcher (replaces the reference's SLURM templates submit.sh / submits/*.sh).
# One main.py process on this host that spawns one worker per GPU:
#   scripts/run_local.sh <gpus> <model> <experiment> [extra main.py options]
# e.g. scripts/run_local.sh 8 imagenet main_ourdist --limit_batches 50
set -euo pipefail
GPUS=${1:?number of GPUs (0 = CPU)}; MODEL=${2:-mnist}; EXP=${3:-main_ourdist}; shift $(( $# < 3 ? $# : 3 ))
ROOT="$(cd "$(dirname "$0")/.." && pwd)"
if [ "$GPUS" -gt 0 ]; then
  exec python "$ROOT/main.py" 1 0 "$GPUS" "$GPUS" 127.0.0.1 lo "$MODEL" "${DATA_ROOT:-/tmp/dlb200_data}" 1 \
       --backend "${BACKEND:-b200}" --experiment "$EXP" --random_input 1 "$@"
else
  exec python "$ROOT/main.py" 1 0 1 1 127.0.0.1 lo "$MODEL" "${DATA_ROOT:-/tmp/dlb200_data}" 0 \
       --backend gloo --experiment "$EXP" --random_input 1 "$@"
fi
