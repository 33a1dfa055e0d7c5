#!/usr/bin/env bash
# One rank per GPU under torchrun (the reference's "1-step" layout: n=D, d=1; rank and size come
# from the environment through envarg://, as they did from mpirun in submits/submit2x*.sh):
#   scripts/run_torchrun.sh <gpus> <model> <experiment> [extra main.py options]
set -euo pipefail
GPUS=${1:?number of GPUs}; MODEL=${2:-imagenet}; EXP=${3:-main_onestep_reduce}; shift $(( $# < 3 ? $# : 3 ))
ROOT="$(cd "$(dirname "$0")/.." && pwd)"
PORT=${MASTER_PORT:-29600}
exec python -m torch.distributed.run --nnodes=1 --nproc-per-node "$GPUS" --master-addr 127.0.0.1 --master-port "$PORT" \
     "$ROOT/main.py" envarg://WORLD_SIZE envarg://RANK 1 "$GPUS" 127.0.0.1 lo "$MODEL" "${DATA_ROOT:-/tmp/dlb200_data}" 1 \
     --backend "${BACKEND:-b200}" --experiment "$EXP" --random_input 1 --master_port $((PORT + 7)) "$@"
