#!/usr/bin/env bash
# N "nodes" on one machine over gloo (how the reference was developed on laptops):
#   scripts/run_cpu_nodes.sh <nodes> <experiment> [extra options]
set -euo pipefail
N=${1:?nodes}; EXP=${2:-experiment3}; shift $(( $# < 2 ? $# : 2 ))
ROOT="$(cd "$(dirname "$0")/.." && pwd)"
for r in $(seq 0 $((N - 1))); do
  python "$ROOT/main.py" "$N" "$r" 1 "$N" 127.0.0.1 lo mnist "${DATA_ROOT:-/tmp/dlb200_data}" 0 \
         --experiment "$EXP" --random_input 1 "$@" &
done
wait
