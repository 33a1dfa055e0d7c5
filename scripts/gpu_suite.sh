#!/usr/bin/env bash
# The measurement suite behind BENCH_NOTES.md: correctness at N ranks, the flagship bench (both
# arms), the allreduce sweep, ResNet-50 bf16, the small-CNN config and the fusion sweeps.
#   [SKIP="step names"] scripts/gpu_suite.sh <N>
# Everything lands in gpurun_out/suite_n<N>/ (one .out/.err per step).
set -u
N=${1:-2}; MODE=${2:-full}
ROOT="$(cd "$(dirname "$0")/.." && pwd)"; cd "$ROOT"
OUT=gpurun_out/suite_n$N; mkdir -p "$OUT"
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
SKIP=" ${SKIP:-} "
step() { local name=$1; shift; case "$SKIP" in *" $name "*) echo "[$name] skipped"; return;; esac; local t0; t0=$(date +%s); "$@" > "$OUT/$name.out" 2> "$OUT/$name.err"; local rc=$?
         echo "[$name] rc=$rc $(( $(date +%s) - t0 ))s"; tail -n 1 "$OUT/$name.out" | cut -c1-400; }
M="python main.py 1 0 $N $N 127.0.0.1 lo"

step mp_worker      timeout 600 $TR --master-port 29701 tests/_mp_worker.py all
step bench_b200     timeout 300 $TR --master-port 29702 bench.py --gpus "$N" --steps 20 --warmup 5
step bench_ref      timeout 500 $TR --master-port 29703 bench.py --impl reference --gpus "$N" --steps 10 --warmup 3
step allreduce_fp32 timeout 600 $TR --master-port 29704 benchmarks/bench_allreduce.py --ring --out "$OUT/allreduce_fp32.json"
step allreduce_bf16 timeout 300 $TR --master-port 29705 benchmarks/bench_allreduce.py --dtype bf16 --min-bytes 65536 --factor 16 --blocks 16,32,48 --out "$OUT/allreduce_bf16.json"
step bench_resnet50 timeout 300 $TR --master-port 29706 bench.py --model resnet50 --dtype bf16 --gpus "$N" --steps 20 --warmup 5
step resnet50_ddp   timeout 150 $TR --master-port 29707 main.py envarg://WORLD_SIZE envarg://RANK 1 "$N" 127.0.0.1 lo resnet50 /tmp/none 1 \
                    --backend nccl --experiment main_ddp --dtype bf16 --random_input 1 --limit_batches 25 --batch_size 128 \
                    --channels_last 1 --cudnn_benchmark 1 --results_root "$OUT/results" --master_port 29750 --job_id resnet_ddp
step bench_mnist    timeout 300 $TR --master-port 29708 bench.py --model mnist --gpus "$N" --steps 200 --warmup 20
step ref_mnist_ddp  timeout 300 $TR --master-port 29709 bench.py --impl reference --model mnist --gpus "$N" --steps 50 --warmup 10
step ref_mnist_ourdist_nccl timeout 240 $TR --master-port 29710 bench.py --impl reference --model mnist --ref-experiment main_ourdist_nccl --gpus "$N" --steps 50 --warmup 10
if true; then
  step googlenet_ddp_nccl timeout 150 $TR --master-port 29711 main.py envarg://WORLD_SIZE envarg://RANK 1 "$N" 127.0.0.1 lo imagenet /tmp/none 1 \
                    --backend nccl --experiment main_ddp --random_input 1 --limit_batches 25 --batch_size 128 \
                    --channels_last 1 --cudnn_benchmark 1 --results_root "$OUT/results" --master_port 29752 --job_id googlenet_ddp
  step fusion_cifar   timeout 900 $M cifar /tmp/none 1 --backend b200 --experiment fusion_experiment_ourdist --random_input 1 \
                    --limit_batches 60 --batch_size 128 --results_root "$OUT/results" --master_port 29760 --job_id cifar
  step fusion_googlenet timeout 900 $M imagenet /tmp/none 1 --backend b200 --experiment fusion_experiment_ourdist --random_input 1 \
                    --limit_batches 14 --batch_size 128 --channels_last 1 --cudnn_benchmark 1 --results_root "$OUT/results" --master_port 29770 --job_id googlenet
  step experiment1_b200 timeout 900 $M cifar /tmp/none 1 --backend b200 --experiment experiment1 --random_input 1 \
                    --limit_batches 60 --batch_size 128 --results_root "$OUT/results" --master_port 29780 --job_id exp1
fi
python benchmarks/summarize_results.py "$OUT"/results/* --json "$OUT/results_summary.json" > "$OUT/results_summary.txt" 2>&1
cat "$OUT/results_summary.txt"
grep -l "Traceback\|Error" "$OUT"/*.err 2>/dev/null | head
