#!/usr/bin/env bash
# The measurement suite behind BENCH_NOTES.md: correctness at N ranks, the flagship bench (both
# arms, and the reference's own NCCL ring), the allreduce sweeps (fp32/bf16 x eager/graph, clocks
# and phase traces in every JSON), ResNet-50 bf16 against DDP-autocast, the small-CNN configs and
# the fusion sweep.
#   [SKIP="step names"] [ONLY="step names"] scripts/gpu_suite.sh <N> [tag]
# Everything lands in gpurun_out/<tag>_n<N>/ (one .out/.err per step).
set -u
N=${1:-2}; TAG=${2:-r2_suite}
ROOT="$(cd "$(dirname "$0")/.." && pwd)"; cd "$ROOT"
OUT=gpurun_out/${TAG}_n$N; mkdir -p "$OUT"
SKIP=" ${SKIP:-} "; ONLY=" ${ONLY:-} "
# run <seconds> <port> <script and arguments>: under torchrun for N > 1, plain python for N = 1
run() { local secs=$1 port=$2; shift 2
        if [ "$N" -gt 1 ]; then timeout "$secs" python -m torch.distributed.run --nnodes=1 --nproc-per-node "$N" --master-addr 127.0.0.1 --master-port "$port" "$@"
        else timeout "$secs" python "$@"; fi; }
step() { local name=$1; shift
         case "$SKIP" in *" $name "*) echo "[$name] skipped"; return;; esac
         if [ "$ONLY" != "  " ]; then case "$ONLY" in *" $name "*) ;; *) return;; esac; fi
         local t0; t0=$(date +%s); "$@" > "$OUT/$name.out" 2> "$OUT/$name.err"; local rc=$?
         echo "[$name] rc=$rc $(( $(date +%s) - t0 ))s"; tail -n 1 "$OUT/$name.out" | cut -c1-300; }
M="python main.py 1 0 $N $N 127.0.0.1 lo"
SW="--blocks 16,32,64 --min-bytes 4096"

[ "$N" -gt 1 ] && step mp_worker      run 600 29701 tests/_mp_worker.py all
step bench_b200     run 300 29702 bench.py --gpus "$N" --steps 20 --warmup 5
step bench_ref      run 500 29703 bench.py --impl reference --gpus "$N" --steps 10 --warmup 3
if [ "$N" -gt 1 ]; then
step ar_fp32_eager  run 400 29704 benchmarks/bench_allreduce.py $SW --out "$OUT/allreduce_fp32_eager.json"
step ar_fp32_graph  run 400 29705 benchmarks/bench_allreduce.py $SW --graph --out "$OUT/allreduce_fp32_graph.json"
step ar_bf16_graph  run 400 29706 benchmarks/bench_allreduce.py $SW --dtype bf16 --graph --out "$OUT/allreduce_bf16_graph.json"
step ar_bf16_eager  run 400 29707 benchmarks/bench_allreduce.py $SW --dtype bf16 --out "$OUT/allreduce_bf16_eager.json"
step bench_ref_ourdist_nccl run 500 29708 bench.py --impl reference --ref-experiment main_ourdist_nccl --gpus "$N" --steps 10 --warmup 3
fi
step bench_resnet50 run 300 29709 bench.py --model resnet50 --dtype bf16 --gpus "$N" --steps 20 --warmup 5
if [ "$N" -gt 1 ]; then
step resnet50_ddp   run 200 29710 main.py envarg://WORLD_SIZE envarg://RANK 1 "$N" 127.0.0.1 lo resnet50 /tmp/none 1 \
                    --backend nccl --experiment main_ddp --dtype bf16 --random_input 1 --limit_batches 25 --batch_size 128 \
                    --channels_last 1 --cudnn_benchmark 1 --results_root "$OUT/results" --master_port 29750 --job_id resnet_ddp
step googlenet_ddp  run 200 29711 main.py envarg://WORLD_SIZE envarg://RANK 1 "$N" 127.0.0.1 lo imagenet /tmp/none 1 \
                    --backend nccl --experiment main_ddp --random_input 1 --limit_batches 25 --batch_size 128 \
                    --channels_last 1 --cudnn_benchmark 1 --results_root "$OUT/results" --master_port 29752 --job_id googlenet_ddp
fi
step bench_mnist    run 300 29712 bench.py --model mnist --gpus "$N" --steps 200 --warmup 20
step bench_cifar    run 300 29713 bench.py --model cifar --gpus "$N" --steps 200 --warmup 20
step ref_mnist_ddp  run 300 29714 bench.py --impl reference --model mnist --gpus "$N" --steps 50 --warmup 10
if [ "$N" -gt 1 ]; then
step ref_mnist_ourdist_nccl run 300 29715 bench.py --impl reference --model mnist --ref-experiment main_ourdist_nccl --gpus "$N" --steps 50 --warmup 10
step fusion_googlenet timeout 900 $M imagenet /tmp/none 1 --backend b200 --experiment fusion_experiment_ourdist --random_input 1 \
                    --limit_batches 14 --batch_size 128 --channels_last 1 --cudnn_benchmark 1 --results_root "$OUT/results" --master_port 29770 --job_id googlenet
step fusion_cifar   timeout 900 $M cifar /tmp/none 1 --backend b200 --experiment fusion_experiment_ourdist --random_input 1 \
                    --limit_batches 60 --batch_size 128 --results_root "$OUT/results" --master_port 29760 --job_id cifar
fi
ls "$OUT"/results >/dev/null 2>&1 && python benchmarks/summarize_results.py "$OUT"/results/* --json "$OUT/results_summary.json" > "$OUT/results_summary.txt" 2>&1 && cat "$OUT/results_summary.txt" | tail -40
grep -l "Traceback\|Error" "$OUT"/*.err 2>/dev/null | head
