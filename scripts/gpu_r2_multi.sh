#!/bin/bash
# N GPUs: multi-GPU correctness worker, allreduce sweeps (eager + graph), bench.  $1 = N, $2 = out tag, $3 = what
N=${1:-2}; out=gpurun_out/${2:-r2m}; what=${3:-test,sweep,bench}
mkdir -p $out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
port=29610
if [[ $what == *test* ]]; then
  timeout 900 $TR --master-port $((port++)) tests/_mp_worker.py all > $out/mp_worker.log 2>&1; echo "mp_worker exit $?" | tee -a $out/mp_worker.log
  tail -4 $out/mp_worker.log
fi
if [[ $what == *sweep* ]]; then
  for dt in fp32 bf16; do
    timeout 600 $TR --master-port $((port++)) benchmarks/bench_allreduce.py --dtype $dt --out $out/allreduce_${dt}_eager.json > $out/allreduce_${dt}_eager.txt 2>&1; echo "sweep $dt eager exit $?"
    timeout 600 $TR --master-port $((port++)) benchmarks/bench_allreduce.py --dtype $dt --graph --out $out/allreduce_${dt}_graph.json > $out/allreduce_${dt}_graph.txt 2>&1; echo "sweep $dt graph exit $?"
  done
  cat $out/allreduce_fp32_eager.txt | tail -14
  cat $out/allreduce_fp32_graph.txt | tail -12
fi
if [[ $what == *bench* ]]; then
  timeout 600 $TR --master-port $((port++)) bench.py --gpus $N --steps 20 --warmup 5 > $out/bench_b200.json 2> $out/bench_b200.err; echo "bench exit $?"
  cat $out/bench_b200.json | cut -c1-1500
  timeout 600 $TR --master-port $((port++)) bench.py --gpus $N --steps 20 --warmup 5 --fuse-pack 0 --no-e2e > $out/bench_b200_nofuse.json 2> $out/bench_b200_nofuse.err; echo "bench nofuse exit $?"
  cat $out/bench_b200_nofuse.json | cut -c1-400
fi
