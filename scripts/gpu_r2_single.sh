#!/bin/bash
# One-GPU check of the tree: GPU tests, smoke, both bench arms.  Results under gpurun_out/$1.
out=gpurun_out/${1:-r2a}
mkdir -p $out
nvidia-smi --query-gpu=index,name,clocks.sm,clocks.max.sm,power.draw --format=csv > $out/smi.txt 2>&1
nproc > $out/nproc.txt
timeout 1500 python -m pytest tests -m gpu -x -q > $out/pytest.log 2>&1; echo "pytest exit $?" >> $out/pytest.log
tail -5 $out/pytest.log
timeout 300 python __graft_entry__.py smoke > $out/smoke.log 2>&1; echo "smoke exit $?" >> $out/smoke.log
tail -3 $out/smoke.log
timeout 600 python bench.py --steps 20 --warmup 5 > $out/bench_b200.json 2> $out/bench_b200.err; echo "bench exit $?"
cat $out/bench_b200.json
timeout 600 python bench.py --impl reference --steps 10 --warmup 3 > $out/bench_ref.json 2> $out/bench_ref.err; echo "ref exit $?"
cat $out/bench_ref.json
