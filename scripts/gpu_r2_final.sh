#!/bin/bash
# Final one-GPU validation of the tree: the whole GPU test suite, smoke, the N=1 benches.
out=gpurun_out/${1:-r2final}; mkdir -p $out
timeout 600 python -m pytest tests -m gpu -x -q > $out/pytest.log 2>&1; echo "pytest exit $?" | tee -a $out/pytest.log; tail -3 $out/pytest.log
timeout 200 python __graft_entry__.py smoke > $out/smoke.log 2>&1; echo "smoke exit $?"; tail -1 $out/smoke.log
timeout 200 python bench.py --steps 20 --warmup 5 > $out/bench_b200.json 2> $out/bench_b200.err; echo "bench exit $?"; cut -c1-400 $out/bench_b200.json
timeout 200 python bench.py --model resnet50 --dtype bf16 --steps 20 --warmup 5 > $out/bench_resnet50.json 2> $out/bench_resnet50.err; echo "resnet exit $?"; cut -c1-300 $out/bench_resnet50.json
timeout 120 python bench.py --model mnist --steps 200 --warmup 20 > $out/bench_mnist.json 2> $out/bench_mnist.err; echo "mnist exit $?"; cut -c1-300 $out/bench_mnist.json
timeout 120 python bench.py --model cifar --steps 200 --warmup 20 > $out/bench_cifar.json 2> $out/bench_cifar.err; echo "cifar exit $?"; cut -c1-300 $out/bench_cifar.json
