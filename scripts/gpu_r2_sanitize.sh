#!/bin/bash
# $1 = tool (memcheck | racecheck | synccheck), results in gpurun_out/r2_sanitize/
out=gpurun_out/r2_sanitize; mkdir -p $out
python profiles/sanitize_once.py > $out/plain.log 2>&1 && \
timeout 600 compute-sanitizer --tool $1 --log-file $out/$1.log python profiles/sanitize_once.py > $out/$1.out 2>&1
echo "sanitizer exit $?"; tail -5 $out/$1.out; tail -12 $out/$1.log
