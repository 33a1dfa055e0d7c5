#!/bin/bash
# ncu on rank 0 of an N-GPU job (the other ranks run plain): NVLink bytes of the real kernels.  $1 = N, $2 = tag
N=${1:-2}; out=gpurun_out/${2:-r2nvl}
mkdir -p $out
rm -f /tmp/nvl_run_*
export NVL_WORLD=$N NVL_PORT=29700 DLB200_WAIT_TIMEOUT_MS=15000
# plain run of the whole job first (also what ncu's own pre-run expects to have succeeded)
for r in $(seq 0 $((N-1))); do NVL_RANK=$r timeout 120 python profiles/nvlink_worker.py > $out/plain_$r.log 2>&1 & done
wait
grep NVL_SEQUENCE $out/plain_0.log | cut -c1-300
# profiled job: peers may be started twice, because ncu may run rank 0's command once before profiling it
for r in $(seq 1 $((N-1))); do
  ( NVL_RANK=$r timeout 200 python profiles/nvlink_worker.py > $out/peer_${r}_a.log 2>&1; NVL_RANK=$r timeout 100 python profiles/nvlink_worker.py > $out/peer_${r}_b.log 2>&1 ) &
done
NVL_RANK=0 timeout 400 ncu --metrics gpu__time_duration.sum,nvlrx__bytes.sum,nvltx__bytes.sum,nvlrx__bytes_data_user.sum,nvltx__bytes_data_user.sum,dram__bytes_read.sum,dram__bytes_write.sum \
   --clock-control none --cache-control none -k regex:'twoshot|oneshot|nvls' -c 8 --csv --log-file $out/rank0_nvlink.csv \
   python profiles/nvlink_worker.py > $out/rank0_ncu.log 2>&1
echo "ncu exit $?"
wait
tail -5 $out/rank0_ncu.log
head -c 3000 $out/rank0_nvlink.csv
