#!/bin/bash
# ncu on rank 0 of an N-GPU job (the other ranks run plain): NVLink bytes of the real kernels.  $1 = N, $2 = tag
N=${1:-2}; out=gpurun_out/${2:-r2nvl}
mkdir -p $out
export NVL_WORLD=$N DLB200_WAIT_TIMEOUT_MS=15000
M_NVL=gpu__time_duration.sum,nvlrx__bytes.sum,nvltx__bytes.sum,nvlrx__bytes_data_user.sum,nvltx__bytes_data_user.sum

plain() {   # the whole job without a profiler
  rm -f /tmp/nvl_run_*; export NVL_PORT=$1
  for r in $(seq 0 $((N-1))); do NVL_RANK=$r timeout 120 python profiles/nvlink_worker.py > $out/plain_$r.log 2>&1 & done
  wait
  grep -c NVL_SEQUENCE $out/plain_0.log
}

attempt() {  # $1 = name, $2 = port base, rest = ncu arguments
  local name=$1 port=$2; shift 2
  rm -f /tmp/nvl_run_*; export NVL_PORT=$port
  NVL_RANK=0 timeout 400 ncu "$@" --clock-control none --cache-control none -k regex:'twoshot|oneshot|nvls' -c 8 \
     --csv --log-file $out/${name}_rank0.csv python profiles/nvlink_worker.py > $out/${name}_rank0.log 2>&1 &
  local ncu_pid=$!
  # peers: ncu may run rank 0's command more than once, so a peer starts again for as long as ncu is alive
  for r in $(seq 1 $((N-1))); do
    ( i=0; while kill -0 $ncu_pid 2>/dev/null && [ $i -lt 3 ]; do i=$((i+1))
        NVL_RANK=$r timeout 150 python profiles/nvlink_worker.py > $out/${name}_peer_${r}_$i.log 2>&1 || break; done ) &
  done
  wait $ncu_pid; echo "$name: ncu exit $?"
  wait
  tail -3 $out/${name}_rank0.log
  head -c 1500 $out/${name}_rank0.csv; echo
}

plain 29700
attempt kern_time 29740 --metrics gpu__time_duration.sum
attempt kern_nvl 29760 --metrics nvlrx__bytes.sum,nvltx__bytes.sum
