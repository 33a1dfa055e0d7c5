#!/bin/bash
# One GPU: kernel tests + graph-timed pack table.
out=gpurun_out/${1:-r2k}
mkdir -p $out
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_fused_pack_gpu.py tests/test_allreduce_gpu.py tests/test_zz_no_alloc.py tests/test_graph_parity_gpu.py tests/test_native_options_gpu.py -x -q > $out/pytest.log 2>&1; echo "pytest exit $?" >> $out/pytest.log
tail -15 $out/pytest.log
python benchmarks/bench_pack.py --tag default > $out/pack_default.json 2> $out/pack.err
python - <<PY
import json
d=json.load(open("$out/pack_default.json")); print(d.get("clocks"))
for r in d["rows"]:
    print(f"  {r['net']:16s} {r['tensors']}->{r['flat']} {r['op']:6s} {r['variant']:6s} s={r['scale']:<5} {r['us']:7.2f} us {r['GBs']:7.0f} GB/s {100*r['frac_hbm']:5.1f}%")
PY
tail -3 $out/pack.err
