#!/bin/bash
out=gpurun_out/${1:-r2p}
mkdir -p $out
python benchmarks/bench_pack.py --tag default > $out/pack_default.json 2> $out/pack.err
DLB200_PACK_U=4 python benchmarks/bench_pack.py --variants tiles --tag U4 > $out/pack_tiles_u4.json 2>> $out/pack.err
DLB200_BULK_CTAS_PER_SM=2 python benchmarks/bench_pack.py --variants bulk --tag bulk2 > $out/pack_bulk2.json 2>> $out/pack.err
python - <<PY
import json,glob
for f in sorted(glob.glob("$out/pack_*.json")):
    try: d=json.load(open(f))
    except Exception as e: print(f, e); continue
    print(f, d.get("clocks"))
    for r in d["rows"]:
        print(f"  {r['net']:16s} {r['tensors']}->{r['flat']} {r['op']:6s} {r['variant']:6s} s={r['scale']:<5} {r['us']:7.2f} us {r['GBs']:7.0f} GB/s {100*r['frac_hbm']:5.1f}%")
PY
tail -3 $out/pack.err
python profiles/pack_once.py > $out/once_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:pack_ -c 12 -o $out/pack_variants python profiles/pack_once.py > $out/ncu.log 2>&1
tail -3 $out/ncu.log
