#!/usr/bin/env python3
"""Pack / unpack kernel variants on the gradient lists of GoogLeNet and ResNet-50, one GPU.

Timing: CUDA events around one replay of a CUDA graph that holds `iters` back-to-back launches
rotating over several independent sets of (tensors, flat buffer) whose total size is a multiple
of the 126 MB L2, so every launch reads and writes memory that is not cached; the per-launch time
is the total divided by `iters` (event resolution and host launch overhead are out of the picture).
Fractions are of the measured HBM copy bandwidth in MEASURED_PEAKS.json.  Prints one JSON line.

    python benchmarks/bench_pack.py [--variants chunks,tiles,bulk] [--iters 40]
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import ClockSampler  # noqa: E402
from distributed_learning_b200.models import GoogLeNet, ResNet50  # noqa: E402
from distributed_learning_b200.ops import fused as K  # noqa: E402
from distributed_learning_b200.parallel.buckets import Bucket  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--variants", default="chunks,tiles,bulk")
    ap.add_argument("--iters", type=int, default=40)
    ap.add_argument("--tag", default="")
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    try:
        hbm = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
        src = "measured"
    except (OSError, KeyError, ValueError):
        hbm, src = 6650.0, "fallback"
    with torch.device("meta"):
        nets = {"googlenet_173t": [p.shape for n, p in GoogLeNet().named_parameters()
                                   if not n.startswith(("model.aux1", "model.aux2"))],
                "resnet50_161t": [p.shape for p in ResNet50().parameters()]}
    out = {"hbm_gbs_denominator": hbm, "denominator_source": src, "tag": args.tag, "rows": [],
           "timing": "events around back-to-back launches rotating over sets larger than L2"}
    clocks = ClockSampler(0).start()
    for net, shapes in nets.items():
        for tname, tdt, fname, fdt in (("fp32", torch.float32, "fp32", torch.float32),
                                       ("bf16", torch.bfloat16, "bf16", torch.bfloat16),
                                       ("fp32", torch.float32, "bf16", torch.bfloat16)):
            one = sum(int(torch.Size(s).numel()) for s in shapes) * (torch.empty((), dtype=tdt).element_size()
                                                                      + torch.empty((), dtype=fdt).element_size())
            nsets = max(3, -(-(400 << 20) // one))
            sets = []
            for _ in range(nsets):
                grads = [torch.randn(s, device=dev).to(tdt) for s in reversed(shapes)]
                b = Bucket(0, grads, dtype=fdt, aligned=True, allocate=False)
                sets.append((K.PackPlan.for_tensors(grads, b.offsets), grads, torch.zeros(b.numel, device=dev, dtype=fdt)))
            nbytes = one
            for variant in args.variants.split(","):
                if variant == "bulk" and tdt != fdt:
                    continue
                for op in ("pack", "unpack"):
                    for scale in ((1.0, 0.125) if op == "pack" else (1.0,)):
                        def run(i):
                            plan, grads, flat = sets[i % nsets]
                            if op == "pack":
                                plan.pack(flat, grads, scale, variant=variant)
                            else:
                                plan.unpack(flat, grads, scale, variant=variant)
                        for i in range(nsets):
                            run(i)
                        torch.cuda.synchronize()
                        # the launches are replayed from a CUDA graph: the host cannot keep a ~10 us
                        # kernel fed from Python (173 pointers are refreshed per call)
                        side = torch.cuda.Stream()
                        graph = torch.cuda.CUDAGraph()
                        with torch.cuda.stream(side):
                            with torch.cuda.graph(graph, stream=side):
                                for i in range(args.iters):
                                    run(i)
                        graph.replay()
                        torch.cuda.synchronize()
                        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                        e0.record()
                        graph.replay()
                        e1.record()
                        e1.synchronize()
                        us = e0.elapsed_time(e1) * 1000.0 / args.iters
                        del graph
                        out["rows"].append({"net": net, "tensors": tname, "flat": fname, "op": op, "variant": variant,
                                            "scale": scale, "us": round(us, 2), "bytes": nbytes,
                                            "GBs": round(nbytes / us / 1e3, 1), "frac_hbm": round(nbytes / us / 1e3 / hbm, 4)})
            del sets
            torch.cuda.empty_cache()
    out["clocks"] = clocks.stop()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
