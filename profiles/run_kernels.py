#!/usr/bin/env python3
"""Launch each kernel of libdlb200 on one GPU at production sizes, timed with CUDA events.

This is the program that is captured under ncu (see profiles/README.md).  Sizes follow the
flagship config: GoogLeNet's used gradients (173 tensors, 25.3 MiB fp32) as one bucket, and a
64 MiB buffer for the bandwidth-bound limits.  Multi-rank kernels run as 4 emulated ranks in one
cooperative launch (the only way to run them on a single GPU).  Prints one JSON line.
"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_learning_b200.models import GoogLeNet  # noqa: E402
from distributed_learning_b200.ops import comm as C  # noqa: E402
from distributed_learning_b200.ops import fused as K  # noqa: E402
from distributed_learning_b200.parallel.buckets import Bucket  # noqa: E402


ONCE = "--once" in sys.argv      # one launch per kernel: the mode captured under ncu
VARIANTS = ("chunks", "tiles", "bulk")


def timed(fn, iters=20, flush=None):
    if ONCE:
        iters = 1
    for _ in range(0 if ONCE else 3):
        fn()
    ts = []
    for _ in range(iters):
        if flush is not None:
            flush.add_(1.0)                  # 512 MiB write: evicts L2 between timed launches
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        e1.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2] * 1000.0         # median, microseconds


def main():
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))
    except OSError:
        pass
    hbm = float(peaks.get("hbm_gbs", 6650.0))
    flush = torch.zeros(128 << 20, device=dev)
    out = {"hbm_gbs_denominator": hbm, "denominator_source": "measured" if peaks else "fallback", "kernels": {}}
    from bench import ClockSampler  # noqa: E402

    clocks = ClockSampler(0).start()

    # ---- K-pack / K-unpack on the gradient lists of GoogLeNet (one bucket) and ResNet-50 ---------------
    from distributed_learning_b200.models import ResNet50  # noqa: E402

    with torch.device("meta"):
        nets = {"googlenet_173t": [p.shape for n, p in GoogLeNet().named_parameters()
                                   if not n.startswith(("model.aux1", "model.aux2"))],
                "resnet50_161t": [p.shape for p in ResNet50().parameters()]}
    for net_name, shapes in nets.items():
        for name, dt in (("fp32", torch.float32), ("bf16", torch.bfloat16)):
            # tensors in the flat dtype (what the training path packs: fp32 grads -> fp32 bucket, bf16 -> bf16)
            grads = [torch.randn(s, device=dev).to(dt) for s in reversed(shapes)]
            b = Bucket(0, grads, dtype=dt, aligned=True, allocate=False)
            plan = K.PackPlan.for_tensors(grads, b.offsets)
            flat = torch.zeros(b.numel, device=dev, dtype=dt)
            nbytes = b.dense_numel * 2 * flat.element_size()
            for variant in VARIANTS:
                for scale in (1.0, 0.125):
                    tag = f"{net_name}_{name}_{variant}" + ("" if scale == 1.0 else "_scaled")
                    us = timed(lambda: plan.pack(flat, grads, scale, variant=variant), flush=flush)
                    out["kernels"][f"pack_{tag}"] = {"us": us, "bytes": nbytes, "GBs": nbytes / us / 1e3,
                                                     "frac_hbm": nbytes / us / 1e3 / hbm}
                us = timed(lambda: plan.unpack(flat, grads, 1.0, variant=variant), flush=flush)
                out["kernels"][f"unpack_{net_name}_{name}_{variant}"] = {"us": us, "bytes": nbytes, "GBs": nbytes / us / 1e3,
                                                                          "frac_hbm": nbytes / us / 1e3 / hbm}
        # mixed: fp32 gradients into a bf16 bucket (comm_dtype bf16 on an fp32 model) -- conversion in flight
        grads = [torch.randn(s, device=dev) for s in reversed(shapes)]
        b = Bucket(0, grads, dtype=torch.bfloat16, aligned=True, allocate=False)
        plan = K.PackPlan.for_tensors(grads, b.offsets)
        flat = torch.zeros(b.numel, device=dev, dtype=torch.bfloat16)
        nbytes = b.dense_numel * 6
        for variant in ("chunks", "tiles"):
            us = timed(lambda: plan.pack(flat, grads, 0.125, variant=variant), flush=flush)
            out["kernels"][f"pack_{net_name}_fp32_to_bf16_{variant}"] = {"us": us, "bytes": nbytes, "GBs": nbytes / us / 1e3,
                                                                       "frac_hbm": nbytes / us / 1e3 / hbm}
            us = timed(lambda: plan.unpack(flat, grads, 1.0, variant=variant), flush=flush)
            out["kernels"][f"unpack_{net_name}_bf16_to_fp32_{variant}"] = {"us": us, "bytes": nbytes, "GBs": nbytes / us / 1e3,
                                                                         "frac_hbm": nbytes / us / 1e3 / hbm}

    # ---- K-sgd on 64 MiB ---------------------------------------------------------------------------------
    n = 16 << 20
    p, m = torch.randn(n, device=dev), torch.zeros(n, device=dev)
    h = K.make_hyper(0.01, 0.5)
    for name, g, sh in (("fp32", torch.randn(n, device=dev), None),
                        ("bf16_shadow", torch.randn(n, device=dev).bfloat16(), torch.zeros(n, device=dev, dtype=torch.bfloat16))):
        nbytes = n * (16 + g.element_size() + (2 if sh is not None else 0))
        us = timed(lambda: K.sgd_flat(p, m, g, h, sh), flush=flush)
        out["kernels"][f"sgd_flat_16Mi_{name}"] = {"us": us, "bytes": nbytes, "GBs": nbytes / us / 1e3,
                                                   "frac_hbm": nbytes / us / 1e3 / hbm}

    # ---- K-ar, 4 emulated ranks x 16 MiB, local memory stands in for peers ----------------------------
    nr, ne = 4, 4 << 20
    bufs = [torch.randn(ne, device=dev) for _ in range(nr)]
    outs = [torch.empty_like(x) for x in bufs]
    flags = C.emulated_flags(nr, dev)
    keep = [(torch.randn(ne, device=dev), torch.zeros(ne, device=dev)) for _ in range(nr)]   # descriptors hold raw pointers
    sgds = [K.fused_sgd_args(pp, mm, h) for pp, mm in keep]
    for name, kw in (("twoshot", dict(algo="twoshot")), ("oneshot", dict(algo="oneshot", outs=outs)),
                     ("twoshot_fused_sgd", dict(algo="twoshot", sgds=sgds)),
                     ("oneshot_fused_sgd", dict(algo="oneshot", outs=outs, sgds=sgds))):
        us = timed(lambda: C.emulated_allreduce(bufs, scale=0.25, flags=flags, nblocks=32, nthreads=512, **kw), flush=flush)
        out["kernels"][f"allreduce_emulated_4x16MiB_{name}"] = {"us": us, "note": "single GPU: HBM stands in for NVLink"}
    del keep
    out["clocks"] = clocks.stop()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
