#!/usr/bin/env python3
"""Allreduce bandwidth sweep (BASELINE.json config 5): 1 KB - 1 GB at N GPUs, the native kernels
against NCCL and against the NVLink roofline.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        benchmarks/bench_allreduce.py [--max-bytes 1073741824] [--dtype fp32] [--out file.json]

For every size each variant (one-shot, two-shot, nvls when the box has multicast, NCCL
``all_reduce``, and with ``--ring`` the reference-style host-driven NCCL ring) is launched
``iters`` times back to back on one stream between two CUDA events after ``warmup`` launches;
the time is the max over ranks.  Bus bandwidth = 2 (N-1)/N * bytes / time, the bytes each GPU
must move over its links per direction; roofline fractions are quoted against 900 GB/s nominal
and the 770 GB/s measured peer copy (B200_PROFILING.md).  Peer memory is not cached in the
reader's L2, so repeated launches on one buffer stay link-bound; the buffer alternates between
two arena regions anyway.  The reference never measured this.
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def timed(fn, warmup, iters, device, graph=False):
    """ms per call of ``fn``: ``iters`` back-to-back launches between two events, max over ranks.
    ``graph=True`` captures the ``iters`` launches in one CUDA graph first, which removes the host's
    per-launch cost (Python + driver) and leaves the device-side time."""
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize(device)
    replay = None
    if graph:
        side = torch.cuda.Stream(device=device)
        side.wait_stream(torch.cuda.current_stream(device))
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=side):
            for _ in range(iters):
                fn()
        replay = g.replay
        replay()
        torch.cuda.synchronize(device)
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    if replay is not None:
        replay()
    else:
        for _ in range(iters):
            fn()
    e1.record()
    torch.cuda.synchronize(device)
    t = torch.tensor([e0.elapsed_time(e1) / iters], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t[0])


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--min-bytes", type=int, default=1024)
    ap.add_argument("--max-bytes", type=int, default=1 << 30)
    ap.add_argument("--factor", type=int, default=4)
    ap.add_argument("--dtype", default="fp32", choices=("fp32", "bf16"))
    ap.add_argument("--blocks", default="8,16,32,48,64")
    ap.add_argument("--unrolls", default="4,8", help="nvls pipeline depths to try (2, 4, 8)")
    ap.add_argument("--algos", default=None, help="comma list; default: every variant the box supports")
    ap.add_argument("--no-nccl", action="store_true", help="skip the NCCL comparison rows")
    ap.add_argument("--ring", action="store_true", help="also time the reference-style host-driven NCCL send/recv ring")
    ap.add_argument("--graph", action="store_true", help="time CUDA-graph replays of the launches (device-side time, no host launch cost)")
    ap.add_argument("--out", default=None)
    args = ap.parse_args()

    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    device = torch.device(f"cuda:{int(os.environ.get('LOCAL_RANK', rank))}")
    torch.cuda.set_device(device)
    dist.init_process_group("cpu:gloo,cuda:nccl", rank=rank, world_size=world)
    from distributed_learning_b200.ops.comm import NativeComm
    from distributed_learning_b200.parallel.allreduce import ring_allreduce_gpu

    dtype = torch.float32 if args.dtype == "fp32" else torch.bfloat16
    esize = 4 if dtype == torch.float32 else 2
    comm = NativeComm(device=device)
    region = (args.max_bytes + 4095) // 4096 * 4096
    comm.allocate(2 * region)
    if rank == 0:
        print(f"arena {comm.mode}, multicast {comm.has_multicast} {comm.mc_note}", flush=True)
    blocks = [int(b) for b in args.blocks.split(",")]
    unrolls = [int(u) for u in args.unrolls.split(",")]
    algos = ["oneshot", "twoshot", "twoshot_bulk"] + (["nvls"] if comm.has_multicast else [])
    if args.algos:
        algos = [a for a in args.algos.split(",") if a != "nvls" or comm.has_multicast]
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from bench import ClockSampler

    clocks = ClockSampler(device.index).start()
    trace = torch.zeros(8, dtype=torch.int64, device=device)
    rows = []
    size = args.min_bytes
    while size <= args.max_bytes:
        numel = size // esize
        iters = 200 if size <= 1 << 20 else (50 if size <= 64 << 20 else 10)
        row = {"bytes": size, "ms": {}, "best": None}
        runners = {}
        comm.view(0, numel, dtype).normal_()
        comm.view(region, numel, dtype).normal_()
        for algo in algos:
            if algo == "oneshot" and size > 64 << 20:
                continue
            for nb in blocks:
                if size <= 64 << 10 and nb > 16:
                    continue
                for un in (unrolls if algo == "nvls" else [0]):
                    if algo == "nvls" and size // world <= 512 * 16 * 2 * nb and un > min(unrolls):
                        continue        # one tile per CTA already covers the slice: deeper unroll changes nothing
                    flip = [0]

                    def run(algo=algo, nb=nb, un=un, tr=None):
                        a, b = (0, region) if flip[0] == 0 else (region, 0)
                        flip[0] ^= 1
                        comm.allreduce(a, numel, dtype, algo, scale=1.0 / world, out_offset=b, nblocks=nb,
                                       stream=torch.cuda.current_stream(device), unroll=un, trace=tr)

                    row["ms"][f"{algo}/{nb}" + (f"/u{un}" if un else "")] = timed(run, 5, iters, device, args.graph)
                    runners[f"{algo}/{nb}" + (f"/u{un}" if un else "")] = run
        x = torch.randn(numel, device=device).to(dtype)
        if args.no_nccl:
            row["ms"]["nccl"] = float("nan")
        else:
            row["ms"]["nccl"] = timed(lambda: dist.all_reduce(x), 5, iters, device, args.graph)
        if args.ring and size >= 4096:
            row["ms"]["nccl_ring_hostdriven"] = timed(lambda: ring_allreduce_gpu(x), 2, max(2, iters // 10), device)
        native = {k: v for k, v in row["ms"].items() if not k.startswith("nccl")}
        best = min(native, key=native.get)
        bus = 2 * (world - 1) / world * size
        row["best"] = best
        row["best_ms"] = native[best]
        row["busbw_GBs"] = bus / native[best] / 1e6
        row["nccl_busbw_GBs"] = bus / row["ms"]["nccl"] / 1e6
        row["frac_of_900"] = row["busbw_GBs"] / 900.0
        row["frac_of_770_measured_copy"] = row["busbw_GBs"] / 770.0
        row["speedup_vs_nccl"] = row["ms"]["nccl"] / native[best]
        # where the time of the best variant goes: %globaltimer at the phase boundaries of CTA 0, rank 0
        dist.barrier()
        runners[best](tr=trace)
        torch.cuda.synchronize(device)
        t = trace.cpu().tolist()
        row["phases_us_rank0"] = {k: (t[i + 1] - t[i]) / 1000.0 for i, k in
                                  enumerate(("pack", "entry_barrier", "data", "exit_barrier", "epilogue"))}
        rows.append(row)
        if rank == 0:
            print(f"{size:>11d} B  best {best:<14s} {native[best] * 1000:9.1f} us  busbw {row['busbw_GBs']:7.1f} GB/s "
                  f"({row['frac_of_900'] * 100:5.1f}% of 900)  nccl {row['ms']['nccl'] * 1000:9.1f} us "
                  f"{row['nccl_busbw_GBs']:7.1f} GB/s  x{row['speedup_vs_nccl']:.2f}", flush=True)
        size *= args.factor
    comm.check()
    result = {"n_gpus": world, "dtype": args.dtype, "arena": comm.mode, "multicast": comm.has_multicast,
              "clocks": clocks.stop(),
              "timing": ("CUDA events around one CUDA-graph replay of the launches" if args.graph else
                         "CUDA events, back-to-back launches") + ", max over ranks", "rows": rows}
    if rank == 0 and args.out:
        os.makedirs(os.path.dirname(os.path.abspath(args.out)), exist_ok=True)
        with open(args.out, "w") as f:
            json.dump(result, f, indent=1)
    comm.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
