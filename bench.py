#!/usr/bin/env python3
"""Headline benchmark: training images/sec of the reference's flagship config on N B200s.

    python bench.py --gpus N --steps K --warmup W            (N=1)
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W          (N>1, one rank per GPU)
    ... --impl reference      runs the UNMODIFIED reference (baseline/_ref) for the same config

Config (the reference's `experiment2` suite): GoogLeNet with aux heads,
batch 128 per GPU (weak scaling), fp32 parameters, SGD lr 0.01 momentum 0.5, NLL loss,
synthetic ImageNet-shaped input, random-init weights.  `--model resnet50 --dtype bf16` and
`--model mnist` select the other BASELINE.json configs.

Rank 0 prints ONE JSON line.  `value` is device-timed (CUDA events around exactly K steps,
barrier + synchronize on both sides, max over ranks) with inputs resident on the device and
rotating over a pool larger than L2.  `e2e` is the same metric through the public API with
every step's input copied from pinned host memory and the loss read back to the host.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# DDP images/sec of the reference on P100s (gloo), GoogLeNet batch 128/GPU
PUBLISHED_DDP = {1: 299.0, 2: 575.2, 4: 1106.5, 8: 2198.2, 16: 4362.5}
# One metric string for BOTH arms (the driver compares them field by field); how each arm times its
# steps is in config.timing.
METRIC = "images/sec (training throughput, data parallel, max over ranks)"
MODEL_ALIASES = {"googlenet": "imagenet", "imagenet": "imagenet", "resnet50": "resnet50", "mnist": "mnist",
                 "basicnet": "mnist", "cifar": "cifar"}


def parse_args():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=20)
    p.add_argument("--warmup", type=int, default=5)
    p.add_argument("--impl", default="b200", choices=("b200", "reference"))
    p.add_argument("--model", default="googlenet", choices=sorted(MODEL_ALIASES))
    p.add_argument("--batch-size", type=int, default=128, help="per GPU")
    p.add_argument("--dtype", default=None, choices=("fp32", "bf16"))
    p.add_argument("--grouping-size", type=int, default=25 * 1024 * 1024)
    p.add_argument("--algo", default="auto")
    p.add_argument("--fused-sgd", type=int, default=1)
    p.add_argument("--fuse-pack", type=int, default=1, help="gather the gradients in the allreduce kernel's prologue (one launch per bucket)")
    p.add_argument("--overlap", type=int, default=1)
    p.add_argument("--tune", type=int, default=1, help="measure the allreduce variants once and pick per bucket size")
    p.add_argument("--channels-last", type=int, default=1, help="NHWC activations/weights (layout only; same arithmetic)")
    p.add_argument("--cudnn-benchmark", type=int, default=1, help="let cuDNN time its algorithms at the first step")
    p.add_argument("--cuda-graph", type=int, default=1, help="capture the whole training step in a CUDA graph")
    p.add_argument("--no-e2e", action="store_true")
    p.add_argument("--ref-experiment", default="main_ddp")
    p.add_argument("--ref-backend", default="nccl")
    p.add_argument("--ref-use-gpu", type=int, default=1, help="0: run the reference arm on the CPU (dry run of the plumbing)")
    return p.parse_args()


class ClockSampler:
    """`nvidia-smi` clocks and throttle reasons sampled every 200 ms while the timed region runs."""

    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")
    NAMES = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")

    def __init__(self, gpu_index: int):
        self.rows = []
        self.proc = None
        self.gpu = gpu_index
        self.thread = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits", "-lms", "200",
                 "-i", str(self.gpu)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None
            return self
        self.thread = threading.Thread(target=self._read, daemon=True)
        self.thread.start()
        return self

    def _read(self):
        for line in self.proc.stdout:
            parts = [x.strip() for x in line.split(",")]
            if len(parts) >= 7:
                self.rows.append(parts)

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        if self.thread:
            self.thread.join(timeout=2)
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
            except ValueError:
                continue
            for name, v in zip(self.NAMES, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def baseline_fields(value, e2e_value, model_type, per_gpu_batch, world, dtype) -> dict:
    """`vs_baseline` is value / the number BASELINE.md publishes for this config: the reference's DDP
    run on P100s over gloo (section 2) -- mostly a statement about hardware generations.  The
    same-hardware context sits next to it: the unmodified reference and torch DDP-NCCL as measured
    on this pool's B200 box (benchmarks/same_hw_reference.json)."""
    out = {"vs_baseline": None, "vs_baseline_is": "value / BASELINE.md section 2 (reference DDP on P100 + gloo, published CSVs)"}
    if not (model_type == "imagenet" and per_gpu_batch == 128 and dtype == "fp32"):
        return out
    if world in PUBLISHED_DDP:
        out["vs_baseline"] = value / PUBLISHED_DDP[world]
    try:
        with open(os.path.join(ROOT, "benchmarks", "same_hw_reference.json")) as f:
            same = json.load(f)["googlenet_fp32_b128"]
    except (OSError, KeyError, ValueError):
        return out
    ref = same["reference_main_ddp_nccl_images_per_s"].get(str(world))
    ddp = same["torch_ddp_nccl_same_layout_images_per_s"].get(str(world))
    if ref:
        out["vs_reference_same_hw"] = value / ref
        if e2e_value:
            out["e2e_vs_reference_same_hw"] = e2e_value / ref
    if ddp:
        out["vs_ddp_nccl_same_hw"] = value / ddp
    out["same_hw_source"] = "benchmarks/same_hw_reference.json (round-1 driver run on this pool)"
    return out


def dist_env():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))


def max_over_ranks(value: float, device=None) -> float:
    import torch
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return value
    t = torch.tensor([value], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)       # cpu tensor: goes over gloo
    return float(t[0])


def main() -> int:
    args = parse_args()
    helpers = sys.modules[__name__]
    rank = dist_env()[0]
    try:
        if args.impl == "reference":
            from benchmarks import bench_reference as arm
        else:
            from benchmarks import bench_native as arm
        result = arm.run(args, helpers)
    except SystemExit:
        raise
    except Exception as e:  # noqa: BLE001
        if args.impl == "reference":
            result = {"impl": "reference", "unavailable": f"{type(e).__name__}: {e}"[:300]}
        else:
            raise
    if rank == 0:
        print(json.dumps(result), flush=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
