#!/usr/bin/env python3
"""A short, fixed sequence of allreduce launches on N GPUs, one process per GPU, for profiling ONE
rank under ncu (scripts/gpu_r2_ncu_multi.sh wraps rank 0 only; the other ranks run plain).

Every rank issues the same launches, so the wrapped rank's kernels meet their peers' at the
device-side barriers while ncu reads the NVLink counters of its GPU.  Only counters that fit in
one pass are collected: a second replay of a kernel that waits for peers could not complete.
Rendezvous uses the port  $NVL_PORT + (how many times this rank's worker has been started), because
ncu may run its command once before profiling it.
"""
import json
import os
import sys
import datetime

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    rank, world = int(os.environ["NVL_RANK"]), int(os.environ["NVL_WORLD"])
    counter = f"/tmp/nvl_run_{rank}"
    idx = int(open(counter).read()) if os.path.exists(counter) else 0
    open(counter, "w").write(str(idx + 1))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(int(os.environ.get("NVL_PORT", 29700)) + idx)
    dev = torch.device(f"cuda:{rank}")
    torch.cuda.set_device(dev)
    dist.init_process_group("gloo", rank=rank, world_size=world, timeout=datetime.timedelta(seconds=60))
    from distributed_learning_b200.ops import fused as K
    from distributed_learning_b200.ops.comm import NativeComm

    comm = NativeComm(device=dev)
    big = 64 << 20
    comm.allocate(2 * big)
    comm.barrier()
    torch.cuda.synchronize()
    dist.barrier()
    h = K.make_hyper(0.01, 0.5)
    n16 = (16 << 20) // 4
    p, m = torch.zeros(n16, device=dev), torch.zeros(n16, device=dev)
    grads = [torch.randn(n16 // 4, device=dev) for _ in range(4)]
    plan = K.PackPlan.for_tensors(grads, [i * (n16 // 4) for i in range(4)])
    seq = [("twoshot", big // 4, 48, 0, None, None), ("oneshot", (1 << 20) // 4, 16, 0, None, None),
           ("twoshot", n16, 32, 0, K.fused_sgd_args(p, m, h), plan)]
    if comm.has_multicast:
        seq += [("nvls", big // 4, 48, 8, None, None), ("nvls", (4 << 20) // 4, 32, 4, None, None),
                ("nvls", n16, 32, 8, K.fused_sgd_args(p, m, h), plan)]
    log = []
    for algo, numel, nb, un, sgd, pk in seq:
        comm.view(0, numel, torch.float32).normal_()
        torch.cuda.synchronize()
        dist.barrier()
        comm.allreduce(0, numel, torch.float32, algo, scale=1.0 / world, out_offset=big, nblocks=nb, unroll=un, sgd=sgd,
                       pack=pk.bind(grads) if pk is not None else None, pack_scale=1.0 / world)
        torch.cuda.synchronize()
        log.append({"algo": algo, "bytes": numel * 4, "nblocks": nb, "unroll": un, "fused_sgd": sgd is not None,
                    "fused_pack": pk is not None})
    try:
        comm.check()
    except Exception as e:  # noqa: BLE001
        print(f"rank {rank}: {e}")
    if rank == 0:
        print("NVL_SEQUENCE " + json.dumps({"world": world, "multicast": comm.has_multicast, "launches": log}))
    dist.barrier()
    comm.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
