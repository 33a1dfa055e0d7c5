#!/usr/bin/env python3
"""Every kernel family of libdlb200 once, at small sizes: the program run under
`compute-sanitizer --tool memcheck|racecheck|synccheck` (one tool per GPU call)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_learning_b200.ops import comm as C  # noqa: E402
from distributed_learning_b200.ops import fused as K  # noqa: E402

dev = torch.device("cuda:0")
torch.manual_seed(0)
sizes = [5, 1000, 8193, 40003]
offs, cur = [], 0
for n in sizes:
    offs.append(cur)
    cur = (cur + n + 3) // 4 * 4
ts = [torch.randn(n, device=dev) for n in sizes]
flat = torch.zeros(cur, device=dev)
for variant in ("chunks", "tiles", "bulk"):
    K.pack(flat, ts, offs, scale=0.5, variant=variant)
    back = [torch.empty_like(t) for t in ts]
    K.unpack(flat, back, offs, scale=2.0, variant=variant)
    torch.cuda.synchronize()
    assert all(torch.equal(a, b) for a, b in zip(ts, back)), variant
h = K.make_hyper(0.01, 0.5)
p, m, g = torch.ones(cur, device=dev), torch.zeros(cur, device=dev), torch.ones(cur, device=dev)
K.sgd_flat(p, m, g, h)
K.sgd_multi([torch.ones(100, device=dev)], [torch.zeros(100, device=dev)], [torch.ones(100, device=dev)], h)
nr = 2
stacked = [torch.randn(nr, n, device=dev) for n in sizes]
plan = K.PackPlan(sizes, offs, [torch.float32] * len(sizes)).bind(stacked)
for algo in ("twoshot", "oneshot", "twoshot_bulk"):
    bufs = [torch.zeros(cur, device=dev) for _ in range(nr)]
    outs = [torch.zeros(cur, device=dev) for _ in range(nr)] if algo == "oneshot" else None
    params = [torch.ones(cur, device=dev) for _ in range(nr)]
    moms = [torch.zeros(cur, device=dev) for _ in range(nr)]
    C.emulated_allreduce(bufs, algo, scale=1.0, outs=outs, nblocks=2, nthreads=512, pack=plan, pack_scale=0.5,
                         sgds=[K.fused_sgd_args(a, b, h) for a, b in zip(params, moms)])
    torch.cuda.synchronize()
    assert torch.equal(params[0], params[1]), algo
print("sanitize_once ok")
