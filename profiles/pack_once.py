#!/usr/bin/env python3
"""One launch of every pack variant on GoogLeNet's and ResNet-50's gradient lists (fp32): the program
captured under ncu (`-k regex:pack_`).  A warm-up launch of each variant comes first (skip with -s)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_learning_b200.models import GoogLeNet, ResNet50  # noqa: E402
from distributed_learning_b200.ops import fused as K  # noqa: E402
from distributed_learning_b200.parallel.buckets import Bucket  # noqa: E402

dev = torch.device("cuda:0")
with torch.device("meta"):
    nets = [[p.shape for n, p in GoogLeNet().named_parameters() if not n.startswith(("model.aux1", "model.aux2"))],
            [p.shape for p in ResNet50().parameters()]]
for shapes in nets:
    grads = [torch.randn(s, device=dev) for s in reversed(shapes)]
    b = Bucket(0, grads, aligned=True, allocate=False)
    plan = K.PackPlan.for_tensors(grads, b.offsets)
    flat = torch.zeros(b.numel, device=dev)
    for variant in ("chunks", "tiles", "bulk"):
        plan.pack(flat, grads, 1.0, variant=variant)
        plan.unpack(flat, grads, 1.0, variant=variant)
torch.cuda.synchronize()
print("pack_once ok")
