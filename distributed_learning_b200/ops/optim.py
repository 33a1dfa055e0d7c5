"""``MultiTensorSGD``: a drop-in ``torch.optim.Optimizer`` on the multi-tensor K-sgd kernel.

``FusedSGD`` (``parallel/native.py``) needs the wrapper's flat buffers.  This class is for
everything else -- a bare model, the DDP baseline, a wrapper of the host-driven backends -- and
updates any list of fp32 CUDA parameters in place with ``dlb_sgd_multi``: one launch per 64
tensors instead of torch's chain of foreach kernels.  Same semantics and state layout as
``torch.optim.SGD`` (reference ``src/main.py:36``), so ``state_dict`` round-trips with it.
"""
from __future__ import annotations

from typing import Iterable

import torch

from . import fused as K


class MultiTensorSGD(torch.optim.Optimizer):
    def __init__(self, params: Iterable, lr: float = 0.01, momentum: float = 0.0, dampening: float = 0.0,
                 weight_decay: float = 0.0, nesterov: bool = False):
        if nesterov and (momentum <= 0 or dampening != 0):
            raise ValueError("Nesterov momentum requires a momentum and zero dampening")
        super().__init__(params, dict(lr=lr, momentum=momentum, dampening=dampening, weight_decay=weight_decay,
                                      nesterov=nesterov, maximize=False, foreach=None, differentiable=False, fused=None))

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        for group in self.param_groups:
            ps, ms, gs = [], [], []
            first_with_dampening = []
            for p in group["params"]:
                if p.grad is None:
                    continue
                if not (p.is_cuda and p.dtype == torch.float32 and p.is_contiguous() and p.grad.is_contiguous()):
                    raise TypeError("MultiTensorSGD handles contiguous fp32 CUDA parameters")
                st = self.state[p]
                if "momentum_buffer" not in st or st["momentum_buffer"] is None:
                    st["momentum_buffer"] = torch.zeros_like(p)
                    if group["dampening"] != 0:
                        first_with_dampening.append(p)
                ps.append(p)
                ms.append(st["momentum_buffer"])
                gs.append(p.grad)
            if not ps:
                continue
            if first_with_dampening:
                # torch seeds the buffer with the raw gradient on the first step (no dampening)
                ids = {id(p) for p in first_with_dampening}
                plain = K.make_hyper(group["lr"], group["momentum"], 0.0, group["weight_decay"], group["nesterov"])
                sel = [i for i, p in enumerate(ps) if id(p) in ids]
                K.sgd_multi([ps[i] for i in sel], [ms[i] for i in sel], [gs[i] for i in sel], plain)
                keep = [i for i in range(len(ps)) if i not in set(sel)]
                ps, ms, gs = [ps[i] for i in keep], [ms[i] for i in keep], [gs[i] for i in keep]
                if not ps:
                    continue
            h = K.make_hyper(group["lr"], group["momentum"], group["dampening"], group["weight_decay"], group["nesterov"])
            K.sgd_multi(ps, ms, gs, h)
        return loss
