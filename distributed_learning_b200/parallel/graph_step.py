"""Whole-step CUDA graph: zero_grad + forward + loss + backward + pack + allreduce + SGD as one
replayable graph.

The reference's step is ~600 Python-launched kernels plus host-side threads and queues; for the
small CNN the host is the bottleneck.  On the native backend everything a step
does is stream work -- the hooks fire at *capture* time and enqueue the pack and the allreduce on
the side stream, which joins the capture through its events, and the cross-GPU barrier keeps its
epoch counters in device memory -- so the entire step can be captured once and replayed with one
launch.  Replays run no Python, no hooks and no allocator.

The constructor runs ``warmup`` eager steps and the capture on the example batch and then restores
parameters, momentum, module buffers, step count and RNG state, so constructing the object changes
nothing: N calls make exactly N updates, as in eager mode (``tests/test_graph_parity_gpu.py``).

Constraints (checked or documented): static shapes; the learning rate and momentum are baked
into the captured kernels (re-capture after changing them); the wrapper must be built with
``cuda_graph=True``; two graphs are captured, one per arena-region parity, and replayed
alternately (so the one-shot variant keeps the region alternation that spares it a trailing
barrier); all ranks must replay the same number of times.
"""
from __future__ import annotations

from typing import Callable, Optional

import torch
import torch.nn.functional as F


class GraphedTrainStep:
    def __init__(self, wrapper, optimizer, example_data: torch.Tensor, example_target: torch.Tensor,
                 loss_fn: Callable = F.nll_loss, warmup: int = 3):
        if not getattr(wrapper, "static_regions", True):
            raise ValueError("build the wrapper with cuda_graph=True before capturing its step")
        self.wrapper, self.optimizer, self.loss_fn = wrapper, optimizer, loss_fn
        self.data = example_data.clone()
        self.target = example_target.clone()
        self.graph: Optional[torch.cuda.CUDAGraph] = None
        self.loss: Optional[torch.Tensor] = None
        self.launches_per_step = 0
        dev = self.data.device
        # The warm-up steps are real steps (they build the buckets, tune, let cuDNN pick algorithms)
        # and must leave no trace: parameters, momentum, BatchNorm statistics, the optimizer's step
        # count and the RNG are put back afterwards, so the first replay is the first update.
        snap = None
        if hasattr(wrapper, "snapshot_state"):
            wrapper.build_now(self.data)
            torch.cuda.synchronize(dev)
            snap = wrapper.snapshot_state()
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            for _ in range(max(1, warmup)):
                self._eager()
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        before = getattr(wrapper, "kernel_launches", 0)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.loss = self._eager()
        self.launches_per_step = getattr(wrapper, "kernel_launches", 0) - before
        # Second graph with the buckets' other arena regions: replayed on odd steps, so consecutive
        # steps alternate regions exactly as eager mode does and the one-shot variant needs no
        # trailing barrier.  Both graphs share one memory pool (they never run concurrently).
        self.graphs = [self.graph]
        self.losses = [self.loss]
        if getattr(wrapper, "alternate_graphs", False):
            wrapper.__dict__["graph_parity"] = 1
            g1 = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g1, pool=self.graph.pool()):
                l1 = self._eager()
            wrapper.__dict__["graph_parity"] = 0
            wrapper.__dict__["kernel_launches"] = before + self.launches_per_step
            self.graphs.append(g1)
            self.losses.append(l1)
        self.calls = 0
        if snap is not None:
            torch.cuda.synchronize(dev)
            wrapper.restore_state(snap)
            torch.cuda.synchronize(dev)

    def _eager(self) -> torch.Tensor:
        self.optimizer.zero_grad()
        out = self.wrapper(self.data)
        loss = self.loss_fn(out.float() if out.dtype != torch.float32 else out, self.target)
        loss.backward()
        self.wrapper.sync_gradients()
        self.optimizer.step()
        return loss.detach()

    def __call__(self, data: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
        """Run one training step on ``(data, target)``; returns the (static) loss tensor."""
        if data.data_ptr() != self.data.data_ptr():
            self.data.copy_(data, non_blocking=True)
        if target.data_ptr() != self.target.data_ptr():
            self.target.copy_(target, non_blocking=True)
        k = self.calls % len(self.graphs)
        self.calls += 1
        self.graphs[k].replay()
        self.loss = self.losses[k]
        if hasattr(self.optimizer, "steps"):
            self.optimizer.steps += 1           # the replay ran the update; no Python optimizer code did
        if hasattr(self.wrapper, "kernel_launches"):
            self.wrapper.__dict__["kernel_launches"] = self.wrapper.kernel_launches + self.launches_per_step
        return self.loss
