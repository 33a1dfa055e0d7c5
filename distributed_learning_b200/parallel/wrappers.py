"""Distributed model wrappers for the host-driven backends (gloo plumbing, nccl baseline).

Parity: reference ``src/ourdist.py`` and the two inline wrappers of ``src/main.py:181-248``.
Every wrapper has the reference's surface: ``wrapper(data)`` / ``forward(data)``,
``sync_gradients()``, ``cleanup()``, and forwards unknown attributes to the model
(``src/ourdist.py:174-178``).  Constructor signature ``(model, reducer, grouping_size,
grad_buff_device)`` is kept.

* ``OurDist`` (``:53-178``): pipelining + fusion ("P+F").  A sender thread waits for each
  bucket to fill (in bucket order), fuses it and ``put``s it; a receiver thread ``get``s the
  results in the same order and unfuses them; ``sync_gradients`` waits for the last bucket.
* ``SeqMergeDist`` (``:180-203``): fusion only, after backward.
* ``SeqDist`` (``:205-225``): per-parameter, after backward.
* ``WarmupDist`` (``:227-248``): cached zero output, exercises data loading and device init.
* ``SingleDist`` / ``make_ddp`` (``src/main.py:181-196,239-248``): controls.

Deliberate differences: gradients are observed with
``register_post_accumulate_grad_hook`` so the gradient exists when the hook runs (Q1/Q2: no
spin-wait, no stale read); parameters that the loss does not reach are found once at the
first forward and left out of the buckets (Q3/Q4) instead of all-reducing uninitialised
memory every step; threads are daemons and ``cleanup`` joins them; a failure in a helper
thread is re-raised in ``sync_gradients`` instead of being printed and swallowed.

The native NVLink wrapper is ``distributed_learning_b200.parallel.native.B200Dist``.
"""
from __future__ import annotations

import threading
from typing import List, Optional

import torch

from .buckets import Bucket, find_unused_parameters, fusion_grouping


class _Wrapper:
    """Common surface: call-through, attribute passthrough, parameter list."""

    def __init__(self, model, reducer):
        self.__dict__["model"] = model
        self.__dict__["reducer"] = reducer

    def forward(self, data):
        return self.model(data)

    def __call__(self, data):
        return self.forward(data)

    def __getattr__(self, attr):
        # only reached when normal lookup fails
        return getattr(self.__dict__["model"], attr)

    def sync_gradients(self) -> None:
        pass

    def cleanup(self) -> None:
        pass


class _BucketedWrapper(_Wrapper):
    def __init__(self, model, reducer, grouping_size, grad_buff_device="cpu", exclude_unused: bool = True,
                 dynamic_unused: bool = False):
        super().__init__(model, reducer)
        self.grouping_size = int(grouping_size)
        self.grad_buff_device = grad_buff_device
        # dynamic_unused: the reference's per-iteration walk (src/ourdist.py:137-172): nothing is
        # excluded, and after every forward the parameters this step's loss does not reach are
        # marked ready with no gradient (fused as zeros, not as uninitialised memory: Q3)
        self.dynamic_unused = bool(dynamic_unused)
        self.exclude_unused = exclude_unused and not self.dynamic_unused
        self.groups: Optional[List[Bucket]] = None
        self.unused: List[torch.Tensor] = []

    def _dead_this_step(self, output) -> set:
        if not self.dynamic_unused:
            return set()
        bucketed = [t for b in self.groups for t in b.tensors]
        dead = find_unused_parameters(output, bucketed)
        for p in dead:
            p.grad = None
        return {id(p) for p in dead}

    def _build_groups(self, output: Optional[torch.Tensor]) -> None:
        params = [p for p in self.model.parameters() if p.requires_grad]
        if output is not None and self.exclude_unused:
            self.unused = find_unused_parameters(output, params)
            dead = {id(p) for p in self.unused}
            params = [p for p in params if id(p) not in dead]
        self.groups = [Bucket(i, ts, device=self.grad_buff_device)
                       for i, ts in enumerate(fusion_grouping(params, self.grouping_size))]
        self._after_build()

    def _after_build(self) -> None:
        pass


class OurDist(_BucketedWrapper):
    """Overlap of gradient reduction with backward, on fused buckets."""

    def __init__(self, model, reducer, grouping_size=0, grad_buff_device="cpu", exclude_unused: bool = True,
                 dynamic_unused: bool = False):
        super().__init__(model, reducer, grouping_size, grad_buff_device, exclude_unused, dynamic_unused)
        self._hooks = []
        self._bucket_events: List[threading.Event] = []
        self._start_send = threading.Event()
        self._start_recv = threading.Event()
        self._done = threading.Event()
        self._done.set()
        self._shutting_down = False
        self._error: Optional[BaseException] = None
        self._threads = [threading.Thread(target=self._guard, args=(self._sender,), daemon=True, name="dlb200-send"),
                         threading.Thread(target=self._guard, args=(self._receiver,), daemon=True, name="dlb200-recv")]
        self._threads_started = False
        self._iteration_armed = False

    # -- set-up --------------------------------------------------------------------------
    def _after_build(self) -> None:
        self._bucket_events = [threading.Event() for _ in self.groups]
        for b in self.groups:
            for t in b.tensors:
                self._hooks.append(t.register_post_accumulate_grad_hook(self._make_hook(b)))
        if not self._threads_started:
            for th in self._threads:
                th.start()
            self._threads_started = True

    def _make_hook(self, bucket: Bucket):
        def hook(_param):
            if not self._iteration_armed:
                return                       # backward outside forward()/sync_gradients(): not ours
            if bucket.add_grad():
                self._bucket_events[bucket.bucket_id].set()
        return hook

    # -- helper threads ------------------------------------------------------------------
    def _guard(self, fn):
        try:
            fn()
        except BaseException as e:  # noqa: BLE001 - reported to the training thread
            self._error = e
            self._done.set()

    def _sender(self):
        while True:
            self._start_send.wait()
            self._start_send.clear()
            if self._shutting_down:
                return
            for b, ev in zip(self.groups, self._bucket_events):
                ev.wait()
                if self._shutting_down:
                    return
                self.reducer.put(b.fuse())

    def _receiver(self):
        while True:
            self._start_recv.wait()
            self._start_recv.clear()
            if self._shutting_down:
                return
            for b in self.groups:
                b.unfuse(self.reducer.get())
            self._done.set()

    # -- step ----------------------------------------------------------------------------
    def forward(self, data):
        output = self.model(data)
        if self.groups is None:
            self._build_groups(output)
        for b, ev in zip(self.groups, self._bucket_events):
            b.reset()
            ev.clear()
        self._done.clear()
        self._iteration_armed = True
        self._start_send.set()
        self._start_recv.set()
        dead = self._dead_this_step(output)
        if dead:
            for b in self.groups:
                for t in b.tensors:
                    if id(t) in dead and b.add_grad():
                        self._bucket_events[b.bucket_id].set()
        return output

    def sync_gradients(self) -> None:
        # backward has returned, so every hook that will fire has fired: a bucket that is still
        # short will never fill (the set of used parameters changed) and the sender would wait on
        # it forever.  Fail instead of hanging.
        if self.groups is not None and self._iteration_armed:
            short = [b for b in self.groups if b.ready != len(b.tensors)]
            if short:
                self._iteration_armed = False
                self._shutting_down = True
                for ev in self._bucket_events:
                    ev.set()
                b = short[0]
                raise RuntimeError(
                    f"bucket {b.bucket_id}: {b.ready}/{len(b.tensors)} gradients arrived; the set of parameters the "
                    "loss depends on changed since the first step (use dynamic_unused=True / --dynamic_unused 1)")
        self._done.wait()
        self._iteration_armed = False
        if self._error is not None:
            err, self._error = self._error, None
            raise RuntimeError("gradient reduction failed in a helper thread") from err

    def cleanup(self) -> None:
        self._shutting_down = True
        self._start_send.set()
        self._start_recv.set()
        for ev in self._bucket_events:
            ev.set()
        for h in self._hooks:
            h.remove()
        self._hooks = []
        if self._threads_started:
            for th in self._threads:
                th.join(timeout=5.0)


class SeqMergeDist(_BucketedWrapper):
    """Fusion without overlap: fuse, reduce and unfuse bucket by bucket after backward."""

    def __init__(self, model, reducer, grouping_size, _grad_buffer_device="cpu", exclude_unused: bool = True,
                 dynamic_unused: bool = False):
        super().__init__(model, reducer, grouping_size, _grad_buffer_device, exclude_unused, dynamic_unused)

    def forward(self, data):
        output = self.model(data)
        if self.groups is None:
            self._build_groups(output)
        self._dead_this_step(output)         # clears stale gradients of parameters unused this step
        return output

    def sync_gradients(self) -> None:
        for b in self.groups or []:
            self.reducer.put(b.fuse())
            b.unfuse(self.reducer.get())


class SeqDist(_Wrapper):
    """Neither fusion nor overlap: one allreduce per parameter after backward."""

    def __init__(self, model, reducer, _grouping_size=0, _grad_buffer_device="cpu"):
        super().__init__(model, reducer)

    def sync_gradients(self) -> None:
        for p in self.model.parameters():
            if p.grad is None:
                continue
            flat = p.grad.view(-1)
            self.reducer.put(flat)
            out = self.reducer.get()
            if out.data_ptr() != flat.data_ptr():
                flat.copy_(out)


class WarmupDist(_Wrapper):
    """Runs the model once to learn the output shape, then returns a cached zero tensor."""

    def __init__(self, model, reducer=None, _grouping_size=0, _grad_buffer_device="cpu"):
        super().__init__(model, reducer)
        self.__dict__["out"] = None

    def forward(self, data):
        if self.out is None:
            with torch.no_grad():
                shape_of = self.model(data)
            self.__dict__["out"] = torch.zeros_like(shape_of, requires_grad=True)
        return self.out


class SingleDist(_Wrapper):
    """Single-device control: no synchronisation (reference ``main_single``)."""

    def __init__(self, model, reducer=None, _grouping_size=0, _grad_buffer_device="cpu"):
        super().__init__(model, reducer)


class DDPDist(_Wrapper):
    """torch ``DistributedDataParallel`` behind the wrapper surface (reference ``main_ddp``)."""

    def __init__(self, model, reducer=None, grouping_size=25 * 1024 * 1024, _grad_buffer_device="cpu",
                 device_ids=None, find_unused_parameters: bool = True, process_group=None):
        from torch.nn.parallel import DistributedDataParallel as DDP

        super().__init__(model, reducer)
        bucket_mb = max(grouping_size, 1) / 1024 / 1024
        self.__dict__["ddp"] = DDP(model, device_ids=device_ids, bucket_cap_mb=bucket_mb,
                                   find_unused_parameters=find_unused_parameters, process_group=process_group)

    def forward(self, data):
        return self.ddp(data)

    def parameters(self):
        return self.model.parameters()
