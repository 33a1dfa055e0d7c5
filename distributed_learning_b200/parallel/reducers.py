"""Reducers: the component between a model wrapper and an allreduce algorithm.

Parity: reference ``src/reducers.py``.  A reducer has ``put(buffer)``, ``get() -> buffer`` and
``cleanup()``; results come back in submission order, which is the only thing that matches a
result to its bucket, so every rank must submit buckets in the same order.

* ``ReduceImmediately`` (reference ``ReduceImmediatelly``, ``:6-19``; the misspelt name is
  kept as an alias): the "1-step" reducer, one rank per device; ``put`` runs the allreduce in
  the caller's thread.
* ``NodeAggregateReducer`` / ``NodeAggregateReducerCPU`` (``:21-69``): the "2-step" reducer.
  Workers of a node hand buffers to the node's main process through ``mp.Queue`` pairs; the
  main process sums one buffer from each local worker, divides by the worker count, runs the
  inter-node allreduce and hands the result back to every worker.
* ``StreamReducer`` (new): device-side reducer for the ``nccl`` baseline.  ``put`` enqueues an
  asynchronous library all_reduce; ``get`` returns the buffer after making the current stream
  wait for it.  No host thread blocks.

On one NVSwitch box the native backend needs neither shape: every GPU reaches every other in
one hop, so "sum inside the node, then ring between nodes" collapses into one direct
all-to-all kernel (``distributed_learning_b200.parallel.native.NativeReducer``).  The
experiment names of both families are kept and map onto it ("1-step / 2-step").
"""
from __future__ import annotations

import queue
from typing import Callable, List, Optional

import torch
import torch.multiprocessing as mp


class ReduceImmediately:
    def __init__(self, reduce_fn: Callable[[torch.Tensor], None]):
        self.reduce_fn = reduce_fn
        self.queue: "queue.Queue[torch.Tensor]" = queue.Queue(maxsize=128)

    def put(self, buff: torch.Tensor) -> None:
        self.reduce_fn(buff)
        self.queue.put(buff)

    def get(self) -> torch.Tensor:
        return self.queue.get()

    def cleanup(self) -> None:
        pass


ReduceImmediatelly = ReduceImmediately  # the reference's spelling


class StreamReducer:
    """Asynchronous library allreduce on device buffers (baseline for ``--backend nccl``)."""

    def __init__(self, group=None):
        import torch.distributed as dist

        self._dist = dist
        self.group = group
        self.pending: "queue.Queue" = queue.Queue()

    def put(self, buff: torch.Tensor) -> None:
        work = self._dist.all_reduce(buff, op=self._dist.ReduceOp.SUM, group=self.group, async_op=True)
        self.pending.put((buff, work))

    def get(self) -> torch.Tensor:
        buff, work = self.pending.get()
        work.wait()                      # for NCCL: current stream waits, host does not block
        buff /= float(self._dist.get_world_size(self.group))
        return buff

    def cleanup(self) -> None:
        pass


class NodeAggregateReducer:
    """Worker-side handle of the 2-step reducer (lives in a spawned worker process)."""

    def __init__(self, to_node_queue, from_node_queue):
        self.to_node_queue = to_node_queue
        self.from_node_queue = from_node_queue

    def put(self, buff: torch.Tensor) -> None:
        self.to_node_queue.put(buff)

    def get(self) -> torch.Tensor:
        return self.from_node_queue.get()

    def cleanup(self) -> None:
        self.to_node_queue.put(None)


class NodeAggregateReducerCPU:
    """Node-side half of the 2-step reducer; ``pump`` runs in the process that owns the
    inter-node process group."""

    def __init__(self, reduce_fn: Callable, ndevs: int, groups=None, ctx=None):
        ctx = ctx or mp
        self.reduce_fn = reduce_fn
        self.ndevs = ndevs
        self.groups = groups
        self.to_node_queues = [ctx.Queue() for _ in range(ndevs)]
        self.from_node_queues = [ctx.Queue() for _ in range(ndevs)]
        self.reducers: List[NodeAggregateReducer] = [
            NodeAggregateReducer(t, f) for t, f in zip(self.to_node_queues, self.from_node_queues)]

    def _drain_after_shutdown(self, first_idx: int, timeout: Optional[float]) -> None:
        # every worker sends exactly one sentinel; swallow the others so no feeder thread
        # is left blocked at exit
        for i, q in enumerate(self.to_node_queues):
            if i == first_idx:
                continue
            try:
                while q.get(timeout=timeout) is not None:
                    pass
            except queue.Empty:
                pass

    def pump(self, poll_timeout: Optional[float] = None) -> int:
        """Serve rounds until a worker sends the shutdown sentinel; returns rounds served."""
        rounds = 0
        while True:
            agg = None
            n = 0
            for i, q in enumerate(self.to_node_queues):
                buff = q.get()
                if buff is None:
                    self._drain_after_shutdown(i, poll_timeout or 5.0)
                    return rounds
                n = buff.numel()
                if agg is None:
                    agg = buff.detach().clone()
                else:
                    agg[:n] += buff
                del buff
            agg /= float(self.ndevs)
            if self.groups is None:
                self.reduce_fn(agg[:n])
            else:
                self.reduce_fn(agg[:n], self.groups)
            for q in self.from_node_queues:
                q.put(agg[:n])
            del agg
            rounds += 1


NodeAgreggateReducerCPU = NodeAggregateReducerCPU  # the reference's spelling
