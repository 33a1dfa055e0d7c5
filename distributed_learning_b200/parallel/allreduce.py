"""Host-driven allreduce algorithms over ``torch.distributed`` point-to-point messages.

These are the *plumbing* and *baseline* algorithms (``--backend gloo`` and ``--backend nccl``);
the product path is the in-kernel NVLink allreduce in ``csrc/allreduce.cu`` bound by
``distributed_learning_b200.ops.comm.NativeComm``.

Parity: reference ``src/allreduce.py``:

* ``built_in_allreduce`` (``:5-7``): library all_reduce(SUM) then divide.
* ``central_allreduce`` (``:9-43``): rank 0 gathers, sums in rank order, divides, sends back.
* ``ring_allreduce`` (``:45-98``): ``size-1`` reduce-scatter steps then ``size-1`` all-gather
  steps over ``torch.chunk`` pieces, sending to ``rank+1`` and receiving from ``rank-1``, then
  one division.  The chunk boundaries, ring direction and order of additions are the same, so
  on CPU the result is bit-identical to the reference's (checked against a plain restatement
  of that arithmetic in ``tests/test_distributed_cpu.py``, and against the reference itself in
  ``tests/test_reference_parity_cpu.py``).
* ``ring_allreduce_gpu`` (``:100-170``): the reference emulates NCCL point-to-point with
  broadcasts in two-rank groups, and special-cases the last rank to avoid deadlock.  NCCL has
  had real send/recv for years, so here each ring step is one grouped
  ``batch_isend_irecv`` (send right + receive left); ``groups`` is accepted and ignored.

All functions average ``send`` in place and accept an optional process ``group``.
"""
from __future__ import annotations

from typing import List, Tuple

import torch
import torch.distributed as dist


def _world(group=None) -> Tuple[int, int]:
    return dist.get_rank(group), dist.get_world_size(group)


def _global_rank(group, group_rank: int) -> int:
    if group is None or group is dist.group.WORLD:
        return group_rank
    return dist.get_global_rank(group, group_rank)


def chunk_bounds(numel: int, size: int) -> List[Tuple[int, int]]:
    """``[begin, end)`` of each of ``size`` ring chunks; same split as ``torch.chunk``
    (chunk length ``ceil(numel/size)``, trailing chunks may be short or empty)."""
    step = -(-numel // size) if numel else 0
    return [(min(i * step, numel), min((i + 1) * step, numel)) for i in range(size)]


def built_in_allreduce(send: torch.Tensor, group=None) -> None:
    dist.all_reduce(send, op=dist.ReduceOp.SUM, group=group)
    send /= float(dist.get_world_size(group))


def central_allreduce(send: torch.Tensor, group=None) -> None:
    """Parameter-server style: rank 0 is the centre."""
    rank, size = _world(group)
    if size <= 1:
        return
    if rank == 0:
        inbox = [torch.empty_like(send) for _ in range(size - 1)]
        pending = [dist.irecv(buf, _global_rank(group, src), group=group) for src, buf in enumerate(inbox, start=1)]
        for req, buf in zip(pending, inbox):
            req.wait()
            send += buf
        send /= float(size)
        for req in [dist.isend(send, _global_rank(group, dst), group=group) for dst in range(1, size)]:
            req.wait()
    else:
        root = _global_rank(group, 0)
        dist.send(send, root, group=group)
        dist.recv(send, root, group=group)


def _ring_pass(flat: torch.Tensor, bounds, rank: int, size: int, first_chunk: int, accumulate: bool,
               scratch: torch.Tensor, right: int, left: int, group, grouped: bool) -> None:
    """One pass of ``size-1`` steps.  At step ``i`` chunk ``first_chunk - i`` goes right and
    chunk ``first_chunk - i - 1`` arrives from the left."""
    in_flight = []
    for i in range(size - 1):
        sb, se = bounds[(first_chunk - i) % size]
        rb, re = bounds[(first_chunk - i - 1) % size]
        out_piece, in_piece = flat[sb:se], scratch[: re - rb]
        if grouped:
            ops = []
            if se > sb:
                ops.append(dist.P2POp(dist.isend, out_piece, right, group))
            if re > rb:
                ops.append(dist.P2POp(dist.irecv, in_piece, left, group))
            for req in (dist.batch_isend_irecv(ops) if ops else []):
                req.wait()
        else:
            if se > sb:
                in_flight.append(dist.isend(out_piece, right, group=group))
            if re > rb:
                dist.recv(in_piece, left, group=group)      # blocking: the data is needed now
        if re > rb:
            if accumulate:
                flat[rb:re] += in_piece
            else:
                flat[rb:re] = in_piece
    for req in in_flight:
        req.wait()


def _ring(send: torch.Tensor, group, grouped: bool) -> None:
    rank, size = _world(group)
    if size <= 1:
        return
    flat = send.view(-1)
    bounds = chunk_bounds(flat.numel(), size)
    scratch = torch.empty(max(e - b for b, e in bounds) or 1, dtype=flat.dtype, device=flat.device)
    right = _global_rank(group, (rank + 1) % size)
    left = _global_rank(group, (rank - 1) % size)
    # reduce-scatter: afterwards rank r owns the full sum of chunk r+1
    _ring_pass(flat, bounds, rank, size, rank, True, scratch, right, left, group, grouped)
    # all-gather: circulate the finished chunks
    _ring_pass(flat, bounds, rank, size, rank + 1, False, scratch, right, left, group, grouped)
    send /= float(size)


def ring_allreduce(send: torch.Tensor, group=None) -> None:
    """Ring allreduce over isend/recv (gloo; CPU tensors)."""
    _ring(send, group, grouped=False)


def ring_allreduce_gpu(send: torch.Tensor, groups=None, group=None) -> None:
    """Ring allreduce over grouped NCCL send/recv (device tensors).  ``groups`` is ignored."""
    _ring(send, group, grouped=True)


ALGORITHMS = {
    "ring": ring_allreduce,
    "ring_gpu": ring_allreduce_gpu,
    "central": central_allreduce,
    "built_in": built_in_allreduce,
}
