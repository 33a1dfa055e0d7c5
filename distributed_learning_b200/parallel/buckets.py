"""Tensor-fusion buckets.

Parity: reference ``src/ourdist.py:13-68``.

``fusion_grouping`` is the reference's ``_fusion_grouping_gen`` rule (``:54-68``): parameters
in *reverse* registration order (the order in which backward produces gradients), greedily
packed while the running byte count stays within ``grouping_size``; every bucket has at least
one tensor, so a tensor larger than the budget gets a bucket of its own and
``grouping_size=0`` gives one tensor per bucket.  The golden bucket counts of SURVEY.md §2.7
are asserted in ``tests/test_buckets.py``.

``Bucket`` replaces the reference's ``Group`` (``:13-51``): a flat ``grad_buffer`` plus the
offset of every tensor in it.  ``fuse``/``unfuse`` are the CPU/plumbing implementations; on
the native path the same layout is consumed by the ``dlb_pack`` / ``dlb_unpack`` kernels
(``csrc/pack.cu``) and the buffer lives in the peer-mapped arena.  Differences from the
reference: no spin-wait on ``grad is None`` (hooks are post-accumulate hooks, SURVEY.md
Q1/Q2), tensors that never receive a gradient are *excluded* rather than filled with
uninitialised memory (Q3), and offsets are padded so every tensor starts on a 16-byte
boundary, which the vectorised kernels rely on.
"""
from __future__ import annotations

from typing import Iterable, Iterator, List, Optional, Sequence

import torch

ALIGN_BYTES = 16


def fusion_grouping(params: Iterable[torch.Tensor], grouping_size: int = 0) -> Iterator[List[torch.Tensor]]:
    """Yield lists of tensors, one list per bucket (reference ``_fusion_grouping_gen``)."""
    to_fuse: List[torch.Tensor] = []
    running = 0
    for p in reversed(list(params)):
        size = p.element_size() * p.numel()
        if not to_fuse or running + size <= grouping_size:
            to_fuse.append(p)
            running += size
        else:
            yield to_fuse
            to_fuse = [p]
            running = size
    if to_fuse:
        yield to_fuse


def bucket_byte_sizes(params: Iterable[torch.Tensor], grouping_size: int) -> List[int]:
    return [sum(t.numel() * t.element_size() for t in g) for g in fusion_grouping(params, grouping_size)]


def _round_up(x: int, m: int) -> int:
    return (x + m - 1) // m * m


class Bucket:
    """A fusion bucket: tensors, their element offsets in a flat buffer, and the buffer.

    ``buffer`` may be passed in (a slice of the peer-mapped arena on the native path);
    otherwise it is allocated on ``device``.  ``aligned=False`` reproduces the reference's
    dense layout (needed for bit-equality of the gloo ring, whose chunking depends on the
    element count).
    """

    def __init__(self, bucket_id: int, tensors: Sequence[torch.Tensor], device="cpu",
                 dtype: Optional[torch.dtype] = None, buffer: Optional[torch.Tensor] = None, aligned: bool = False,
                 allocate: bool = True):
        self.bucket_id = bucket_id
        self.tensors = list(tensors)
        self.dtype = dtype or self.tensors[0].dtype
        esize = torch.empty((), dtype=self.dtype).element_size()
        step = max(1, ALIGN_BYTES // esize) if aligned else 1
        self.offsets: List[int] = []
        cur = 0
        for t in self.tensors:
            self.offsets.append(cur)
            cur = _round_up(cur + t.numel(), step)
        self.numel = cur
        self.dense_numel = sum(t.numel() for t in self.tensors)
        if buffer is None and allocate:
            buffer = torch.zeros(self.numel, dtype=self.dtype, device=device)
        elif buffer is not None and buffer.numel() < self.numel:
            raise ValueError("bucket buffer too small")
        # allocate=False: layout only; the owner assigns grad_buffer later (native arena)
        self.grad_buffer = buffer[: self.numel] if buffer is not None else None
        self.ready = 0

    @property
    def nbytes(self) -> int:
        return self.numel * self.grad_buffer.element_size()

    # -- plumbing implementations (reference Group.fuse / Group.unfuse) -------------------
    def fuse(self, scale: Optional[float] = None) -> torch.Tensor:
        buf = self.grad_buffer
        for t, off in zip(self.tensors, self.offsets):
            n = t.numel()
            if t.grad is not None:
                buf[off:off + n].copy_(t.grad.reshape(-1))
            else:
                buf[off:off + n].zero_()
        if scale is not None:
            buf.mul_(scale)
        return buf

    def unfuse(self, tensor: Optional[torch.Tensor] = None) -> None:
        src = self.grad_buffer if tensor is None else tensor
        for t, off in zip(self.tensors, self.offsets):
            n = t.numel()
            if t.grad is None:
                t.grad = torch.empty_like(t)
            t.grad.view(-1).copy_(src[off:off + n])

    def grad_views(self, buffer: Optional[torch.Tensor] = None) -> List[torch.Tensor]:
        """Zero-copy 'unpack': views of the (reduced) buffer shaped like each tensor."""
        src = self.grad_buffer if buffer is None else buffer
        return [src[off:off + t.numel()].view(t.shape) for t, off in zip(self.tensors, self.offsets)]

    # -- readiness counting (reference Group.add_grad) -------------------------------------
    def add_grad(self) -> bool:
        self.ready += 1
        assert self.ready <= len(self.tensors), "a gradient was reported twice in one iteration"
        return self.ready == len(self.tensors)

    def reset(self) -> None:
        self.ready = 0


def make_buckets(params: Iterable[torch.Tensor], grouping_size: int, device="cpu", dtype=None,
                 aligned: bool = False) -> List[Bucket]:
    return [Bucket(i, ts, device=device, dtype=dtype, aligned=aligned)
            for i, ts in enumerate(fusion_grouping(params, grouping_size))]


def find_unused_parameters(output: torch.Tensor, params: Iterable[torch.Tensor]) -> List[torch.Tensor]:
    """Parameters not reachable from ``output`` in the autograd graph.

    Parity: reference ``OurDist._find_unused`` (``src/ourdist.py:137-156``) walks
    ``grad_fn.next_functions`` every iteration.  Same walk here, but the callers do it once
    (the graph of these models is static) and cache the result.
    """
    params = list(params)
    by_id = {id(p): p for p in params}
    seen_params = set()
    root = output.grad_fn
    if root is None:
        return params
    stack, seen = [root], {root}
    while stack:
        fn = stack.pop()
        var = getattr(fn, "variable", None)
        if var is not None and id(var) in by_id:
            seen_params.add(id(var))
        for nxt, _ in fn.next_functions:
            if nxt is not None and nxt not in seen:
                seen.add(nxt)
                stack.append(nxt)
    return [p for p in params if id(p) not in seen_params and p.requires_grad]
