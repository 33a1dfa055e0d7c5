"""Property tests (hypothesis) of the host-side layout arithmetic the kernels rely on."""
import pytest
import torch

hypothesis = pytest.importorskip("hypothesis")
from hypothesis import given, settings, strategies as st  # noqa: E402

from distributed_learning_b200.parallel.allreduce import chunk_bounds  # noqa: E402
from distributed_learning_b200.parallel.buckets import Bucket, fusion_grouping  # noqa: E402


@settings(max_examples=200, deadline=None)
@given(st.integers(0, 10_000), st.integers(1, 33))
def test_chunk_bounds_match_torch_chunk(numel, size):
    bounds = chunk_bounds(numel, size)
    assert len(bounds) == size and bounds[0][0] == 0 and bounds[-1][1] == numel
    assert all(b0 <= e0 == b1 for (b0, e0), (b1, _) in zip(bounds, bounds[1:]))
    pieces = torch.chunk(torch.arange(numel), size) if numel else ()
    assert [e - b for b, e in bounds if e > b] == [p.numel() for p in pieces]


@settings(max_examples=150, deadline=None)
@given(st.lists(st.integers(1, 5000), min_size=1, max_size=40), st.integers(0, 40_000))
def test_fusion_grouping_invariants(sizes, budget):
    params = [torch.empty(n) for n in sizes]
    groups = list(fusion_grouping(params, budget))
    flat = [t for g in groups for t in g]
    assert [id(t) for t in flat] == [id(t) for t in reversed(params)]          # reverse order, nothing lost
    for g in groups:
        nbytes = sum(t.numel() * 4 for t in g)
        assert len(g) >= 1 and (nbytes <= budget or len(g) == 1)                # over budget only when alone
    for g, nxt in zip(groups, groups[1:]):                                       # greedy: the next tensor did not fit
        assert sum(t.numel() * 4 for t in g) + nxt[0].numel() * 4 > budget


@settings(max_examples=150, deadline=None)
@given(st.lists(st.integers(1, 3000), min_size=1, max_size=30), st.sampled_from([torch.float32, torch.bfloat16]))
def test_aligned_bucket_layout(sizes, dtype):
    tensors = [torch.empty(n) for n in sizes]
    b = Bucket(0, tensors, dtype=dtype, aligned=True, allocate=False)
    per_vec = 16 // torch.empty((), dtype=dtype).element_size()
    assert b.numel % per_vec == 0                                                # the allreduce kernels need whole 16-byte vectors
    ends = [o + t.numel() for o, t in zip(b.offsets, tensors)]
    assert all(o % per_vec == 0 for o in b.offsets)                              # every tensor starts on a 16-byte boundary
    assert all(e <= o2 for e, o2 in zip(ends, b.offsets[1:])) and ends[-1] <= b.numel   # no overlap
    assert b.numel - b.dense_numel < per_vec * len(tensors)                      # padding is bounded
