"""The native host-side input generator (csrc/hostgen.cpp) and its use by the prefetcher."""
import torch

from distributed_learning_b200.utils.data import NativeHostRandom, PinnedPrefetcher


def test_fill_is_deterministic_and_independent_of_thread_count():
    shape = (4, 3, 50, 37)          # odd sizes: ranges do not divide evenly between threads
    outs = []
    for nthreads in (1, 2, 5):
        src = NativeHostRandom(shape, 10, seed=11, nthreads=nthreads)
        a, ta = next(src)
        b, tb = next(src)
        outs.append((a, ta, b, tb))
    for a, ta, b, tb in outs[1:]:
        assert torch.equal(a, outs[0][0]) and torch.equal(ta, outs[0][1])
        assert torch.equal(b, outs[0][2]) and torch.equal(tb, outs[0][3])
    a, ta, b, tb = outs[0]
    assert not torch.equal(a, b), "consecutive batches must differ"
    assert 0.0 <= float(a.min()) and float(a.max()) < 1.0
    assert abs(float(a.mean()) - 0.5) < 0.02 and abs(float(a.var()) - 1 / 12) < 0.01
    assert int(ta.min()) >= 0 and int(ta.max()) < 10 and ta.dtype == torch.int64
    other, _ = next(NativeHostRandom(shape, 10, seed=12, nthreads=1))
    assert not torch.equal(other, a), "another seed gives another stream"


def test_bf16_output_and_like_constructor():
    src = NativeHostRandom.like((torch.zeros(2, 1, 28, 28), torch.tensor([0, 9])), dtype=torch.bfloat16, nthreads=2)
    x, y = next(src)
    assert x.dtype == torch.bfloat16 and x.shape == (2, 1, 28, 28)
    assert 0.0 <= float(x.float().min()) and float(x.float().max()) < 1.0
    assert src.num_classes == 10 and int(y.max()) < 10


def test_prefetcher_passes_a_fill_source_through_on_cpu():
    src = NativeHostRandom((3, 1, 8, 8), 5, seed=1, nthreads=1)
    want = [next(NativeHostRandom((3, 1, 8, 8), 5, seed=1, nthreads=1))]
    pref = PinnedPrefetcher(src, "cpu")
    x, y = next(pref)
    assert torch.equal(x, want[0][0]) and torch.equal(y, want[0][1])
    x2, _ = next(pref)
    assert not torch.equal(x, x2)
