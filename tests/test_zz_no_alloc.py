"""``dlb_allreduce_emulated_static``: the emulated allreduce for callers that cannot allocate.

Some users of libdlb200 cannot hand the library a scratch buffer because they cannot allocate
device memory at all: an extension living inside a process whose memory pool it does not own, a
health check run when the device is already full, a C program that links the library and nothing
that wraps ``cudaMalloc``.  For them the library keeps the barrier flags in a fixed array of its
own.  This test drives that entry point through the C ABI (ctypes, raw pointers) and checks it
against the existing entry point, which takes caller-allocated scratch: same inputs, same kernel,
so the outputs must be equal bit for bit -- the tolerance the other allreduce tests use.  It makes
two calls in a row, the second with a different rank count and algorithm, because the array is
shared and still holds the first call's flags when the second call starts.

The file name sorts last on purpose: with ``pytest -x`` every other test has run before this one.
"""
import ctypes
from ctypes import c_float, c_int, c_int64, c_size_t, c_void_p

import pytest
import torch

pytestmark = pytest.mark.gpu


def _bind():
    from distributed_learning_b200.ops._lib import lib

    L = lib()
    fn = L.dlb_allreduce_emulated_static
    fn.restype = c_int
    fn.argtypes = [c_int, ctypes.POINTER(c_void_p), ctypes.POINTER(c_void_p), c_int, c_int64, c_int, c_float, c_int,
                   c_int, c_void_p]
    L.dlb_emulated_static_scratch_bytes.restype = c_size_t
    L.dlb_emulated_static_scratch_bytes.argtypes = []
    return L, fn


def _ptrs(tensors):
    arr = (c_void_p * len(tensors))()
    for i, t in enumerate(tensors):
        arr[i] = t.data_ptr()
    return arr


def _run_both(fn, algo, nranks, numel, dtype, seed, nblocks, nthreads):
    """Returns (outputs of the no-alloc entry point, outputs of the existing entry point)."""
    from distributed_learning_b200.ops import comm as C
    from distributed_learning_b200.ops._lib import check, dtype_code

    dev = torch.device("cuda:0")
    g = torch.Generator(device=dev).manual_seed(seed)
    inputs = [torch.randn(numel, device=dev, generator=g).to(dtype) for _ in range(nranks)]
    code = {"twoshot": 0, "oneshot": 1}[algo]
    stream = torch.cuda.current_stream().cuda_stream

    a_bufs = [x.clone() for x in inputs]
    a_outs = [torch.zeros_like(x) for x in inputs] if algo == "oneshot" else None
    rc = fn(code, _ptrs(a_bufs), _ptrs(a_outs) if a_outs is not None else None, nranks, numel, dtype_code(dtype),
            1.0 / nranks, nblocks, nthreads, stream)
    check(rc)

    b_bufs = [x.clone() for x in inputs]
    b_outs = [torch.zeros_like(x) for x in inputs] if algo == "oneshot" else None
    C.emulated_allreduce(b_bufs, algo, scale=1.0 / nranks, outs=b_outs, nblocks=nblocks, nthreads=nthreads)
    torch.cuda.synchronize()
    return (a_outs if a_outs is not None else a_bufs), (b_outs if b_outs is not None else b_bufs), inputs


def test_static_scratch_entry_point_matches_the_allocating_one_twice_in_a_row():
    L, fn = _bind()
    assert L.dlb_emulated_static_scratch_bytes() == 35072
    torch.cuda.set_device(0)
    # first call: 4 ranks, two-shot, fp32; second call right after: 8 ranks, one-shot, bf16 --
    # a different layout of the same shared array, which the entry point must have cleared
    calls = [("twoshot", 4, 40000, torch.float32, 1, 4, 128),
             ("oneshot", 8, 24000, torch.bfloat16, 2, 3, 256)]
    for algo, nranks, numel, dtype, seed, nblocks, nthreads in calls:
        got, want, inputs = _run_both(fn, algo, nranks, numel, dtype, seed, nblocks, nthreads)
        ref = inputs[0].float().clone()
        for x in inputs[1:]:
            ref += x.float()
        ref = (ref * (1.0 / nranks)).to(dtype)
        for r in range(nranks):
            assert torch.equal(got[r], want[r]), f"{algo}: rank {r} differs from the existing entry point"
            assert torch.equal(got[r], ref), f"{algo}: rank {r} differs from the fp32 reference sum"


def test_static_scratch_entry_point_rejects_sizes_beyond_its_limits():
    from distributed_learning_b200.ops._lib import DLB_E_LIMIT

    L, fn = _bind()
    torch.cuda.set_device(0)
    bufs = [torch.zeros(64, device="cuda:0") for _ in range(9)]
    stream = torch.cuda.current_stream().cuda_stream
    assert fn(0, _ptrs(bufs), None, 9, 64, 0, 1.0, 2, 64, stream) == DLB_E_LIMIT      # 9 ranks > 8
    assert b"at most 8" in L.dlb_last_error()
    assert fn(0, _ptrs(bufs[:2]), None, 2, 64, 0, 1.0, 65, 64, stream) == DLB_E_LIMIT  # 65 CTAs > 64
    assert b"at most 64" in L.dlb_last_error()
    torch.cuda.synchronize()
    assert all(torch.count_nonzero(b) == 0 for b in bufs), "a rejected call must not launch anything"
