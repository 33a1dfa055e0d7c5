"""The pack prologue fused into the allreduce kernels (one launch per bucket), on one GPU.

N ranks are emulated in one cooperative launch; every tensor is stored as ``[nranks, numel]`` so
that row r is emulated rank r's gradient.  The fused launch must give bit for bit what the two
launches it replaces give (``dlb_pack`` into every rank's buffer, then the allreduce), and both
must equal a plain fp32 PyTorch sum in rank order.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu

SIZES = [1, 3, 7, 8, 64, 1000, 8191, 8192, 8193, 40003, 150000]


def _layout(sizes, esize):
    step = 16 // esize
    offs, cur = [], 0
    for n in sizes:
        offs.append(cur)
        cur = (cur + n + step - 1) // step * step
    return offs, cur


def _reference(stacked, offs, total, nranks, flat_dtype, pack_scale, scale):
    bufs = []
    for r in range(nranks):
        flat = torch.zeros(total, device=stacked[0].device if stacked[0] is not None else "cuda:0", dtype=flat_dtype)
        for t, o in zip(stacked, offs):
            if t is not None:
                flat[o:o + t.shape[1]] = (t[r].float() * pack_scale).to(flat_dtype)
        bufs.append(flat)
    acc = bufs[0].float().clone()
    for b in bufs[1:]:
        acc += b.float()
    return (acc * scale).to(flat_dtype), bufs


@pytest.mark.parametrize("algo", ["twoshot", "oneshot", "twoshot_bulk"])
@pytest.mark.parametrize("nranks", [2, 3, 4, 8])
@pytest.mark.parametrize("flat_dtype,grad_dtype", [(torch.float32, torch.float32), (torch.bfloat16, torch.bfloat16),
                                                   (torch.bfloat16, torch.float32), (torch.float32, torch.bfloat16)])
def test_fused_pack_equals_pack_then_allreduce(algo, nranks, flat_dtype, grad_dtype):
    from distributed_learning_b200.ops import comm as C
    from distributed_learning_b200.ops import fused as K

    dev = torch.device("cuda:0")
    torch.manual_seed(nranks)
    stacked = [torch.randn(nranks, n, device=dev).to(grad_dtype) for n in SIZES]
    stacked[4] = None                                     # a parameter without gradient: packs zeros
    sizes = SIZES
    esize = torch.empty((), dtype=flat_dtype).element_size()
    offs, total = _layout(sizes, esize)
    pack_scale, scale = 1.0 / nranks, 1.0
    want, packed = _reference(stacked, offs, total, nranks, flat_dtype, pack_scale, scale)

    plan = K.PackPlan(sizes, offs, [grad_dtype] * len(sizes)).bind(stacked)
    nan = float("nan")
    for nblocks, nthreads in ((3, 128), (8, 512)):
        if algo == "twoshot_bulk" and nthreads != 512:
            continue                       # one 16-byte vector of the 8 KB tile per thread
        bufs = [torch.full((total,), nan, device=dev, dtype=flat_dtype) for _ in range(nranks)]
        outs = [torch.full((total,), nan, device=dev, dtype=flat_dtype) for _ in range(nranks)] if algo == "oneshot" else None
        C.emulated_allreduce(bufs, algo, scale=scale, outs=outs, nblocks=nblocks, nthreads=nthreads, pack=plan,
                             pack_scale=pack_scale)
        # the two launches it replaces
        bufs2 = [p.clone() for p in packed]
        outs2 = [torch.empty_like(b) for b in bufs2] if algo == "oneshot" else None
        # (for the bulk-copy variant the reference is the plain two-shot kernel: same sums, other engine)
        C.emulated_allreduce(bufs2, "twoshot" if algo == "twoshot_bulk" else algo, scale=scale, outs=outs2, nblocks=nblocks,
                             nthreads=nthreads)
        torch.cuda.synchronize()
        got, got2 = (outs, outs2) if algo == "oneshot" else (bufs, bufs2)
        for r in range(nranks):
            assert torch.equal(got[r], got2[r]), f"rank {r}: fused pack differs from pack + allreduce"
            assert torch.equal(got[r], want), f"rank {r}: differs from the fp32 reference"
        if algo == "oneshot":           # the inputs were packed in place and not reduced over
            for r in range(nranks):
                assert torch.equal(bufs[r], packed[r])


def test_fused_pack_with_fused_sgd_is_one_launch_for_the_whole_gradient_path():
    from distributed_learning_b200.ops import comm as C
    from distributed_learning_b200.ops import fused as K

    dev = torch.device("cuda:0")
    torch.manual_seed(0)
    nranks, sizes = 4, [5, 4096, 70001]
    stacked = [torch.randn(nranks, n, device=dev) for n in sizes]
    offs, total = _layout(sizes, 4)
    want, _ = _reference(stacked, offs, total, nranks, torch.float32, 0.25, 1.0)
    p0 = torch.randn(total, device=dev)
    ref_p = torch.nn.Parameter(p0.clone())
    opt = torch.optim.SGD([ref_p], lr=0.01, momentum=0.5)
    ref_p.grad = want.clone()
    opt.step()
    hyper = K.make_hyper(0.01, 0.5)
    for algo in ("twoshot", "oneshot", "twoshot_bulk"):
        params = [p0.clone() for _ in range(nranks)]
        moms = [torch.zeros(total, device=dev) for _ in range(nranks)]
        bufs = [torch.zeros(total, device=dev) for _ in range(nranks)]
        plan = K.PackPlan(sizes, offs, [torch.float32] * 3).bind(stacked)
        C.emulated_allreduce(bufs, algo, scale=1.0, nblocks=4, nthreads=512 if algo == "twoshot_bulk" else 256, pack=plan, pack_scale=0.25,
                             sgds=[K.fused_sgd_args(p, m, hyper) for p, m in zip(params, moms)])
        torch.cuda.synchronize()
        for r in range(nranks):
            assert torch.equal(params[r], ref_p.detach()), f"{algo}: rank {r}"


def test_fused_pack_rejects_a_layout_it_cannot_vectorise():
    from distributed_learning_b200.ops import comm as C
    from distributed_learning_b200.ops import fused as K
    from distributed_learning_b200.ops._lib import DlbError

    dev = torch.device("cuda:0")
    stacked = [torch.randn(2, 10, device=dev), torch.randn(2, 10, device=dev)]
    plan = K.PackPlan([10, 10], [0, 10], [torch.float32] * 2).bind(stacked)      # second tensor starts mid-vector
    bufs = [torch.zeros(32, device=dev) for _ in range(2)]
    with pytest.raises(DlbError, match="16-byte"):
        C.emulated_allreduce(bufs, "twoshot", pack=plan)


def test_phase_trace_is_monotonic():
    from distributed_learning_b200.ops import comm as C

    dev = torch.device("cuda:0")
    bufs = [torch.randn(1 << 16, device=dev) for _ in range(4)]
    trace = torch.zeros(4 * 8, dtype=torch.int64, device=dev)
    C.emulated_allreduce(bufs, "twoshot", scale=0.25, nblocks=4, nthreads=256, trace=trace)
    torch.cuda.synchronize()
    t = trace.view(4, 8)[:, :6].cpu()
    assert bool((t[:, 1:] >= t[:, :-1]).all()) and bool((t[:, 5] > t[:, 0]).all())


@pytest.mark.parametrize("nranks", [2, 4, 8])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_bulk_copy_twoshot_equals_twoshot_over_many_tiles(nranks, dtype):
    """Sizes that give a CTA zero, one and many tiles, a short last tile, and repeated launches on one
    flag buffer (mbarrier phases and epochs carry over nothing between launches)."""
    from distributed_learning_b200.ops import comm as C

    dev = torch.device("cuda:0")
    per = 4 if dtype == torch.float32 else 8
    flags = C.emulated_flags(nranks, dev)
    for numel in (per, 512 * per, 5000 * per, (1 << 20) + 24 * per):
        for nblocks in (1, 5):
            g = torch.Generator(device=dev).manual_seed(numel + nblocks)
            src = [torch.randn(numel, device=dev, generator=g).to(dtype) for _ in range(nranks)]
            a = [x.clone() for x in src]
            b = [x.clone() for x in src]
            C.emulated_allreduce(a, "twoshot_bulk", scale=1.0 / nranks, flags=flags, nblocks=nblocks, nthreads=512)
            C.emulated_allreduce(b, "twoshot", scale=1.0 / nranks, nblocks=nblocks, nthreads=512)
            torch.cuda.synchronize()
            for r in range(nranks):
                assert torch.equal(a[r], b[r]), (numel, nblocks, r)
