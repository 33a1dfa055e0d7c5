"""The cross-GPU allreduce kernels on ONE GPU: N emulated ranks in one cooperative launch
(``dlb_allreduce_emulated``), compared with ``torch.sum`` in fp32; and the single-rank paths of
``NativeComm`` / ``B200Dist``.  Real multi-GPU runs are in ``test_multigpu.py``."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _dev():
    return torch.device("cuda:0")


def _inputs(nranks, numel, dtype, seed=0):
    g = torch.Generator(device=_dev()).manual_seed(seed)
    return [torch.randn(numel, device=_dev(), generator=g).to(dtype) for _ in range(nranks)]


def _expected(bufs, scale, dtype):
    acc = bufs[0].float().clone()
    for b in bufs[1:]:
        acc += b.float()           # same rank order as the kernel: bit-exact in fp32
    return (acc * scale).to(dtype)


@pytest.mark.parametrize("algo", ["twoshot", "oneshot"])
@pytest.mark.parametrize("nranks", [2, 3, 4, 8])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("numel", [8, 64, 4096, 100000, 1 << 20])
def test_emulated_allreduce(algo, nranks, dtype, numel):
    from distributed_learning_b200.ops import comm as C

    per_vec = 16 // torch.empty((), dtype=dtype).element_size()
    numel = (numel + per_vec - 1) // per_vec * per_vec
    bufs = _inputs(nranks, numel, dtype)
    want = _expected(bufs, 1.0 / nranks, dtype)
    originals = [b.clone() for b in bufs]
    outs = [torch.zeros_like(b) for b in bufs] if algo == "oneshot" else None
    C.emulated_allreduce(bufs, algo, scale=1.0 / nranks, outs=outs, nblocks=4, nthreads=128)
    torch.cuda.synchronize()
    results = outs if algo == "oneshot" else bufs
    for r in results:
        assert torch.equal(r, want)
    if algo == "oneshot":
        for b, o in zip(bufs, originals):
            assert torch.equal(b, o), "one-shot must not modify its inputs"


def test_emulated_allreduce_many_back_to_back_launches_reuse_flags():
    """Epoch counters only grow: the same flag pads serve many launches with different grids."""
    from distributed_learning_b200.ops import comm as C

    nranks = 4
    flags = C.emulated_flags(nranks, _dev())
    for it in range(40):
        numel = 4 * (1 + 37 * it)
        bufs = _inputs(nranks, numel, torch.float32, seed=it)
        want = _expected(bufs, 1.0, torch.float32)
        algo = "twoshot" if it % 2 == 0 else "oneshot"
        outs = [torch.empty_like(b) for b in bufs] if algo == "oneshot" else None
        C.emulated_allreduce(bufs, algo, scale=1.0, outs=outs, flags=flags, nblocks=1 + it % 6, nthreads=64)
        for r in (outs if outs is not None else bufs):
            assert torch.equal(r, want)


@pytest.mark.parametrize("algo", ["twoshot", "oneshot"])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_emulated_allreduce_fused_sgd(algo, dtype):
    """The fused epilogue must equal allreduce followed by torch.optim.SGD on every rank."""
    from distributed_learning_b200.ops import comm as C
    from distributed_learning_b200.ops import fused as K

    nranks, numel = 4, 50000 // 8 * 8
    h = K.make_hyper(0.01, 0.5)
    params = [torch.randn(numel, device=_dev()) for _ in range(nranks)]
    params = [params[0].clone() for _ in range(nranks)]            # replicas start equal
    moms = [torch.zeros(numel, device=_dev()) for _ in range(nranks)]
    shadows = [torch.zeros(numel, device=_dev(), dtype=torch.bfloat16) for _ in range(nranks)]
    ref_p = torch.nn.Parameter(params[0].clone())
    opt = torch.optim.SGD([ref_p], lr=0.01, momentum=0.5)
    flags = C.emulated_flags(nranks, _dev())
    for step in range(3):
        bufs = _inputs(nranks, numel, dtype, seed=100 + step)
        avg = _expected(bufs, 1.0 / nranks, dtype)
        ref_p.grad = avg.float()
        opt.step()
        sgds = [K.fused_sgd_args(params[r], moms[r], h, shadows[r]) for r in range(nranks)]
        outs = [torch.empty_like(b) for b in bufs] if algo == "oneshot" else None
        C.emulated_allreduce(bufs, algo, scale=1.0 / nranks, outs=outs, sgds=sgds, flags=flags, nblocks=3, nthreads=128)
        torch.cuda.synchronize()
        for r in range(nranks):
            assert torch.equal((outs or bufs)[r], avg)
    for r in range(nranks):
        assert torch.allclose(params[r], ref_p.detach(), rtol=1e-6, atol=1e-7)
        assert torch.equal(params[r], params[0]), "replicas must stay bit-identical"
        assert torch.equal(shadows[r], params[r].to(torch.bfloat16))


def test_native_comm_single_rank():
    from distributed_learning_b200.ops.comm import NativeComm

    comm = NativeComm(device=_dev())
    comm.allocate(1 << 20)
    assert comm.mode in ("vmm", "ipc")
    v = comm.view(0, 1024, torch.float32)
    v.copy_(torch.arange(1024, device=_dev(), dtype=torch.float32))
    comm.barrier()
    comm.allreduce(0, 1024, torch.float32, "twoshot", scale=2.0)
    comm.allreduce(0, 1024, torch.float32, "oneshot", scale=0.5, out_offset=8192)
    comm.check()
    assert torch.equal(comm.view(8192, 1024, torch.float32), torch.arange(1024, device=_dev(), dtype=torch.float32))
    comm.close()


def test_topology_query():
    from distributed_learning_b200.parallel.topology import describe_topology, query_topology

    topo = query_topology()
    assert topo["devices"] and topo["devices"][0]["sms"] > 0
    assert "gpu0" in describe_topology(topo)


@pytest.mark.parametrize("fused_sgd", [False, True])
@pytest.mark.parametrize("model_dtype", ["fp32", "bf16"])
def test_b200dist_single_gpu_matches_plain_training(fused_sgd, model_dtype):
    """B200Dist + FusedSGD on one GPU follows plain PyTorch training of the same model."""
    import torch.nn.functional as F

    from distributed_learning_b200.models import CifarNet
    from distributed_learning_b200.parallel.native import B200Dist

    dev = _dev()
    torch.manual_seed(5)
    ref = CifarNet().to(dev).train()
    mine = CifarNet().to(dev).train()
    mine.load_state_dict(ref.state_dict())
    for m in (ref, mine):                      # dropout would need identical RNG streams
        m.dropout1.p = 0.0
        m.dropout2.p = 0.0
    wrapper = B200Dist(mine, None, 1 << 20, dev, fused_sgd=fused_sgd, model_dtype=model_dtype)
    opt = wrapper.make_optimizer(lr=0.01, momentum=0.5)
    ref_opt = torch.optim.SGD(ref.parameters(), lr=0.01, momentum=0.5)
    xs = [torch.rand(16, 3, 32, 32, device=dev) for _ in range(4)]
    ys = [torch.randint(10, (16,), device=dev) for _ in range(4)]
    losses, ref_losses = [], []
    for x, y in zip(xs, ys):
        opt.zero_grad()
        loss = F.nll_loss(wrapper(x).float(), y)
        loss.backward()
        wrapper.sync_gradients()
        opt.step()
        losses.append(loss.item())
        ref_opt.zero_grad()
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=model_dtype == "bf16"):
            out = ref(x)
        rl = F.nll_loss(out.float(), y)
        rl.backward()
        ref_opt.step()
        ref_losses.append(rl.item())
    tol = 5e-2 if model_dtype == "bf16" else 1e-4
    assert losses == pytest.approx(ref_losses, rel=tol, abs=tol)
    if model_dtype == "fp32":
        for (n, p), q in zip(mine.named_parameters(), ref.parameters()):
            assert torch.allclose(p, q, rtol=1e-4, atol=1e-5), n
    assert wrapper.kernel_launches > 0
    sd = opt.state_dict()
    assert len(sd["state"]) == len(list(mine.parameters()))
    wrapper.cleanup()


@pytest.mark.parametrize("model_dtype", ["fp32", "bf16"])
def test_graphed_train_step_matches_eager(model_dtype):
    """The whole step captured in a CUDA graph gives the same parameters as eager stepping."""
    import torch.nn.functional as F

    from distributed_learning_b200.models import CifarNet
    from distributed_learning_b200.parallel.graph_step import GraphedTrainStep
    from distributed_learning_b200.parallel.native import B200Dist

    dev = _dev()
    xs = [torch.rand(16, 3, 32, 32, device=dev) for _ in range(6)]
    ys = [torch.randint(10, (16,), device=dev) for _ in range(6)]

    def run(graphed):
        torch.manual_seed(7)
        net = CifarNet().to(dev).train()
        net.dropout1.p = net.dropout2.p = 0.0
        w = B200Dist(net, None, 1 << 20, dev, model_dtype=model_dtype, cuda_graph=graphed)
        opt = w.make_optimizer(lr=0.01, momentum=0.5)
        losses = []
        if graphed:
            # the warm-up steps inside GraphedTrainStep leave no trace: N calls are N updates
            step = GraphedTrainStep(w, opt, xs[0], ys[0], warmup=2)
            for x, y in zip(xs, ys):
                losses.append(step(x, y).item())
        else:
            def eager(x, y):
                opt.zero_grad()
                loss = F.nll_loss(w(x).float(), y)
                loss.backward()
                w.sync_gradients()
                opt.step()
                return loss.item()
            for x, y in zip(xs, ys):
                losses.append(eager(x, y))
        flat = torch.cat([p.detach().float().reshape(-1) for p in net.parameters()])
        w.cleanup()
        return losses, flat

    l_g, p_g = run(True)
    l_e, p_e = run(False)
    tol = 2e-2 if model_dtype == "bf16" else 1e-5
    assert l_g == pytest.approx(l_e, rel=tol, abs=tol)
    assert torch.allclose(p_g, p_e, rtol=tol, atol=tol)


def test_b200dist_unpack_copy_mode_gives_own_gradient_tensors():
    """``unpack="copy"`` runs the dlb_unpack kernel: gradients land in fresh tensors, not arena views."""
    import torch.nn.functional as F

    from distributed_learning_b200.models import CifarNet
    from distributed_learning_b200.parallel.native import B200Dist

    dev = _dev()
    torch.manual_seed(3)
    ref = CifarNet().to(dev).train()
    mine = CifarNet().to(dev).train()
    mine.load_state_dict(ref.state_dict())
    for m in (ref, mine):
        m.dropout1.p = m.dropout2.p = 0.0
    w = B200Dist(mine, None, 64 * 1024, dev, fused_sgd=False, unpack="copy")
    opt = w.make_optimizer(lr=0.0)                       # lr 0: only look at the gradients
    x, y = torch.rand(8, 3, 32, 32, device=dev), torch.randint(10, (8,), device=dev)
    opt.zero_grad()
    F.nll_loss(w(x), y).backward()
    w.sync_gradients()
    F.nll_loss(ref(x), y).backward()
    torch.cuda.synchronize()
    lo, hi = w.comm.local.data_ptr(), w.comm.local.data_ptr() + w.comm.nbytes
    for p, q in zip(mine.parameters(), ref.parameters()):
        assert torch.allclose(p.grad, q.grad, rtol=1e-5, atol=1e-6)
        assert not (lo <= p.grad.data_ptr() < hi)
    w.cleanup()
