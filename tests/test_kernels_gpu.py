"""Numerics of the single-GPU kernels against plain PyTorch fp32 references (needs a GPU)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _dev():
    return torch.device("cuda:0")


def _layout(sizes, esize):
    step = 16 // esize
    offs, cur = [], 0
    for n in sizes:
        offs.append(cur)
        cur = (cur + n + step - 1) // step * step
    return offs, cur


@pytest.mark.parametrize("variant", ["auto", "chunks", "tiles", "bulk"])
@pytest.mark.parametrize("flat_dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("src_dtype", [torch.float32, torch.bfloat16, torch.float16])
def test_pack_unpack_matches_torch(flat_dtype, src_dtype, variant):
    from distributed_learning_b200.ops import fused as K

    if variant == "bulk" and src_dtype != flat_dtype:
        pytest.skip("the bulk-copy variant moves bytes: same dtype on both sides")
    dev = _dev()
    torch.manual_seed(0)
    sizes = [1, 3, 7, 8, 64, 1000, 8191, 8192, 8193, 100003, 300000]
    tensors = [torch.randn(n, device=dev).to(src_dtype) for n in sizes]
    offs, total = _layout(sizes, torch.empty((), dtype=flat_dtype).element_size())
    for scale in (0.125, 1.0):
        flat = torch.full((total,), 7.0, device=dev, dtype=flat_dtype)
        K.pack(flat, tensors, offs, scale=scale, variant=variant)
        torch.cuda.synchronize()
        for t, o in zip(tensors, offs):
            want = (t.float() * scale).to(flat_dtype)
            assert torch.equal(flat[o:o + t.numel()], want), (t.numel(), scale)
        # unpack into fresh tensors (cast back, scale 1/scale)
        back = [torch.zeros_like(t) for t in tensors]
        K.unpack(flat, back, offs, scale=1.0 / scale, variant=variant)
        torch.cuda.synchronize()
        for t, b, o in zip(tensors, back, offs):
            want = (flat[o:o + t.numel()].float() / scale).to(src_dtype)
            assert torch.equal(b, want), (t.numel(), scale)


@pytest.mark.parametrize("variant", ["tiles", "bulk"])
def test_pack_variants_leave_holes_alone_and_zero_the_pads(variant):
    """Layout with holes between tensors and with storage that is not 16-byte aligned."""
    from distributed_learning_b200.ops import fused as K

    dev = _dev()
    torch.manual_seed(5)
    base = torch.randn(70000, device=dev)
    a, b, c = base[:5001], base[5001:5001 + 20000], torch.randn(9, device=dev)     # b's storage is misaligned
    offs = [0, 8192, 40960]                # a: 0..5001 (+pad to 5004), hole, b, hole, c
    flat = torch.full((41000,), 7.0, device=dev)
    K.pack(flat, [a, b, c], offs, scale=1.0, variant=variant)
    torch.cuda.synchronize()
    assert torch.equal(flat[:5001], a) and torch.equal(flat[8192:28192], b) and torch.equal(flat[40960:40969], c)
    assert torch.count_nonzero(flat[5001:5004]) == 0 and torch.count_nonzero(flat[40969:40972]) == 0
    assert torch.all(flat[5004:8192] == 7.0) and torch.all(flat[28192:40960] == 7.0) and torch.all(flat[40972:] == 7.0)
    back = [torch.zeros_like(t) for t in (a, b, c)]
    K.unpack(flat, back, offs, variant=variant)
    torch.cuda.synchronize()
    assert all(torch.equal(x, y) for x, y in zip(back, (a, b, c)))


def test_pack_none_tensor_and_unaligned_views_and_many_tensors():
    from distributed_learning_b200.ops import fused as K

    dev = _dev()
    torch.manual_seed(1)
    base = torch.randn(5000, device=dev)
    view = base[1:4098]                      # data pointer not 16-byte aligned
    n_many = 700                             # more than one table (DLB_MAX_TENSORS = 512)
    many = [torch.randn(17 + i, device=dev) for i in range(n_many)]
    tensors = [view, None] + many
    sizes = [view.numel(), 40] + [t.numel() for t in many]
    offs, total = _layout(sizes, 4)
    flat = torch.full((total,), -1.0, device=dev)
    K.pack(flat, tensors, offs, scale=1.0, dtypes=[torch.float32] * len(sizes), numels=sizes)
    torch.cuda.synchronize()
    assert torch.equal(flat[offs[0]:offs[0] + view.numel()], view)
    assert torch.count_nonzero(flat[offs[1]:offs[1] + 40]) == 0
    for t, o in zip(many, offs[2:]):
        assert torch.equal(flat[o:o + t.numel()], t)


@pytest.mark.parametrize("grad_dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("n", [5, 8192, 100003, 3_000_001])
def test_sgd_flat_matches_torch_sgd(grad_dtype, n):
    from distributed_learning_b200.ops import fused as K

    dev = _dev()
    torch.manual_seed(2)
    p0 = torch.randn(n, device=dev)
    ref_p = torch.nn.Parameter(p0.clone())
    opt = torch.optim.SGD([ref_p], lr=0.01, momentum=0.5)
    p, m = p0.clone(), torch.zeros(n, device=dev)
    shadow = torch.zeros(n, device=dev, dtype=torch.bfloat16)
    h = K.make_hyper(0.01, 0.5)
    for _ in range(4):
        g = torch.randn(n, device=dev).to(grad_dtype)
        ref_p.grad = g.float()
        opt.step()
        K.sgd_flat(p, m, g, h, shadow)
    torch.cuda.synchronize()
    assert torch.allclose(p, ref_p.detach(), rtol=1e-6, atol=1e-7)
    assert torch.allclose(m, opt.state[ref_p]["momentum_buffer"], rtol=1e-6, atol=1e-7)
    assert torch.equal(shadow, p.to(torch.bfloat16))


def test_sgd_flat_weight_decay_nesterov():
    from distributed_learning_b200.ops import fused as K

    dev = _dev()
    torch.manual_seed(3)
    n = 50000
    p0 = torch.randn(n, device=dev)
    ref_p = torch.nn.Parameter(p0.clone())
    opt = torch.optim.SGD([ref_p], lr=0.05, momentum=0.9, weight_decay=1e-2, nesterov=True)
    p, m = p0.clone(), torch.zeros(n, device=dev)
    h = K.make_hyper(0.05, 0.9, weight_decay=1e-2, nesterov=True)
    for _ in range(3):
        g = torch.randn(n, device=dev)
        ref_p.grad = g.clone()
        opt.step()
        K.sgd_flat(p, m, g, h)
    torch.cuda.synchronize()
    assert torch.allclose(p, ref_p.detach(), rtol=1e-5, atol=1e-6)


def test_sgd_multi_matches_torch_sgd():
    from distributed_learning_b200.ops import fused as K

    dev = _dev()
    torch.manual_seed(4)
    sizes = [3, 64, 8192, 70001] + [10 + i for i in range(100)]     # more than one launch (64 per launch)
    ps = [torch.randn(n, device=dev) for n in sizes]
    ref = [torch.nn.Parameter(p.clone()) for p in ps]
    opt = torch.optim.SGD(ref, lr=0.01, momentum=0.5)
    ms = [torch.zeros_like(p) for p in ps]
    h = K.make_hyper(0.01, 0.5)
    for step in range(3):
        gs = [torch.randn_like(p) for p in ps]
        gs[1] = None if step == 1 else gs[1]                         # a parameter without gradient is skipped
        for r, g in zip(ref, gs):
            r.grad = None if g is None else g.clone()
        opt.step()
        K.sgd_multi(ps, ms, gs, h)
    torch.cuda.synchronize()
    for p, r in zip(ps, ref):
        assert torch.allclose(p, r.detach(), rtol=1e-6, atol=1e-7)


@pytest.mark.parametrize("kw", [dict(lr=0.01, momentum=0.5), dict(lr=0.05, momentum=0.9, dampening=0.1, weight_decay=1e-3),
                                dict(lr=0.02, momentum=0.9, nesterov=True), dict(lr=0.1)])
def test_multi_tensor_sgd_optimizer_matches_torch(kw):
    """The drop-in optimizer on a real model: same parameters and same state_dict as torch.optim.SGD."""
    import torch.nn.functional as F

    from distributed_learning_b200.models import CifarNet
    from distributed_learning_b200.ops.optim import MultiTensorSGD

    dev = _dev()
    torch.manual_seed(9)
    a, b = CifarNet().to(dev), CifarNet().to(dev)
    b.load_state_dict(a.state_dict())
    for m in (a, b):
        m.dropout1.p = m.dropout2.p = 0.0
    oa, ob = MultiTensorSGD(a.parameters(), **kw), torch.optim.SGD(b.parameters(), **kw)
    for step in range(4):
        x, y = torch.rand(8, 3, 32, 32, device=dev), torch.randint(10, (8,), device=dev)
        ob.zero_grad()
        F.nll_loss(b(x), y).backward()
        for p, q in zip(a.parameters(), b.parameters()):      # identical gradients for both optimizers:
            p.grad = q.grad.clone()                            # cuDNN's backward is not run-to-run deterministic
        oa.step()
        ob.step()
    for p, q in zip(a.parameters(), b.parameters()):
        assert torch.allclose(p, q, rtol=1e-5, atol=1e-6)
    if kw.get("momentum"):
        sa, sb = oa.state_dict()["state"], ob.state_dict()["state"]
        assert all(torch.allclose(sa[k]["momentum_buffer"], sb[k]["momentum_buffer"], rtol=1e-5, atol=1e-6) for k in sb)
        ob.load_state_dict(oa.state_dict())          # layouts are interchangeable
