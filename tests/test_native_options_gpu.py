"""Options of the native wrapper on one GPU: dampening, per-step unused-parameter detection, NVTX."""
import copy

import pytest
import torch
import torch.nn as nn
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


class _TwoHeads(nn.Module):
    def __init__(self):
        super().__init__()
        self.body = nn.Linear(16, 16)
        self.a = nn.Linear(16, 4)
        self.b = nn.Linear(16, 4)
        self.use_b = False

    def forward(self, x):
        h = torch.relu(self.body(x))
        return F.log_softmax((self.b if self.use_b else self.a)(h), dim=1)


def _native_step(wrapper, opt, x, y):
    opt.zero_grad()
    loss = F.nll_loss(wrapper(x), y)
    loss.backward()
    wrapper.sync_gradients()
    opt.step()
    return loss


def test_fused_sgd_with_dampening_matches_torch():
    from distributed_learning_b200.parallel.native import B200Dist

    dev = torch.device("cuda:0")
    torch.manual_seed(1)
    ref = _TwoHeads().to(dev)
    mine = copy.deepcopy(ref)
    kw = dict(lr=0.1, momentum=0.7, dampening=0.3, weight_decay=1e-3)
    ropt = torch.optim.SGD(ref.parameters(), **kw)
    w = B200Dist(mine, None, 1 << 20, dev)
    opt = w.make_optimizer(**kw)
    for i in range(4):
        x, y = torch.rand(8, 16, device=dev), torch.randint(4, (8,), device=dev)
        ropt.zero_grad()
        F.nll_loss(ref(x), y).backward()
        ropt.step()
        _native_step(w, opt, x, y)
    torch.cuda.synchronize()
    for (n, p), q in zip(ref.named_parameters(), mine.parameters()):
        if n.startswith("b."):
            continue                      # never used: torch leaves it alone, the wrapper leaves it out
        assert torch.allclose(p, q, rtol=1e-6, atol=1e-7), n
    w.cleanup()


def test_nesterov_needs_zero_dampening_and_graph_refuses_dampening():
    from distributed_learning_b200.parallel.native import B200Dist

    dev = torch.device("cuda:0")
    w = B200Dist(_TwoHeads().to(dev), None, 1 << 20, dev)
    with pytest.raises(ValueError):
        w.make_optimizer(lr=0.1, momentum=0.5, dampening=0.1, nesterov=True)
    w.cleanup()
    g = B200Dist(_TwoHeads().to(dev), None, 1 << 20, dev, cuda_graph=True)
    with pytest.raises(ValueError):
        g.make_optimizer(lr=0.1, momentum=0.5, dampening=0.1)
    g.cleanup()
    with pytest.raises(ValueError):
        B200Dist(_TwoHeads().to(dev), None, 1 << 20, dev, cuda_graph=True, dynamic_unused=True)


def test_dynamic_unused_trains_when_the_used_set_changes():
    """Default mode refuses a changed used-parameter set; dynamic_unused follows it, giving the
    parameters the loss did not reach a zero gradient that step (what the reference's per-iteration
    walk does, minus the uninitialised memory)."""
    from distributed_learning_b200.parallel.native import B200Dist

    dev = torch.device("cuda:0")
    torch.manual_seed(2)
    ref = _TwoHeads().to(dev)
    mine = copy.deepcopy(ref)
    ropt = torch.optim.SGD(ref.parameters(), lr=0.1, momentum=0.5)
    w = B200Dist(mine, None, 1 << 20, dev, dynamic_unused=True, fused_sgd=True)
    opt = w.make_optimizer(lr=0.1, momentum=0.5)
    for use_b in (False, True, True, False):
        ref.use_b = mine.use_b = use_b
        x, y = torch.rand(8, 16, device=dev), torch.randint(4, (8,), device=dev)
        ropt.zero_grad()
        F.nll_loss(ref(x), y).backward()
        for p in ref.parameters():
            if p.grad is None:
                p.grad = torch.zeros_like(p)       # the comparison model: unused = zero gradient
        ropt.step()
        _native_step(w, opt, x, y)
    torch.cuda.synchronize()
    for (n, p), q in zip(ref.named_parameters(), mine.parameters()):
        assert torch.allclose(p, q, rtol=1e-6, atol=1e-7), n
    w.cleanup()

    strict = B200Dist(copy.deepcopy(ref), None, 1 << 20, dev)
    sopt = strict.make_optimizer(lr=0.1, momentum=0.5)
    strict.model.use_b = False
    x, y = torch.rand(8, 16, device=dev), torch.randint(4, (8,), device=dev)
    _native_step(strict, sopt, x, y)
    strict.model.use_b = True
    sopt.zero_grad()
    F.nll_loss(strict(x), y).backward()
    with pytest.raises(RuntimeError, match="used parameters change"):
        strict.sync_gradients()
    strict.cleanup()


def test_nvtx_ranges_do_not_disturb_a_step():
    from distributed_learning_b200.parallel.native import B200Dist
    from distributed_learning_b200.utils.timing import PhaseTimer

    dev = torch.device("cuda:0")
    w = B200Dist(_TwoHeads().to(dev), None, 1 << 20, dev)
    w.__dict__["nvtx"] = True
    opt = w.make_optimizer()
    t = PhaseTimer(dev, nvtx=True)
    t.start("batch")
    loss = _native_step(w, opt, torch.rand(8, 16, device=dev), torch.randint(4, (8,), device=dev))
    t.end("batch")
    t.end_step("x", {})
    assert float(loss) == float(loss) and t.rows()[0][1]["batch"] > 0
    w.cleanup()


@pytest.mark.parametrize("algo", ["twoshot", "oneshot"])
@pytest.mark.parametrize("dtype", ["fp32", "bf16"])
@pytest.mark.parametrize("graph", [False, True])
def test_whole_fused_gradient_path_on_one_gpu(algo, dtype, graph):
    """comm_kernel_when_alone: the wrapper launches the fused kernel of the multi-GPU path (pack
    prologue + barrier + reduce + SGD epilogue, one launch per bucket) with a world of one."""
    from distributed_learning_b200.parallel.graph_step import GraphedTrainStep
    from distributed_learning_b200.parallel.native import B200Dist

    dev = torch.device("cuda:0")
    torch.manual_seed(4)
    ref = _TwoHeads().to(dev)
    mine = copy.deepcopy(ref)
    ropt = torch.optim.SGD(ref.parameters(), lr=0.1, momentum=0.5)
    w = B200Dist(mine, None, 600, dev, algo=algo, model_dtype=dtype, comm_kernel_when_alone=True, cuda_graph=graph)
    opt = w.make_optimizer(lr=0.1, momentum=0.5)
    batches = [(torch.rand(8, 16, device=dev), torch.randint(4, (8,), device=dev)) for _ in range(4)]
    step = None
    for x, y in batches:
        ropt.zero_grad()
        F.nll_loss(ref(x), y).backward()
        ropt.step()
        if graph:
            step = step or GraphedTrainStep(w, opt, x, y, warmup=2)
            step(x, y)
        else:
            _native_step(w, opt, x, y)
    torch.cuda.synchronize()
    assert len(w.buckets) > 1
    launches_per_step = w.kernel_launches / (len(batches) + (3 if graph else 0))
    assert launches_per_step == len(w.buckets), "one kernel per bucket does pack, reduce and update"
    tol = dict(rtol=1e-6, atol=1e-7) if dtype == "fp32" else dict(rtol=3e-2, atol=3e-3)
    for (n, p), q in zip(ref.named_parameters(), mine.parameters()):
        if not n.startswith("b."):
            assert torch.allclose(p, q.float(), **tol), n
    w.comm.check()
    w.cleanup()
