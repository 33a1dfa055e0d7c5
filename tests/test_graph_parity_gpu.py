"""CUDA-graph mode makes the same updates as eager mode.

``GraphedTrainStep`` warms up with real training steps before it captures; those steps must leave
no trace (parameters, momentum, BatchNorm statistics, step count), otherwise the first batch is
applied several times and the run drifts from the eager trajectory and from the checkpoint's
``batch_count``.
"""
import pytest
import torch
import torch.nn as nn
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


class _Net(nn.Module):
    def __init__(self):
        super().__init__()
        self.c1 = nn.Conv2d(3, 8, 3, padding=1)
        self.bn = nn.BatchNorm2d(8)
        self.c2 = nn.Conv2d(8, 16, 3, padding=1)
        self.fc = nn.Linear(16 * 8 * 8, 10)

    def forward(self, x):
        x = F.max_pool2d(F.relu(self.bn(self.c1(x))), 2)
        x = F.max_pool2d(F.relu(self.c2(x)), 2)
        return F.log_softmax(self.fc(x.flatten(1)), dim=1)


def _run(graph: bool, batches, dtype="fp32"):
    from distributed_learning_b200.parallel.graph_step import GraphedTrainStep
    from distributed_learning_b200.parallel.native import B200Dist

    dev = torch.device("cuda:0")
    torch.manual_seed(7)
    model = _Net().to(dev).train()
    wrapper = B200Dist(model, None, 1 << 20, dev, cuda_graph=graph, model_dtype=dtype)
    opt = wrapper.make_optimizer(lr=0.05, momentum=0.5)
    losses = []
    if graph:
        step = GraphedTrainStep(wrapper, opt, batches[0][0], batches[0][1], warmup=3)
        for x, y in batches:
            losses.append(step(x, y).clone())
    else:
        for x, y in batches:
            opt.zero_grad()
            out = wrapper(x)
            loss = F.nll_loss(out.float(), y)
            loss.backward()
            wrapper.sync_gradients()
            opt.step()
            losses.append(loss.detach().clone())
    torch.cuda.synchronize()
    state = {k: v.detach().float().clone() for k, v in model.state_dict().items()}
    mom = opt.state_dict()["state"]
    steps = opt.steps
    wrapper.cleanup()
    return torch.stack(losses).float().cpu(), state, mom, steps


@pytest.mark.parametrize("dtype", ["fp32", "bf16"])
def test_graph_mode_matches_eager_after_n_batches(dtype):
    dev = torch.device("cuda:0")
    g = torch.Generator(device=dev).manual_seed(3)
    batches = [(torch.rand(16, 3, 32, 32, device=dev, generator=g), torch.randint(10, (16,), device=dev, generator=g))
               for _ in range(4)]
    l_e, s_e, m_e, n_e = _run(False, batches, dtype)
    l_g, s_g, m_g, n_g = _run(True, batches, dtype)
    assert n_e == n_g == 4, "the warm-up steps of the graph must not count as updates"
    tol = dict(rtol=1e-5, atol=1e-6) if dtype == "fp32" else dict(rtol=2e-2, atol=2e-3)
    assert torch.allclose(l_e, l_g, **tol), (l_e, l_g)
    for k in s_e:
        assert torch.allclose(s_e[k], s_g[k], **tol), f"{k} differs between eager and graph mode"
    for k in m_e:
        assert torch.allclose(m_e[k]["momentum_buffer"].float(), m_g[k]["momentum_buffer"].float(), **tol)
