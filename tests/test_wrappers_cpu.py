"""Single-process CPU tests of the host-driven wrappers (no process group: the reduce function is a
stand-in), of the failure path, and of the fusion-sweep folder layout through main.py."""
import os
import subprocess
import sys

import pytest
import torch
import torch.nn as nn
import torch.nn.functional as F

from distributed_learning_b200.parallel.reducers import ReduceImmediately
from distributed_learning_b200.parallel.wrappers import OurDist, SeqDist, SeqMergeDist, SingleDist, WarmupDist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class TwoHeads(nn.Module):
    """One head is never used by the loss (like GoogLeNet's aux heads)."""

    def __init__(self):
        super().__init__()
        self.body = nn.Linear(8, 16)
        self.head = nn.Linear(16, 4)
        self.unused_head = nn.Linear(16, 4)

    def forward(self, x):
        return F.log_softmax(self.head(torch.relu(self.body(x))), dim=1)


def _halve(buf):            # "allreduce" over two identical ranks that averages: identity; make it visible instead
    buf *= 0.5


@pytest.mark.parametrize("cls,kwargs", [(OurDist, dict(grouping_size=64)), (OurDist, dict(grouping_size=0)),
                                        (OurDist, dict(grouping_size=1 << 20)), (SeqMergeDist, dict(grouping_size=128))])
def test_bucketed_wrappers_reduce_every_used_gradient_once(cls, kwargs):
    torch.manual_seed(0)
    net, ref = TwoHeads(), TwoHeads()
    ref.load_state_dict(net.state_dict())
    w = cls(net, ReduceImmediately(_halve), kwargs["grouping_size"])
    x, y = torch.rand(5, 8), torch.randint(4, (5,))
    for _ in range(3):                                   # several iterations: events and counters are re-armed
        for m in (net, ref):
            m.zero_grad(set_to_none=True)
        F.nll_loss(w(x), y).backward()
        w.sync_gradients()
        F.nll_loss(ref(x), y).backward()
        for (n, p), q in zip(net.named_parameters(), ref.parameters()):
            if n.startswith("unused_head"):
                assert p.grad is None                    # left out, not filled with uninitialised memory
            else:
                assert torch.allclose(p.grad, q.grad * 0.5), n
    assert len(w.unused) == 2
    assert all(not any(t is u for u in w.unused) for g in w.groups for t in g.tensors)
    w.cleanup()


def test_seqdist_and_controls():
    torch.manual_seed(1)
    net = TwoHeads()
    w = SeqDist(net, ReduceImmediately(_halve))
    x, y = torch.rand(3, 8), torch.randint(4, (3,))
    F.nll_loss(w(x), y).backward()
    before = net.head.weight.grad.clone()
    w.sync_gradients()
    assert torch.allclose(net.head.weight.grad, before * 0.5)
    assert w.head is net.head                            # attribute passthrough (reference __getattr__)
    s = SingleDist(net)
    s.sync_gradients()
    s.cleanup()
    wu = WarmupDist(net)
    out1, out2 = wu(x), wu(x)
    assert out1 is out2 and out1.requires_grad and float(out1.abs().sum()) == 0.0
    F.nll_loss(out1, y).backward()                       # the loop can still call backward


def test_ourdist_reraises_a_failure_of_its_helper_thread():
    def broken(_buf):
        raise ValueError("link down")

    net = TwoHeads()
    w = OurDist(net, ReduceImmediately(broken), 1 << 20)
    F.nll_loss(w(torch.rand(2, 8)), torch.randint(4, (2,))).backward()
    with pytest.raises(RuntimeError, match="helper thread"):
        w.sync_gradients()
    w.cleanup()


def test_fusion_sweep_folder_layout_cpu(tmp_path, free_port):
    """`fusion_experiment_ourdist`: warm-up, then one sub-folder per bucket size in KiB
    (reference src/main.py:287-295)."""
    from distributed_learning_b200.utils.data import write_fake_mnist

    data = write_fake_mnist(str(tmp_path / "data"), n=512)
    cmd = [sys.executable, os.path.join(ROOT, "main.py"), "1", "0", "1", "1", "127.0.0.1", "lo", "mnist", data, "0",
           "--experiment", "fusion_experiment_ourdist", "--limit_batches", "2", "--batch_size", "16",
           "--results_root", str(tmp_path / "results"), "--master_port", str(free_port())]
    r = subprocess.run(cmd, cwd=ROOT, env=dict(os.environ, DLB200_VERBOSE="1"), stdout=subprocess.PIPE,
                       stderr=subprocess.STDOUT, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:]
    folder = tmp_path / "results" / "fusion_experiment_ourdist_1_job"
    assert (folder / "warmup_0_0_times.csv").exists()
    for size_k in (256, 1024, 4096, 16384, 65536):
        assert (folder / str(size_k) / "ourdist_0_0_times.csv").exists(), size_k
        assert f"grouping_size={size_k * 1024}" in (folder / str(size_k) / "ourdist_config.txt").read_text()


def test_b200_backend_needs_gpu_flag():
    from distributed_learning_b200.utils.config import parse_args

    with pytest.raises(SystemExit):
        parse_args(["1", "0", "1", "1", "a", "i", "mnist", "/r", "0", "--backend", "b200"])
    with pytest.raises(SystemExit):
        parse_args(["2", "0", "8", "16", "a", "i", "mnist", "/r", "1", "--backend", "b200"])


# ---- a changing set of used parameters ---------------------------------------------------------------
class _TwoHeads(torch.nn.Module):
    """The loss reaches head `a` or head `b` depending on a flag: the used-parameter set changes."""

    def __init__(self):
        super().__init__()
        self.body = torch.nn.Linear(6, 6)
        self.a = torch.nn.Linear(6, 3)
        self.b = torch.nn.Linear(6, 3)
        self.use_b = False

    def forward(self, x):
        h = torch.relu(self.body(x))
        return torch.log_softmax((self.b if self.use_b else self.a)(h), dim=1)


class _LocalReducer:
    def __init__(self):
        import queue

        self.q = queue.Queue()

    def put(self, buf):
        self.q.put(buf)

    def get(self):
        return self.q.get()

    def cleanup(self):
        pass


def test_ourdist_raises_instead_of_hanging_when_the_used_set_changes():
    import pytest

    from distributed_learning_b200.parallel.wrappers import OurDist

    torch.manual_seed(0)
    m = _TwoHeads()
    w = OurDist(m, _LocalReducer(), 1 << 20)
    x = torch.rand(4, 6)
    w(x).sum().backward()
    w.sync_gradients()                      # first step: head b is found unused and left out
    m.use_b = True
    for p in m.parameters():
        p.grad = None
    w(x).sum().backward()                   # now head a gets no gradient: its bucket never fills
    with pytest.raises(RuntimeError, match="set of parameters"):
        w.sync_gradients()
    w.cleanup()


def test_dynamic_unused_follows_a_changing_used_set():
    from distributed_learning_b200.parallel.wrappers import OurDist, SeqMergeDist

    for cls in (OurDist, SeqMergeDist):
        torch.manual_seed(0)
        m = _TwoHeads()
        w = cls(m, _LocalReducer(), 1 << 20, dynamic_unused=True)
        x = torch.rand(4, 6)
        for use_b in (False, True, False):
            m.use_b = use_b
            for p in m.parameters():
                p.grad = None
            w(x).sum().backward()
            w.sync_gradients()
            used, unused = (m.b, m.a) if use_b else (m.a, m.b)
            assert used.weight.grad is not None and float(used.weight.grad.abs().sum()) > 0
            assert unused.weight.grad is not None and float(unused.weight.grad.abs().sum()) == 0, \
                "a parameter the loss did not reach gets a zero gradient, not stale or uninitialised memory"
        w.cleanup()
