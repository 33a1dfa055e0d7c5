"""``NativeReducer``: the native allreduce behind the reference's put/get reducer surface, combined
with the host-driven wrappers (needs a GPU; world size 1 here, real ranks in ``_mp_worker.py``)."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("wrapper_name", ["OurDist", "SeqMergeDist", "SeqDist"])
def test_host_driven_wrappers_over_native_reducer(wrapper_name):
    from distributed_learning_b200.models import CifarNet
    from distributed_learning_b200.parallel import wrappers
    from distributed_learning_b200.parallel.native import NativeReducer

    dev = torch.device("cuda:0")
    torch.manual_seed(2)
    net, ref = CifarNet().to(dev), CifarNet().to(dev)
    ref.load_state_dict(net.state_dict())
    for m in (net, ref):
        m.dropout1.p = m.dropout2.p = 0.0
    reducer = NativeReducer(capacity_bytes=8 << 20, device=dev)
    cls = getattr(wrappers, wrapper_name)
    w = cls(net, reducer, 1 << 20, dev)
    x, y = torch.rand(8, 3, 32, 32, device=dev), torch.randint(10, (8,), device=dev)
    for _ in range(2):
        for m in (net, ref):
            m.zero_grad(set_to_none=True)
        F.nll_loss(w(x), y).backward()
        w.sync_gradients()
        F.nll_loss(ref(x), y).backward()
        torch.cuda.synchronize()
        for p, q in zip(net.parameters(), ref.parameters()):
            assert torch.allclose(p.grad, q.grad, rtol=1e-5, atol=1e-6)      # one rank: the average is the gradient
    w.cleanup()
    reducer.cleanup()


def test_native_reducer_rejects_what_it_cannot_reduce():
    from distributed_learning_b200.parallel.native import NativeReducer

    dev = torch.device("cuda:0")
    r = NativeReducer(capacity_bytes=1 << 20, device=dev)
    with pytest.raises(TypeError):
        r.put(torch.zeros(8, device=dev, dtype=torch.float16))
    with pytest.raises(ValueError):
        r.put(torch.zeros(1 << 20, device=dev))                  # 4 MiB > 1 MiB capacity
    with pytest.raises(ValueError):
        r.put(torch.zeros(8, 8, device=dev).t())                 # not contiguous
    r.cleanup()
