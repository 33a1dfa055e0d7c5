"""The optional fp32-master entry of a checkpoint, with a stand-in wrapper (the real one needs CUDA)."""
import torch

from distributed_learning_b200.models import CifarNet
from distributed_learning_b200.utils.checkpoint import load_checkpoint, save_checkpoint


class FakeNativeWrapper:
    """Has the two hooks ``B200Dist`` offers: bf16 parameters in the model, fp32 masters on the side."""

    def __init__(self, net):
        self.__dict__["model"] = net
        self.masters = {n: p.detach().float() + 1e-4 for n, p in net.named_parameters()}   # low bits bf16 cannot hold
        self.reloaded = 0

    def master_state_dict(self):
        return {k: v.clone() for k, v in self.masters.items()}

    def load_master_state_dict(self, sd):
        self.masters = {k: v.clone() for k, v in sd.items()}

    def parameters_were_reloaded(self):
        self.reloaded += 1


def test_master_weights_roundtrip_through_a_checkpoint(tmp_path):
    torch.manual_seed(0)
    net = CifarNet().to(torch.bfloat16)
    w = FakeNativeWrapper(net)
    want = w.master_state_dict()
    path = save_checkpoint(str(tmp_path / "ck.pt"), w, None, batch_count=3)
    w2 = FakeNativeWrapper(CifarNet().to(torch.bfloat16))
    payload = load_checkpoint(path, w2, None)
    assert set(payload["master_fp32"]) == set(want) and w2.reloaded == 1
    for k, v in want.items():
        assert torch.equal(w2.masters[k], v)                       # exact fp32 values, not bf16-rounded
        assert not torch.equal(v, v.to(torch.bfloat16).float())    # ...which bf16 alone would have lost
    for a, b in zip(net.parameters(), w2.model.parameters()):
        assert torch.equal(a, b)


def test_plain_models_get_no_master_entry(tmp_path):
    net = CifarNet()
    path = save_checkpoint(str(tmp_path / "ck.pt"), net, None, batch_count=0)
    assert "master_fp32" not in torch.load(path, weights_only=True)
