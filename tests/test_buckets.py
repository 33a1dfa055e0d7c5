"""Bucket rule (reference src/ourdist.py:54-68); golden values from SURVEY.md section 2.7."""
import pytest
import torch

from distributed_learning_b200.models import BasicNet, CifarNet, GoogLeNet, ResNet50
from distributed_learning_b200.parallel.buckets import (Bucket, bucket_byte_sizes, find_unused_parameters,
                                                         fusion_grouping, make_buckets)

KIB, MIB = 1024, 1024 * 1024

GOLDEN = {
    "basicnet": {0: 8, 256 * KIB: 3, MIB: 3, 4 * MIB: 3, 16 * MIB: 1, 25 * MIB: 1, 64 * MIB: 1},
    "googlenet": {0: 187, 256 * KIB: 67, MIB: 34, 4 * MIB: 13, 16 * MIB: 4, 25 * MIB: 3, 64 * MIB: 1},
    "resnet50": {0: 161, 256 * KIB: 89, MIB: 66, 4 * MIB: 32, 16 * MIB: 8, 25 * MIB: 5, 64 * MIB: 2},
}


@pytest.fixture(scope="module")
def nets():
    with torch.device("meta"):
        return {"basicnet": BasicNet(), "googlenet": GoogLeNet(), "resnet50": ResNet50()}


@pytest.mark.parametrize("name", sorted(GOLDEN))
def test_golden_bucket_counts(nets, name):
    params = list(nets[name].parameters())
    for size, want in GOLDEN[name].items():
        assert len(bucket_byte_sizes(params, size)) == want, (name, size)


def test_googlenet_default_bucket_sizes(nets):
    sizes = [s / MIB for s in bucket_byte_sizes(nets["googlenet"].parameters(), 25 * MIB)]
    assert [round(s, 1) for s in sizes] == [20.0, 24.5, 5.2]


def test_resnet50_64mib_bucket_sizes(nets):
    sizes = [s / MIB for s in bucket_byte_sizes(nets["resnet50"].parameters(), 64 * MIB)]
    assert [round(s, 1) for s in sizes] == [62.9, 34.6]


def test_rule_reverse_order_and_oversize_tensor_alone():
    ps = [torch.zeros(n) for n in (10, 1000, 20, 30)]           # bytes 40, 4000, 80, 120
    groups = list(fusion_grouping(ps, 256))
    assert [[t.numel() for t in g] for g in groups] == [[30, 20], [1000], [10]]
    assert [len(g) for g in fusion_grouping(ps, 0)] == [1, 1, 1, 1]


def test_bucket_fuse_unfuse_roundtrip_and_alignment():
    ps = [torch.nn.Parameter(torch.randn(n)) for n in (5, 13, 64)]
    for p in ps:
        p.grad = torch.randn_like(p)
    b = Bucket(0, ps, aligned=True)
    assert all(o % 4 == 0 for o in b.offsets) and b.numel % 4 == 0
    want = [p.grad.clone() for p in ps]
    buf = b.fuse(scale=0.5)
    for p, o, w in zip(ps, b.offsets, want):
        assert torch.equal(buf[o:o + p.numel()], w * 0.5)
        p.grad = None
    b.unfuse(buf * 2)
    for p, w in zip(ps, want):
        assert torch.equal(p.grad, w)
    dense = Bucket(0, ps, aligned=False)
    assert dense.numel == 5 + 13 + 64


def test_unused_parameter_detection_googlenet_aux_heads():
    net = GoogLeNet(num_classes=10).train()
    out = net(torch.rand(2, 3, 64, 64))
    unused = find_unused_parameters(out, net.parameters())
    assert len(unused) == 14
    names = {n for n, p in net.named_parameters() if any(p is u for u in unused)}
    assert all(n.startswith(("model.aux1", "model.aux2")) for n in names)
    assert find_unused_parameters(CifarNet()(torch.rand(2, 3, 32, 32)), CifarNet().parameters()) != []  # other net's params


def test_make_buckets_counts():
    assert len(make_buckets(BasicNet().parameters(), 0)) == 8
