"""CPU tests: models, CLI, timing CSV, data pipeline, checkpoint, library loading."""
import os

import pytest
import torch

from distributed_learning_b200.models import BasicNet, CifarNet, GoogLeNet, ResNet50, create_network
from distributed_learning_b200.utils import timing
from distributed_learning_b200.utils.config import parse_args, picklable_config
from distributed_learning_b200.utils.data import (DataPartitioner, MNISTDataset, get_partition_dataset,
                                                   random_data_generator, write_fake_image_folder, write_fake_mnist)
from distributed_learning_b200.utils.logging import eval_arg

ARGV = ["2", "1", "1", "2", "127.0.0.1", "lo", "mnist", "/tmp/x", "0"]


# ---- models -----------------------------------------------------------------------------------
def _shapes(m):
    return [(n, tuple(p.shape)) for n, p in m.named_parameters()]


def test_basicnet_matches_reference_shape():
    m = BasicNet()
    assert len(list(m.parameters())) == 8
    assert sum(p.numel() for p in m.parameters()) == 1_199_882
    assert m(torch.rand(2, 1, 28, 28)).shape == (2, 10)
    c = CifarNet()
    assert c.fc1.in_features == 64 * 14 * 14
    out = c(torch.rand(3, 3, 32, 32))
    assert out.shape == (3, 10) and torch.allclose(out.exp().sum(1), torch.ones(3), atol=1e-5)


def test_googlenet_and_resnet_match_torchvision_parameter_layout():
    tv = pytest.importorskip("torchvision.models")
    with torch.device("meta"):
        mine, theirs = GoogLeNet(), tv.googlenet(weights=None, init_weights=False)
        assert [(n.replace("model.", "", 1), s) for n, s in _shapes(mine)] == _shapes(theirs)
        assert sum(p.numel() for p in mine.parameters()) == 13_004_888
        mine, theirs = ResNet50(), tv.resnet50(weights=None)
        assert [(n.replace("model.", "", 1), s) for n, s in _shapes(mine)] == _shapes(theirs)
        assert sum(p.numel() for p in mine.parameters()) == 25_557_032


def test_googlenet_forward_equals_torchvision_with_same_weights():
    tv = pytest.importorskip("torchvision.models")
    torch.manual_seed(0)
    theirs = tv.googlenet(weights=None, init_weights=False, num_classes=10).eval()
    mine = GoogLeNet(num_classes=10).eval()
    mine.model.load_state_dict(theirs.state_dict())
    x = torch.rand(2, 3, 96, 96)
    with torch.no_grad():
        assert torch.allclose(mine(x), torch.log_softmax(theirs(x), 1), atol=1e-5)


def test_registry():
    assert set(create_network) >= {"mnist", "imagenet", "cifar", "resnet50"}


# ---- CLI -----------------------------------------------------------------------------------------
def test_cli_reference_defaults_and_positionals():
    c = parse_args(ARGV)
    assert (c.size, c.rank, c.node_dev, c.total_dev, c.master_addr, c.ifname) == (2, 1, 1, 2, "127.0.0.1", "lo")
    assert (c.job_id, c.experiment, c.limit_batches, c.backend, c.batch_size, c.random_input) == \
        ("job", "main_ourdist", 3, "gloo", 128, 0)
    assert c.epoch_count == 100 and c.grouping_size == 25 * 1024 * 1024
    assert c.devices == ["cpu"] and c.create_network is BasicNet
    g = parse_args(["1", "0", "4", "4", "a", "i", "imagenet", "/r", "1", "--backend", "b200"])
    assert g.devices == ["cuda:0", "cuda:1", "cuda:2", "cuda:3"]
    import pickle
    pickle.dumps(picklable_config(c))
    pickle.dumps(c)                      # the whole config crosses `spawn` in the 2-step experiments


def test_cli_envarg(monkeypatch):
    monkeypatch.setenv("MY_SIZE", "4")
    monkeypatch.setenv("MY_RANK", "3")
    c = parse_args(["envarg://MY_SIZE", "envarg://MY_RANK", "1", "4"] + ARGV[4:])
    assert (c.size, c.rank) == (4, 3)
    assert eval_arg("plain") == "plain"
    with pytest.raises(KeyError):
        eval_arg("envarg://DOES_NOT_EXIST_XYZ")


def test_cli_rejects_bad_values():
    with pytest.raises(SystemExit):
        parse_args(ARGV[:6] + ["nosuchmodel"] + ARGV[7:])
    with pytest.raises(SystemExit):
        parse_args(["2", "0", "2", "3"] + ARGV[4:])


# ---- timing --------------------------------------------------------------------------------------
def test_timing_csv_is_byte_compatible(tmp_path):
    t = timing.PhaseTimer(mode="host")
    for k in range(2):
        t.start("batch")
        for ph in ("get_data", "data2dev", "zero_grad", "forward", "backprop", "sync", "optimizer_step"):
            t.start(ph)
            t.end(ph)
        t.end("batch")
        t.end_step("ourdist", {"batch_count": k, "data_len": 32})
    path = tmp_path / "t.csv"
    t.writeout(str(path))
    lines = path.read_text().splitlines()
    assert lines[0] == ("experiment_name, get_data, data2dev, zero_grad, forward, backprop, sync, "
                        "optimizer_step, batch, batch_count, data_len")
    assert lines[1].startswith("ourdist, ") and lines[1].endswith(", 0, 32") and len(lines) == 3
    assert len(lines[1].split(", ")) == 11


def test_timing_module_level_api_and_means(tmp_path):
    timing.set_default_timer(timing.PhaseTimer(mode="host"))
    timing.start_timer("a")
    timing.end_timer("a")
    timing.start_timer("a")
    timing.end_timer("a")
    timing.end_timing_experiment("x", {"batch_count": 0})
    means = timing.get_default_timer().phase_means()
    assert means["a"] >= 0
    timing.writeout_timer(str(tmp_path / "m.csv"))
    assert (tmp_path / "m.csv").read_text().startswith("experiment_name, a, batch_count")
    with pytest.raises(KeyError):
        timing.end_timer("never_started")


# ---- data ------------------------------------------------------------------------------------------
def test_partitioner_disjoint_whole_batches():
    torch.manual_seed(1234)
    parts = DataPartitioner(list(range(1000)), total_dev=3, batch_size=32).partitions
    assert all(len(p) == 320 for p in parts)
    assert len(set(torch.cat(parts).tolist())) == 960
    with pytest.raises(ValueError):
        DataPartitioner(list(range(10)), 4, 8)


def test_mnist_reader_and_loader(tmp_path):
    root = write_fake_mnist(str(tmp_path), n=256)
    ds = MNISTDataset(os.path.join(root, "mnist"))
    img, label = ds[0]
    assert img.shape == (1, 28, 28) and 0 <= label < 10 and len(ds) == 256
    torch.manual_seed(1234)
    loader = get_partition_dataset["mnist"](0, 1, 2, 2, root, 16, num_workers=0)
    x, y = next(iter(loader))
    assert x.shape == (16, 1, 28, 28) and y.dtype == torch.int64 and len(loader) == 8


def test_image_folder_and_synthetic_fallback(tmp_path):
    pytest.importorskip("torchvision")
    root = write_fake_image_folder(str(tmp_path), n=16, classes=2, size=40)
    x, y = next(iter(get_partition_dataset["imagenet"](0, 0, 1, 1, root, 4, num_workers=0)))
    assert x.shape == (4, 3, 224, 224)
    x, y = next(iter(get_partition_dataset["cifar"](0, 0, 1, 1, str(tmp_path / "none"), 8, num_workers=0)))
    assert x.shape == (8, 3, 32, 32)


def test_random_data_generator_shapes():
    gen = random_data_generator((torch.zeros(4, 1, 28, 28), torch.tensor([0, 9, 3, 1])))
    x, y = next(gen)
    assert x.shape == (4, 1, 28, 28) and int(y.max()) <= 9 and 0 <= float(x.min()) and float(x.max()) < 1


# ---- checkpoint ------------------------------------------------------------------------------------
def test_checkpoint_roundtrip(tmp_path):
    from distributed_learning_b200.utils.checkpoint import load_checkpoint, save_checkpoint

    torch.manual_seed(0)
    m = CifarNet()
    opt = torch.optim.SGD(m.parameters(), lr=0.01, momentum=0.5)
    m(torch.rand(2, 3, 32, 32)).sum().backward()
    opt.step()
    path = save_checkpoint(str(tmp_path / "c" / "ck.pt"), m, opt, batch_count=7, epoch=1, config=parse_args(ARGV))
    m2 = CifarNet()
    opt2 = torch.optim.SGD(m2.parameters(), lr=0.01, momentum=0.5)
    payload = load_checkpoint(path, m2, opt2)
    assert payload["batch_count"] == 7 and payload["epoch"] == 1 and payload["config"]["experiment"] == "main_ourdist"
    for a, b in zip(m.parameters(), m2.parameters()):
        assert torch.equal(a, b)
    sa, sb = opt.state_dict()["state"], opt2.state_dict()["state"]
    assert all(torch.equal(sa[k]["momentum_buffer"], sb[k]["momentum_buffer"]) for k in sa)


# ---- native library (host-only parts work without a GPU) ---------------------------------------------
def test_library_builds_loads_and_reports_errors():
    from distributed_learning_b200.ops import _lib

    L = _lib.lib()
    assert L.dlb_version() >= 100
    assert L.dlb_emulated_flags_bytes(4) == 4 * (_lib.DLB_SIGNAL_BYTES + 4 * _lib.DLB_STATE_WORDS)
    rc = L.dlb_pack(None, 7, None, None, None, None, 0, 1.0, None)       # bad dtype -> DLB_E_ARG + message
    assert rc == _lib.DLB_E_ARG and b"dtype" in L.dlb_last_error()
    with pytest.raises(_lib.DlbError):
        _lib.check(rc)


def test_fd_passing_between_processes():
    """SCM_RIGHTS plumbing used to share VMM handles: send a pipe's read end through a socket."""
    import ctypes
    import multiprocessing as mp

    from distributed_learning_b200.ops import _lib

    L = _lib.lib()
    name = f"dlb200-test-{os.getpid()}".encode()
    lfd = ctypes.c_int()
    _lib.check(L.dlb_fd_listen(name, ctypes.byref(lfd)))

    def child(name):
        from distributed_learning_b200.ops import _lib as l2
        r, w = os.pipe()
        os.write(w, b"hello")
        l2.check(l2.lib().dlb_fd_send(name, r, 42, 5000))

    p = mp.get_context("fork").Process(target=child, args=(name,))
    p.start()
    fd, tag = ctypes.c_int(), ctypes.c_int()
    _lib.check(L.dlb_fd_accept(lfd.value, ctypes.byref(fd), ctypes.byref(tag), 10000))
    p.join()
    assert tag.value == 42 and os.read(fd.value, 5) == b"hello"
    L.dlb_fd_close(fd.value)
    L.dlb_fd_close(lfd.value)


def test_legacy_partition_api(tmp_path):
    from distributed_learning_b200.utils.data import partition_mnist

    root = write_fake_mnist(str(tmp_path), n=512)
    torch.manual_seed(1234)
    loaders, bs = partition_mnist(2, 4, root, 16, node_id=1)
    assert len(loaders) == 2 and bs == 16 and next(iter(loaders[1]))[0].shape == (16, 1, 28, 28)


def test_summarize_results(tmp_path):
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "benchmarks"))
    from summarize_results import summarise_folder

    head = "experiment_name, get_data, data2dev, zero_grad, forward, backprop, sync, optimizer_step, batch, batch_count, data_len\n"
    for w, batch in ((0, 10.0), (1, 20.0)):
        with open(tmp_path / f"ourdist_0_{w}_times.csv", "w") as f:
            f.write(head)
            for k in range(8):
                f.write(f"ourdist, 1.0, 0.1, 0.1, 3.0, 4.0, {w + 1.0}, 0.5, {batch if k >= 5 else 999.0}, {k}, 128\n")
    s = summarise_folder(str(tmp_path), skip=5)["ourdist"]
    assert s["workers"] == 2 and s["batch_ms"] == 20.0                  # mean over batches >= 5, max over workers
    assert s["images_per_sec"] == pytest.approx(2 * 128 / 20.0 * 1000.0)
    assert s["phases_ms"]["sync"] == 2.0


def test_launcher_port_and_backend_selection(monkeypatch):
    from distributed_learning_b200 import launcher

    c = parse_args(ARGV + ["--master_port", "31234"])
    assert launcher.master_port(c) == 31234
    monkeypatch.setenv("MASTER_PORT", "30000")
    assert launcher.master_port(parse_args(ARGV)) == 30000
    monkeypatch.delenv("MASTER_PORT")
    assert launcher.master_port(parse_args(ARGV)) == 29501              # the reference's fixed port
    g = parse_args(["1", "0", "1", "1", "a", "i", "mnist", "/r", "1", "--backend", "b200"])
    # the product backend never creates an NCCL communicator: gloo carries the bootstrap
    assert launcher._group_backend(g, one_process_per_gpu=True) == "gloo"
    assert launcher._group_backend(g, one_process_per_gpu=False) == "gloo"
    n = parse_args(["1", "0", "1", "1", "a", "i", "mnist", "/r", "1", "--backend", "nccl"])
    assert launcher._group_backend(n, one_process_per_gpu=True) == "cpu:gloo,cuda:nccl"
    assert launcher._group_backend(parse_args(ARGV), True) == "gloo"


# ---- variant / grid selection of the native communicator (pure host logic) -------------------------
def _fake_comm(world, multicast):
    from distributed_learning_b200.ops.comm import NativeComm

    c = object.__new__(NativeComm)
    c.world, c.rank, c.mc_ptr = world, 0, (1 if multicast else None)
    c.tuned, c.tuned_ms = {}, {}
    c._gather = lambda obj: [obj] * world          # stands in for the collective: every rank sees rank 0's object
    return c


def test_choose_heuristics_and_fused_cost_model(tmp_path, monkeypatch):
    from distributed_learning_b200.ops._lib import dtype_code

    c8 = _fake_comm(8, True)
    assert c8.choose(64 << 10, torch.float32)[0] == "oneshot"
    assert c8.choose(16 << 20, torch.float32) == ("nvls", 16, 4)
    assert c8.choose(16 << 20, torch.float32, hbm_bytes=7 * (16 << 20)) == ("nvls", 64, 4), "a fused kernel takes the full grid"
    c2 = _fake_comm(2, True)
    assert c2.choose(4 << 20, torch.float32)[0] == "oneshot" and c2.choose(64 << 20, torch.float32)[0] == "twoshot"
    assert _fake_comm(8, False).choose(16 << 20, torch.float32)[0] == "twoshot"
    assert c8.choose(16 << 20, torch.float32, algo="nvls_deep") == ("nvls", 16, 8)
    assert _fake_comm(8, False).choose(16 << 20, torch.float32, algo="nvls")[0] == "twoshot"

    # measured table: the bare allreduce is fastest on 8 CTAs, the fused kernel is not
    key = (dtype_code(torch.float32), (24 << 20).bit_length())
    c8.tuned[key] = ("nvls", 8, 4)
    c8.tuned_ms[key] = {("nvls", 8, 4): 0.060, ("nvls", 16, 4): 0.062, ("nvls", 64, 4): 0.075, ("twoshot", 64, 0): 0.080}
    assert c8.choose(24 << 20, torch.float32) == ("nvls", 8, 4)
    assert c8.choose(24 << 20, torch.float32, hbm_bytes=7 * (24 << 20)) == ("nvls", 64, 4)
    assert c8.choose(24 << 20, torch.float32, hbm_bytes=1 << 10) == ("nvls", 8, 4), "negligible extra traffic changes nothing"

    # the table survives a round trip through the cache file
    monkeypatch.setenv("DLB200_TUNE_CACHE", str(tmp_path / "tune.json"))
    c8._save_tuned()
    fresh = _fake_comm(8, True)
    assert fresh._load_tuned() and fresh.tuned == c8.tuned and fresh.tuned_ms == c8.tuned_ms
    (tmp_path / "tune.json").write_text('{"0:25": ["nvls", 8, 4]}')          # a file of an older layout
    assert not _fake_comm(8, True)._load_tuned()


def test_broadcast_tensors_stages_every_byte_exactly_once():
    """Host logic of NativeComm.broadcast_tensors with a stand-in for the kernel: the "broadcast"
    adds 1 to every staged byte, so afterwards every byte of every tensor must be its old value + 1 --
    whatever the arena size (chunking), tensor sizes (odd byte counts) and memory formats."""
    from distributed_learning_b200.ops.comm import NativeComm

    for cap in (64, 1000, 1 << 16):
        c = object.__new__(NativeComm)
        c.world, c.rank, c.data_offset = 2, 0, 32
        c.nbytes = 32 + cap
        c.local = torch.zeros(c.nbytes, dtype=torch.uint8)
        calls = []

        def fake_broadcast(offset, nbytes, root=0, _c=c, _calls=calls):
            assert offset == 0 and nbytes % 16 == 0 and nbytes <= _c.nbytes - _c.data_offset
            _c.local[_c.data_offset:_c.data_offset + nbytes] += 1
            _calls.append(nbytes)

        c.broadcast = fake_broadcast
        torch.manual_seed(cap)
        tensors = [torch.randint(0, 200, (n,), dtype=torch.uint8) for n in (1, 15, 16, 17, 999, 5000)]
        tensors.append(torch.randn(3, 5, 7))
        tensors.append(torch.randn(2, 3, 4, 5).contiguous(memory_format=torch.channels_last))   # not contiguous
        tensors.append(torch.randn(6, 4).t())                                                   # a transposed view
        before = [t.clone() for t in tensors]
        moved = c.broadcast_tensors(tensors, root=0)
        assert moved == sum(t.numel() * t.element_size() for t in tensors)
        for t, b in zip(tensors, before):
            got = t.contiguous().view(-1).view(torch.uint8)
            want = b.contiguous().view(-1).view(torch.uint8) + 1
            assert torch.equal(got, want), (cap, tuple(t.shape))
        assert len(calls) >= (1 if cap > 20000 else 2)
