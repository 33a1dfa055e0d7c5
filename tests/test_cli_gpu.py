"""The command line on the native backend with one GPU, checkpoint/resume through FusedSGD, and the
failure detection of the device-side barrier (needs a GPU)."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = "experiment_name, get_data, data2dev, zero_grad, forward, backprop, sync, optimizer_step, batch, batch_count, data_len"


def _main(args, timeout=600):
    cmd = [sys.executable, os.path.join(ROOT, "main.py")] + args
    r = subprocess.run(cmd, cwd=ROOT, env=dict(os.environ, DLB200_VERBOSE="2"), stdout=subprocess.PIPE,
                       stderr=subprocess.STDOUT, text=True, timeout=timeout)
    assert r.returncode == 0, r.stdout[-4000:]
    return r.stdout


@pytest.mark.parametrize("graph", [0, 1])
def test_main_native_backend_single_gpu(tmp_path, free_port, graph):
    results = tmp_path / "results"
    out = _main(["1", "0", "1", "1", "127.0.0.1", "lo", "cifar", str(tmp_path / "data"), "1", "--backend", "b200",
                 "--experiment", "experiment1", "--random_input", "1", "--limit_batches", "8", "--batch_size", "32",
                 "--results_root", str(results), "--master_port", str(free_port()), "--cuda_graph", str(graph)])
    assert "topology:" in out and "NVIDIA" in out          # bring-up discovers the GPUs instead of a NIC
    folder = results / "experiment1_1_job"
    if graph == 0:      # the 1-step suite too: its main_ddp runs torch DDP over NCCL inside the same group
        _main(["1", "0", "1", "1", "127.0.0.1", "lo", "cifar", str(tmp_path / "data"), "1", "--backend", "b200",
               "--experiment", "experiment2", "--random_input", "1", "--limit_batches", "4", "--batch_size", "32",
               "--results_root", str(results), "--master_port", str(free_port())])
        for name in ("ddp", "onestep_reduce", "onestep_central"):
            assert len((results / "experiment2_1_job" / f"{name}_0_0_times.csv").read_text().splitlines()) == 5, name
        out = _main(["1", "0", "1", "1", "127.0.0.1", "lo", "cifar", str(tmp_path / "data"), "1", "--backend", "b200",
                     "--experiment", "main_seq", "--random_input", "1", "--limit_batches", "3", "--batch_size", "32",
                     "--results_root", str(results), "--master_port", str(free_port())])
        assert "bucket7: 1 tensors" in out          # SeqDist on b200 = one bucket per parameter tensor
        assert (results / "main_seq_1_job" / "seq_0_0_times.csv").exists()
    for name in ("warmup", "ourdist", "seq_merge", "overlap", "central_node_reduce"):
        rows = (folder / f"{name}_0_0_times.csv").read_text().splitlines()
        assert rows[0] == HEADER and len(rows) == 9, name
        losses = (folder / f"{name}_0_0_loss.txt").read_text().splitlines()
        assert len(losses) == 8 and losses[0].startswith("Worker 0:0 loss for batch 0: ")
        assert (folder / f"{name}_config.txt").exists()


def test_native_checkpoint_resume_is_exact(tmp_path, free_port):
    """8 batches in one go == 4 batches, checkpoint, resume for 4 more (same synthetic stream is not
    guaranteed across processes, so compare the checkpointed state instead: momentum and step)."""
    ck = tmp_path / "ck.pt"
    common = ["1", "0", "1", "1", "127.0.0.1", "lo", "cifar", str(tmp_path / "data"), "1", "--backend", "b200",
              "--experiment", "main_ourdist", "--random_input", "1", "--batch_size", "16", "--results_root",
              str(tmp_path / "results"), "--checkpoint", str(ck)]
    _main(common + ["--limit_batches", "4", "--master_port", str(free_port())])
    first = torch.load(ck, weights_only=False)
    assert first["batch_count"] == 4
    state = first["optimizer"]["state"]
    assert len(state) == 8 and all("momentum_buffer" in s for s in state.values())      # torch.optim.SGD layout
    assert first["optimizer"]["param_groups"][0]["lr"] == pytest.approx(0.01)
    _main(common + ["--limit_batches", "6", "--resume", "1", "--job_id", "resumed", "--master_port", str(free_port())])
    second = torch.load(ck, weights_only=False)
    assert second["batch_count"] == 6
    lines = (tmp_path / "results" / "main_ourdist_1_resumed" / "ourdist_0_0_loss.txt").read_text().splitlines()
    assert [ln.split(" loss for batch ")[1].split(":")[0] for ln in lines] == ["4", "5"]
    # the resumed run really started from the saved weights: they changed, but not by much
    a, b = first["model"]["fc2.weight"], second["model"]["fc2.weight"]
    assert not torch.equal(a, b) and (a - b).abs().max() < 0.1
    # a checkpoint written by FusedSGD loads into torch.optim.SGD
    from distributed_learning_b200.models import CifarNet
    net = CifarNet()
    opt = torch.optim.SGD(net.parameters(), lr=0.1, momentum=0.9)
    opt.load_state_dict(second["optimizer"])
    assert opt.param_groups[0]["momentum"] == pytest.approx(0.5)


def test_barrier_timeout_reports_a_missing_peer():
    """A rank whose peer never arrives must not hang: the wait gives up and sets the error word."""
    from distributed_learning_b200.ops._lib import (DLB_E_TIMEOUT, DLB_SIGNAL_BYTES, DLB_STATE_WORDS, DlbError, check,
                                                     lib, ptr_array)

    L = lib()
    dev = torch.device("cuda:0")
    sig = [torch.zeros(DLB_SIGNAL_BYTES, dtype=torch.uint8, device=dev) for _ in range(2)]
    bufs = [torch.ones(1024, device=dev) for _ in range(2)]
    state = torch.zeros(DLB_STATE_WORDS, dtype=torch.int32, device=dev)
    check(L.dlb_set_wait_timeout_ms(50))
    try:
        stream = torch.cuda.current_stream().cuda_stream
        # rank 0 of 2 on its own: "rank 1" is just local memory nobody writes flags into
        check(L.dlb_allreduce_twoshot(ptr_array([b.data_ptr() for b in bufs]), ptr_array([s.data_ptr() for s in sig]),
                                      state.data_ptr(), 0, 2, 1024, 0, 1.0, 2, 64, None, stream))
        assert L.dlb_comm_check(state.data_ptr(), stream) == DLB_E_TIMEOUT
        with pytest.raises(DlbError):
            check(L.dlb_comm_check(state.data_ptr(), stream))
        assert b"never arrived" in L.dlb_last_error()
    finally:
        check(L.dlb_set_wait_timeout_ms(600000))


def test_graphed_step_requires_static_regions():
    from distributed_learning_b200.models import CifarNet
    from distributed_learning_b200.parallel.graph_step import GraphedTrainStep
    from distributed_learning_b200.parallel.native import B200Dist

    dev = torch.device("cuda:0")
    w = B200Dist(CifarNet().to(dev), None, 1 << 20, dev)            # built without cuda_graph=True
    opt = w.make_optimizer()
    with pytest.raises(ValueError):
        GraphedTrainStep(w, opt, torch.rand(4, 3, 32, 32, device=dev), torch.zeros(4, dtype=torch.long, device=dev))
    w.cleanup()
