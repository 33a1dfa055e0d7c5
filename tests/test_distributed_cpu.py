"""Multi-process CPU tests over gloo: allreduce algorithms, wrappers x reducers, the 2-step pump,
and whole experiments through main.py (BASELINE.json config 1)."""
import os
import subprocess
import sys

import pytest
import torch
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import _cpu_workers as W  # noqa: E402


def _spawn(fn, world, port, tmp_path):
    mp.spawn(fn, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    assert sorted(os.listdir(tmp_path)) == [f"ok{r}" for r in range(world)]


@pytest.mark.parametrize("world", [2, 3])
def test_allreduce_algorithms_bit_exact(world, free_port, tmp_path):
    _spawn(W.allreduce_worker, world, free_port(), tmp_path)


def test_wrappers_and_reducers_agree(free_port, tmp_path):
    _spawn(W.wrapper_worker, 2, free_port(), tmp_path)


def test_node_aggregate_pump():
    from distributed_learning_b200.parallel.reducers import NodeAggregateReducerCPU

    ctx = mp.get_context("spawn")
    calls = []

    def reduce_fn(buf):          # stands in for the inter-node allreduce
        calls.append(buf.clone())

    node = NodeAggregateReducerCPU(reduce_fn, 2, ctx=ctx)
    procs = [ctx.Process(target=W.pump_worker_child, args=(node.reducers[i], i, [1.0, 5.0, 9.0])) for i in range(2)]
    for p in procs:
        p.start()
    rounds = node.pump()
    for p in procs:
        p.join()
    assert rounds == 3 and [p.exitcode for p in procs] == [0, 0]
    assert torch.allclose(calls[1], torch.full((6,), 5.5))


def _main(args, port, results, timeout=900):
    cmd = [sys.executable, os.path.join(ROOT, "main.py")] + args + ["--results_root", results, "--master_port", str(port)]
    env = dict(os.environ, DLB200_VERBOSE="1")
    return subprocess.Popen(cmd, cwd=ROOT, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)


def _losses(path):
    return [line.split(": ")[1] for line in open(path).read().splitlines()]


def test_experiment3_two_nodes_cpu_bit_identical_losses(free_port, tmp_path):
    """Config 1: the CNN on 2 gloo ranks; every algorithm gives the same loss trajectory to the
    last bit."""
    from distributed_learning_b200.utils.data import write_fake_mnist

    data = write_fake_mnist(str(tmp_path / "data"), n=1024)
    port, results = free_port(), str(tmp_path / "results")
    common = ["1", "2", "127.0.0.1", "lo", "mnist", data, "0", "--experiment", "experiment3", "--limit_batches", "4",
              "--batch_size", "16"]
    procs = [_main(["2", str(r)] + common, port, results) for r in range(2)]
    outs = [p.communicate(timeout=900)[0] for p in procs]
    assert [p.returncode for p in procs] == [0, 0], outs[0][-3000:] + outs[1][-3000:]
    folder = os.path.join(results, "experiment3_2_job")
    names = ["ddp", "onestep_reduce", "onestep_seq_merge", "onestep_overlap"]
    base = _losses(os.path.join(folder, "ddp_0_0_loss.txt"))
    assert len(base) == 4
    for n in names:
        assert _losses(os.path.join(folder, f"{n}_0_0_loss.txt")) == base, n
        head = open(os.path.join(folder, f"{n}_1_0_times.csv")).readline().strip()
        assert head == ("experiment_name, get_data, data2dev, zero_grad, forward, backprop, sync, optimizer_step, "
                        "batch, batch_count, data_len")
        assert os.path.exists(os.path.join(folder, f"{n}_config.txt"))
        worst = open(os.path.join(folder, f"{n}_max_over_ranks_times.csv")).read().splitlines()
        assert worst[0] == head and len(worst) == 5
        mine = [float(x) for x in open(os.path.join(folder, f"{n}_0_0_times.csv")).read().splitlines()[2].split(", ")[1:9]]
        assert all(w >= m - 1e-9 for w, m in zip([float(x) for x in worst[2].split(", ")[1:9]], mine))
    assert os.path.exists(os.path.join(folder, "warmup_0_0_times.csv"))


def test_experiment1_two_step_with_two_workers_and_resume(free_port, tmp_path):
    """2-step process model (d=2 workers + node pump) on one node, then checkpoint/resume."""
    from distributed_learning_b200.utils.data import write_fake_mnist

    data = write_fake_mnist(str(tmp_path / "data"), n=1024)
    results = str(tmp_path / "results")
    ck = str(tmp_path / "ck.pt")
    common = ["1", "0", "2", "2", "127.0.0.1", "lo", "mnist", data, "0", "--batch_size", "16"]
    p = _main(common + ["--experiment", "main_ourdist", "--limit_batches", "4", "--checkpoint", ck], free_port(), results)
    out = p.communicate(timeout=900)[0]
    assert p.returncode == 0, out[-3000:]
    folder = os.path.join(results, "main_ourdist_2_job")
    l0, l1 = (_losses(os.path.join(folder, f"ourdist_0_{w}_loss.txt")) for w in range(2))
    assert len(l0) == 4 and len(l1) == 4
    payload = torch.load(ck, weights_only=False)
    assert payload["batch_count"] == 4 and "conv1.weight" in payload["model"]
    # resume: two more batches, continuing the count
    p = _main(common + ["--experiment", "main_seq_merge", "--limit_batches", "6", "--checkpoint", ck, "--resume", "1",
                        "--job_id", "resumed"], free_port(), results)
    out = p.communicate(timeout=900)[0]
    assert p.returncode == 0, out[-3000:]
    lines = open(os.path.join(results, "main_seq_merge_2_resumed", "seq_merge_0_0_loss.txt")).read().splitlines()
    assert [ln.split(" loss for batch ")[1].split(":")[0] for ln in lines] == ["4", "5"]
    assert torch.load(ck, weights_only=False)["batch_count"] == 6


def test_unknown_experiment_is_rejected(free_port, tmp_path):
    p = _main(["1", "0", "1", "1", "127.0.0.1", "lo", "mnist", "/tmp/none", "0", "--experiment", "experiment_3"],
              free_port(), str(tmp_path))
    out = p.communicate(timeout=300)[0]
    assert p.returncode != 0 and "unknown experiment" in out


def test_registry_has_all_reference_names():
    from distributed_learning_b200.experiments import EXPERIMENTS

    names = {"main_warmup", "main_ourdist", "main_ourdist_nccl", "main_seq", "main_seq_merge", "main_overlap",
             "main_ddp", "main_central_reduce", "main_onestep_reduce", "main_onestep_central", "main_onestep_overlap",
             "main_onestep_seq_merge", "main_single", "experiment1", "experiment2", "experiment3", "experiment_nccl",
             "experiment_ourdist_nccl", "experiment_single", "experiment_onestep_central", "fusion_experiment_ddp",
             "fusion_experiment_ourdist"}
    assert names <= set(EXPERIMENTS) and len(names) == 22
