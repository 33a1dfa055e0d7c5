"""Real multi-GPU runs of the native backend (skipped with fewer than two GPUs)."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run_workers(nproc, what, free_port, timeout=600):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", str(free_port()),
           os.path.join(ROOT, "tests", "_mp_worker.py"), what]
    r = subprocess.run(cmd, cwd=ROOT, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=timeout)
    assert r.returncode == 0 and "MP_WORKER_OK" in r.stdout, r.stdout[-4000:]
    return r.stdout


def test_allreduce_and_training_two_gpus(free_port):
    """One torchrun for both parts: on a fresh box most of the time is process start-up."""
    out = _run_workers(2, "all", free_port)
    assert "arena mode vmm" in out or "arena mode ipc" in out


def test_all_gpus(free_port):
    n = torch.cuda.device_count()
    if n <= 2:
        pytest.skip("covered by the two-GPU tests")
    _run_workers(min(n, 8), "all", free_port)


def test_main_cli_native_backend(free_port, tmp_path):
    """`main.py 1 0 d d ... --backend b200` spawns one worker per GPU and writes the reference's files."""
    n = min(torch.cuda.device_count(), 2)
    cmd = [sys.executable, os.path.join(ROOT, "main.py"), "1", "0", str(n), str(n), "127.0.0.1", "lo", "cifar",
           str(tmp_path / "data"), "1", "--backend", "b200", "--experiment", "main_ourdist", "--random_input", "1",
           "--limit_batches", "6", "--batch_size", "32", "--results_root", str(tmp_path / "results"),
           "--master_port", str(free_port())]
    r = subprocess.run(cmd, cwd=ROOT, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-4000:]
    folder = tmp_path / "results" / f"main_ourdist_{n}_job"
    for w in range(n):
        rows = (folder / f"ourdist_0_{w}_times.csv").read_text().splitlines()
        assert rows[0].startswith("experiment_name, get_data, data2dev, zero_grad, forward, backprop, sync, optimizer_step, batch")
        assert len(rows) == 7
        assert len((folder / f"ourdist_0_{w}_loss.txt").read_text().splitlines()) == 6
