"""Oracle test: the UNMODIFIED reference and this framework, same command line, same seed, same
fabricated MNIST files, two gloo ranks on the CPU -- the loss trajectories must agree at the
precision the reference prints (4 decimals of ``tensor(...)``).  SURVEY.md section 4: "Oracle = the
reference itself on CPU/gloo".  Skipped where the reference sources are not present."""
import os
import re
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CANDIDATES = [os.path.join(ROOT, "baseline", "_ref", "src", "main.py"), "/root/reference/src/main.py"]
REF_MAIN = next((p for p in CANDIDATES if os.path.exists(p)), None)

pytestmark = pytest.mark.skipif(REF_MAIN is None, reason="reference sources not available")


def _run_pair(main_py, extra, cwd, env=None):
    procs = []
    for r in range(2):
        cmd = [sys.executable, main_py, "2", str(r), "1", "2", "127.0.0.1", "lo", "mnist"] + extra
        procs.append(subprocess.Popen(cmd, cwd=cwd, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True))
    outs = [p.communicate(timeout=900)[0] for p in procs]
    assert [p.returncode for p in procs] == [0, 0], outs[0][-2000:] + outs[1][-2000:]


def _losses(path):
    vals = []
    for line in open(path).read().splitlines():
        m = re.search(r"loss for batch (\d+): (?:tensor\()?([-0-9.eE]+)", line)
        vals.append((int(m.group(1)), float(m.group(2))))
    return vals


# one experiment is enough here: test_distributed_cpu.py shows that all of this framework's algorithms
# give bit-identical trajectories among themselves
@pytest.mark.parametrize("experiment,name", [("main_onestep_reduce", "onestep_reduce")])
def test_same_loss_trajectory_as_the_reference(tmp_path, free_port, experiment, name):
    from distributed_learning_b200.utils.data import write_fake_mnist

    data = write_fake_mnist(str(tmp_path / "data"), n=1024)
    common = [data, "0", "--experiment", experiment, "--limit_batches", "5", "--batch_size", "16", "--job_id", "p"]
    ref_cwd = tmp_path / "ref"
    ref_cwd.mkdir()
    _run_pair(REF_MAIN, common, str(ref_cwd))                      # the reference always uses port 29501
    ours_cwd = tmp_path / "ours"
    ours_cwd.mkdir()
    _run_pair(os.path.join(ROOT, "main.py"), common + ["--master_port", str(free_port())], str(ours_cwd),
              env=dict(os.environ, DLB200_VERBOSE="1"))
    for rank in range(2):
        ref = _losses(ref_cwd / "results" / f"{experiment}_2_p" / f"{name}_{rank}_0_loss.txt")
        ours = _losses(ours_cwd / "results" / f"{experiment}_2_p" / f"{name}_{rank}_0_loss.txt")
        assert len(ref) == len(ours) == 5
        for (kb, vb), (ko, vo) in zip(ref, ours):
            assert kb == ko
            assert vo == pytest.approx(vb, abs=6e-5), (rank, kb, vb, vo)     # the reference prints 4 decimals
