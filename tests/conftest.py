import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")
    config.addinivalue_line("markers", "multigpu: needs at least two CUDA devices")


def pytest_collection_modifyitems(config, items):
    import torch

    have = torch.cuda.is_available()
    ngpu = torch.cuda.device_count() if have else 0
    for item in items:
        if "gpu" in item.keywords and not have:
            item.add_marker(pytest.mark.skip(reason="no CUDA device"))
        if "multigpu" in item.keywords and ngpu < 2:
            item.add_marker(pytest.mark.skip(reason="needs >= 2 GPUs"))


@pytest.fixture(scope="session", autouse=True)
def _deterministic_cudnn():
    """Several GPU tests train two copies of a model and compare them within a tolerance; cuDNN's
    default backward kernels use atomics and are not run-to-run reproducible."""
    import torch

    if torch.cuda.is_available():
        torch.backends.cudnn.deterministic = True
        torch.backends.cudnn.benchmark = False
    yield


@pytest.fixture(scope="session")
def free_port():
    import socket

    def _get():
        with socket.socket() as s:
            s.bind(("127.0.0.1", 0))
            return s.getsockname()[1]
    return _get
