"""distributed_learning_b200 -- Blackwell-native decentralised data-parallel training.

A from-scratch framework with the capabilities of zdule/distributed_learning (a PyTorch
re-implementation of Horovod-style gradient averaging): the same ``main.py`` command line,
experiments, wrappers, bucket rule and result files, with the gradient path (pack / scale /
cast, allreduce over NVLink 5 / NVSwitch peer memory, unpack, SGD-momentum) as hand-written
sm_100a CUDA kernels.  See DESIGN.md for the architecture and SURVEY.md for the parity map.

Layout: ``models/`` networks, ``ops/`` kernel bindings (``csrc/`` holds the CUDA/C++ sources,
``include/dlb200.h`` the C ABI), ``parallel/`` buckets, wrappers, reducers, allreduce
algorithms, topology and the native communicator, ``utils/`` CLI, timing, data, checkpoint.
"""
__version__ = "0.1.0"

from .utils.logging import Level, print_d, eval_arg  # noqa: F401
