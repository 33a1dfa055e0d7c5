"""Bindings of the hand-written sm_100a kernels (``csrc/``, C ABI in ``include/dlb200.h``).

``fused``: K-pack / K-unpack / K-sgd.  ``comm``: the peer-mapped arena and the in-kernel
allreduce (one-shot, two-shot, NVLS) with its device-side barrier.  ``build``: compiles
``libdlb200.so`` in-tree.  Nothing here falls back to PyTorch ops: if the library is missing
the import of a kernel raises.
"""
from ._lib import DlbError, available, lib, library_path  # noqa: F401
