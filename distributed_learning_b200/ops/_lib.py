"""ctypes loader for ``libdlb200.so`` (C ABI in ``include/dlb200.h``).

The library is built in-tree by ``ops/build.py``.  ``lib()`` loads it once; when it is missing
or older than the sources under ``csrc/`` and ``include/`` it is (re)built first.  There is no silent fallback: on a machine with a GPU a
missing library is an error, because a training run that quietly used PyTorch ops instead of the
kernels would measure the wrong thing.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import POINTER, c_char_p, c_float, c_int, c_int64, c_size_t, c_uint64, c_void_p
from typing import Optional

import torch

from . import build as _build

DLB_MAX_RANKS = 8
DLB_MAX_BLOCKS = 64
DLB_MAX_TENSORS = 512
DLB_SIGNAL_BYTES = 2 * DLB_MAX_BLOCKS * DLB_MAX_RANKS * 4
DLB_STATE_WORDS = DLB_MAX_BLOCKS + 8

DLB_F32, DLB_BF16, DLB_F16 = 0, 1, 2
DLB_E_ARG, DLB_E_UNSUPPORTED, DLB_E_SYS, DLB_E_TIMEOUT, DLB_E_LIMIT = -1, -2, -3, -4, -5

DTYPE_CODE = {torch.float32: DLB_F32, torch.bfloat16: DLB_BF16, torch.float16: DLB_F16}


class SgdHyper(ctypes.Structure):
    _fields_ = [("lr", c_float), ("momentum", c_float), ("dampening", c_float), ("weight_decay", c_float),
                ("grad_scale", c_float), ("nesterov", c_int)]


class FusedSgd(ctypes.Structure):
    _fields_ = [("param", c_void_p), ("momentum_buf", c_void_p), ("shadow_bf16", c_void_p), ("hyper", SgdHyper)]


class AllreduceDesc(ctypes.Structure):
    """``dlb_allreduce_desc_t`` (include/dlb200.h)."""
    _fields_ = [("algo", c_int), ("rank", c_int), ("nranks", c_int),
                ("bufs", POINTER(c_void_p)), ("out", c_void_p), ("mc_buf", c_void_p), ("sigs", POINTER(c_void_p)),
                ("state", c_void_p), ("numel", c_int64), ("dtype", c_int), ("scale", c_float),
                ("nblocks", c_int), ("nthreads", c_int), ("unroll", c_int), ("end_barrier", c_int),
                ("sgd", POINTER(FusedSgd)),
                ("tensors", POINTER(c_void_p)), ("tensor_dtypes", POINTER(c_int)), ("offsets", POINTER(c_int64)),
                ("numels", POINTER(c_int64)), ("ntensors", c_int), ("pack_scale", c_float), ("trace", c_void_p)]


class EmulatedDesc(ctypes.Structure):
    """``dlb_emulated_desc_t`` (include/dlb200.h)."""
    _fields_ = [("algo", c_int), ("nranks", c_int),
                ("bufs", POINTER(c_void_p)), ("outs", POINTER(c_void_p)), ("sgds", POINTER(FusedSgd)), ("flags", c_void_p),
                ("numel", c_int64), ("dtype", c_int), ("scale", c_float), ("nblocks", c_int), ("nthreads", c_int),
                ("tensors", POINTER(c_void_p)), ("tensor_dtypes", POINTER(c_int)), ("offsets", POINTER(c_int64)),
                ("numels", POINTER(c_int64)), ("ntensors", c_int), ("pack_scale", c_float), ("trace", c_void_p)]


ALGO_CODE = {"twoshot": 0, "oneshot": 1, "nvls": 2, "twoshot_bulk": 3}


class DlbError(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(f"libdlb200 error {code}: {message}")
        self.code = code


_LIB: Optional[ctypes.CDLL] = None

_PROTOS = {
    "dlb_last_error": (c_char_p, []),
    "dlb_version": (c_int, []),
    "dlb_topology_json": (c_int, [c_char_p, c_size_t]),
    "dlb_device_count": (c_int, [POINTER(c_int)]),
    "dlb_can_access_peer": (c_int, [c_int, c_int, POINTER(c_int)]),
    "dlb_enable_peer_access": (c_int, [c_int, c_int]),
    "dlb_mem_alloc": (c_int, [c_size_t, POINTER(c_void_p)]),
    "dlb_mem_free": (c_int, [c_void_p]),
    "dlb_ipc_export": (c_int, [c_void_p, c_char_p]),
    "dlb_ipc_open": (c_int, [c_char_p, POINTER(c_void_p)]),
    "dlb_ipc_close": (c_int, [c_void_p]),
    "dlb_vmm_granularity": (c_int, [c_int, c_int, POINTER(c_size_t)]),
    "dlb_vmm_alloc": (c_int, [c_int, c_size_t, POINTER(c_uint64), POINTER(c_void_p), POINTER(c_size_t)]),
    "dlb_vmm_export_fd": (c_int, [c_uint64, POINTER(c_int)]),
    "dlb_vmm_import_fd": (c_int, [c_int, POINTER(c_uint64)]),
    "dlb_vmm_map": (c_int, [c_int, c_uint64, c_size_t, POINTER(c_void_p)]),
    "dlb_vmm_unmap": (c_int, [c_void_p, c_size_t]),
    "dlb_vmm_release": (c_int, [c_uint64]),
    "dlb_fd_listen": (c_int, [c_char_p, POINTER(c_int)]),
    "dlb_fd_accept": (c_int, [c_int, POINTER(c_int), POINTER(c_int), c_int]),
    "dlb_fd_send": (c_int, [c_char_p, c_int, c_int, c_int]),
    "dlb_fd_close": (c_int, [c_int]),
    "dlb_mc_supported": (c_int, [c_int, POINTER(c_int)]),
    "dlb_mc_create": (c_int, [c_int, c_size_t, POINTER(c_uint64), POINTER(c_size_t)]),
    "dlb_mc_add_device": (c_int, [c_uint64, c_int]),
    "dlb_mc_bind": (c_int, [c_uint64, c_size_t, c_uint64, c_size_t]),
    "dlb_mc_map": (c_int, [c_int, c_uint64, c_size_t, POINTER(c_void_p)]),
    "dlb_pack": (c_int, [c_void_p, c_int, POINTER(c_void_p), POINTER(c_int), POINTER(c_int64), POINTER(c_int64),
                         c_int, c_float, c_void_p]),
    "dlb_unpack": (c_int, [c_void_p, c_int, POINTER(c_void_p), POINTER(c_int), POINTER(c_int64), POINTER(c_int64),
                           c_int, c_float, c_void_p]),
    "dlb_pack_ex": (c_int, [c_void_p, c_int, POINTER(c_void_p), POINTER(c_int), POINTER(c_int64), POINTER(c_int64),
                            c_int, c_float, c_int, c_void_p]),
    "dlb_unpack_ex": (c_int, [c_void_p, c_int, POINTER(c_void_p), POINTER(c_int), POINTER(c_int64), POINTER(c_int64),
                              c_int, c_float, c_int, c_void_p]),
    "dlb_sgd_flat": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_int64, POINTER(SgdHyper), c_void_p]),
    "dlb_sgd_multi": (c_int, [POINTER(c_void_p), POINTER(c_void_p), POINTER(c_void_p), POINTER(c_int),
                              POINTER(c_int64), c_int, POINTER(SgdHyper), c_void_p]),
    "dlb_allreduce_twoshot": (c_int, [POINTER(c_void_p), POINTER(c_void_p), c_void_p, c_int, c_int, c_int64, c_int,
                                      c_float, c_int, c_int, POINTER(FusedSgd), c_void_p]),
    "dlb_allreduce_oneshot": (c_int, [POINTER(c_void_p), c_void_p, POINTER(c_void_p), c_void_p, c_int, c_int, c_int64,
                                      c_int, c_float, c_int, c_int, c_int, POINTER(FusedSgd), c_void_p]),
    "dlb_allreduce_nvls": (c_int, [c_void_p, POINTER(c_void_p), POINTER(c_void_p), c_void_p, c_int, c_int, c_int64,
                                   c_int, c_float, c_int, c_int, POINTER(FusedSgd), c_void_p]),
    "dlb_allreduce_nvls_deep": (c_int, [c_void_p, POINTER(c_void_p), POINTER(c_void_p), c_void_p, c_int, c_int, c_int64,
                                        c_int, c_float, c_int, c_int, POINTER(FusedSgd), c_void_p]),
    "dlb_allreduce": (c_int, [POINTER(AllreduceDesc), c_void_p]),
    "dlb_allreduce_emulated_ex": (c_int, [POINTER(EmulatedDesc), c_void_p]),
    "dlb_broadcast": (c_int, [POINTER(c_void_p), POINTER(c_void_p), c_void_p, c_int, c_int, c_int, c_int64, c_int, c_int, c_void_p]),
    "dlb_barrier": (c_int, [POINTER(c_void_p), c_void_p, c_int, c_int, c_int, c_void_p]),
    "dlb_set_wait_timeout_ms": (c_int, [c_int]),
    "dlb_comm_check": (c_int, [c_void_p, c_void_p]),
    "dlb_allreduce_emulated": (c_int, [c_int, POINTER(c_void_p), POINTER(c_void_p), POINTER(FusedSgd), c_void_p, c_int,
                                       c_int64, c_int, c_float, c_int, c_int, c_void_p]),
    "dlb_emulated_flags_bytes": (c_size_t, [c_int]),
    "dlb_host_fill_uniform": (c_int, [c_void_p, c_int, c_int64, c_uint64, c_uint64, c_int]),
    "dlb_host_fill_randint": (c_int, [c_void_p, c_int64, c_int64, c_uint64, c_uint64]),
}


def library_path() -> str:
    return os.environ.get("DLB200_LIB", _build.LIB_PATH)


def lib() -> ctypes.CDLL:
    """The loaded library; builds it first if the in-tree copy is missing and nvcc exists."""
    global _LIB
    if _LIB is not None:
        return _LIB
    path = library_path()
    if path == _build.LIB_PATH and not _build.is_current():
        # missing, or older than csrc/ / include/: never load a stale library silently
        try:
            _build.build_library()
        except Exception as e:  # noqa: BLE001
            if not os.path.exists(path):
                raise RuntimeError(
                    f"{path} is missing and could not be built ({e}). Run `python -m distributed_learning_b200.ops.build`."
                ) from e
            import warnings

            warnings.warn(f"{path} is older than its sources and could not be rebuilt ({e}); loading it as it is")
    handle = ctypes.CDLL(path)
    for name, (restype, argtypes) in _PROTOS.items():
        try:
            fn = getattr(handle, name)
        except AttributeError:
            continue    # optional entry points added later are bound by their own modules
        fn.restype, fn.argtypes = restype, argtypes
    _LIB = handle
    return handle


def available() -> bool:
    try:
        lib()
        return True
    except (RuntimeError, OSError):
        return False


def check(rc: int) -> None:
    if rc != 0:
        raise DlbError(rc, lib().dlb_last_error().decode("utf-8", "replace"))


def stream_handle(stream: Optional["torch.cuda.Stream"] = None) -> int:
    return (stream or torch.cuda.current_stream()).cuda_stream


def dtype_code(dtype: torch.dtype) -> int:
    try:
        return DTYPE_CODE[dtype]
    except KeyError:
        raise TypeError(f"unsupported dtype {dtype} (fp32, bf16, fp16)") from None


def ptr_array(ptrs) -> "ctypes.Array":
    arr = (c_void_p * len(ptrs))()
    for i, p in enumerate(ptrs):
        arr[i] = p
    return arr
