"""Python bindings of the single-GPU kernels: K-pack, K-unpack, K-sgd.

``PackPlan`` keeps the ctypes tables of one bucket (dtypes, offsets, element counts) so a step
only refreshes the tensor pointers -- the gradients autograd hands out live at new addresses
every iteration -- and makes one call.  See ``csrc/pack.cu`` / ``csrc/sgd.cu`` for the kernels and
``include/dlb200.h`` for the ABI.
"""
from __future__ import annotations

import ctypes
from ctypes import c_int, c_int64, c_void_p
from typing import Optional, Sequence

import torch

from ._lib import FusedSgd, SgdHyper, check, dtype_code, lib, stream_handle


PACK_VARIANTS = {"auto": 0, "chunks": 1, "tiles": 2, "bulk": 3}


class PackPlan:
    """Tables for packing/unpacking one list of tensors into/from one flat buffer."""

    def __init__(self, numels: Sequence[int], offsets: Sequence[int], dtypes: Sequence[torch.dtype]):
        n = len(numels)
        self.n = n
        self.numels = (c_int64 * n)(*[int(x) for x in numels])
        self.offsets = (c_int64 * n)(*[int(x) for x in offsets])
        self.dtypes = (c_int * n)(*[dtype_code(d) for d in dtypes])
        self.ptrs = (c_void_p * n)()

    @classmethod
    def for_tensors(cls, tensors: Sequence[torch.Tensor], offsets: Sequence[int]) -> "PackPlan":
        return cls([t.numel() for t in tensors], offsets, [t.dtype for t in tensors])

    def _set_ptrs(self, tensors: Sequence[Optional[torch.Tensor]]) -> None:
        ptrs = self.ptrs
        for i, t in enumerate(tensors):
            if t is None:
                ptrs[i] = None
            else:
                if not t.is_contiguous():
                    raise ValueError("pack/unpack need contiguous tensors")
                ptrs[i] = t.data_ptr()

    def bind(self, tensors: Sequence[Optional[torch.Tensor]]) -> "PackPlan":
        """Point the table at this step's tensors (for the allreduce kernels' fused pack prologue)."""
        self._set_ptrs(tensors)
        return self

    def pack(self, flat: torch.Tensor, tensors: Sequence[Optional[torch.Tensor]], scale: float = 1.0, stream=None,
             variant: str = "auto") -> None:
        """``flat[offset_i : offset_i+n_i] = cast(tensor_i * scale)``; a ``None`` tensor packs zeros.
        ``variant`` names the kernel (``PACK_VARIANTS``; see ``csrc/pack.cu``)."""
        self._set_ptrs(tensors)
        check(lib().dlb_pack_ex(flat.data_ptr(), dtype_code(flat.dtype), self.ptrs, self.dtypes, self.offsets,
                                self.numels, self.n, float(scale), PACK_VARIANTS[variant], stream_handle(stream)))

    def unpack(self, flat: torch.Tensor, tensors: Sequence[Optional[torch.Tensor]], scale: float = 1.0, stream=None,
               variant: str = "auto") -> None:
        """``tensor_i = cast(flat[offset_i : offset_i+n_i] * scale)``; ``None`` tensors are skipped."""
        self._set_ptrs(tensors)
        check(lib().dlb_unpack_ex(flat.data_ptr(), dtype_code(flat.dtype), self.ptrs, self.dtypes, self.offsets,
                                  self.numels, self.n, float(scale), PACK_VARIANTS[variant], stream_handle(stream)))


def pack(flat: torch.Tensor, tensors: Sequence[Optional[torch.Tensor]], offsets: Sequence[int], scale: float = 1.0,
         dtypes: Optional[Sequence[torch.dtype]] = None, numels: Optional[Sequence[int]] = None, stream=None,
         variant: str = "auto") -> None:
    """One-off pack (builds the tables on the fly)."""
    numels = numels or [t.numel() for t in tensors]
    dtypes = dtypes or [t.dtype for t in tensors]
    PackPlan(numels, offsets, dtypes).pack(flat, tensors, scale, stream, variant)


def unpack(flat: torch.Tensor, tensors: Sequence[torch.Tensor], offsets: Sequence[int], scale: float = 1.0, stream=None,
           variant: str = "auto") -> None:
    PackPlan.for_tensors(tensors, offsets).unpack(flat, tensors, scale, stream, variant)


def make_hyper(lr: float, momentum: float = 0.0, dampening: float = 0.0, weight_decay: float = 0.0,
               nesterov: bool = False, grad_scale: float = 1.0) -> SgdHyper:
    return SgdHyper(float(lr), float(momentum), float(dampening), float(weight_decay), float(grad_scale), int(nesterov))


def sgd_flat(param: torch.Tensor, momentum_buf: torch.Tensor, grad: torch.Tensor, hyper: SgdHyper,
             shadow_bf16: Optional[torch.Tensor] = None, stream=None) -> None:
    """In-place SGD-momentum on flat fp32 ``param``/``momentum_buf`` from a flat fp32/bf16 ``grad``."""
    n = param.numel()
    if param.dtype != torch.float32 or momentum_buf.dtype != torch.float32:
        raise TypeError("master parameters and momentum must be fp32")
    if grad.numel() < n or momentum_buf.numel() < n or (shadow_bf16 is not None and shadow_bf16.numel() < n):
        raise ValueError("sgd_flat: buffer shorter than the parameters")
    check(lib().dlb_sgd_flat(param.data_ptr(), momentum_buf.data_ptr(), grad.data_ptr(), dtype_code(grad.dtype),
                             shadow_bf16.data_ptr() if shadow_bf16 is not None else None, n,
                             ctypes.byref(hyper), stream_handle(stream)))


def sgd_multi(params: Sequence[torch.Tensor], momentum_bufs: Sequence[torch.Tensor],
              grads: Sequence[Optional[torch.Tensor]], hyper: SgdHyper, stream=None) -> None:
    """SGD-momentum over a list of (not flattened) fp32 parameters; ``None`` gradients are skipped."""
    n = len(params)
    P = (c_void_p * n)(*[p.data_ptr() for p in params])
    M = (c_void_p * n)(*[m.data_ptr() for m in momentum_bufs])
    G = (c_void_p * n)(*[g.data_ptr() if g is not None else None for g in grads])
    D = (c_int * n)(*[dtype_code(g.dtype) if g is not None else 0 for g in grads])
    N = (c_int64 * n)(*[p.numel() for p in params])
    check(lib().dlb_sgd_multi(P, M, G, D, N, n, ctypes.byref(hyper), stream_handle(stream)))


def fused_sgd_args(param: torch.Tensor, momentum_buf: torch.Tensor, hyper: SgdHyper,
                   shadow_bf16: Optional[torch.Tensor] = None) -> FusedSgd:
    """Descriptor of a fused-SGD epilogue for the allreduce kernels."""
    return FusedSgd(param.data_ptr(), momentum_buf.data_ptr(),
                    shadow_bf16.data_ptr() if shadow_bf16 is not None else None, hyper)
