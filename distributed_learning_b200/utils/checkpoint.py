"""Checkpoint / resume in torch ``state_dict`` format.

The reference has no checkpointing at all (no ``state_dict``, ``torch.save`` or
``load`` anywhere in ``src/``); the north star requires one.  A checkpoint is one file written
by worker 0 of node 0 (all replicas hold identical state after every step):

``{"format": 1, "model": model.state_dict(), "optimizer": optimizer.state_dict(),
   "batch_count": int, "epoch": int, "config": {...plain values...}, "rng": {...},
   "master_fp32": {...}}``  (the last entry only carries data for bf16 models on the native path:
   their fp32 master weights)

The optimizer entry has ``torch.optim.SGD``'s layout (``param_groups`` +
``state[i]["momentum_buffer"]``) whichever optimizer wrote it, so a run can be resumed under a
different backend (``FusedSGD`` on the native path, ``torch.optim.SGD`` elsewhere).  Files are
written to a temporary name and renamed, so a crash mid-write never leaves a torn checkpoint.
"""
from __future__ import annotations

import os
from typing import Any, Dict, Optional

import torch

FORMAT = 1


def _unwrap(model):
    # wrappers of this package keep the bare network in .model
    m = getattr(model, "__dict__", {}).get("model", None)
    return m if m is not None else model


def save_checkpoint(path: str, model, optimizer, batch_count: int, epoch: int = 0,
                    config: Optional[Any] = None, extra: Optional[Dict[str, Any]] = None) -> str:
    net = _unwrap(model)
    cfg = {}
    if config is not None:
        for k, v in vars(config).items():
            if isinstance(v, (int, float, str, bool, type(None))) or (
                    isinstance(v, (list, tuple)) and all(isinstance(x, (int, float, str)) for x in v)):
                cfg[k] = v
    payload = {
        "format": FORMAT,
        "model": {k: v.detach().to("cpu", copy=True) for k, v in net.state_dict().items()},
        "optimizer": _to_cpu(optimizer.state_dict()) if optimizer is not None else None,
        "batch_count": int(batch_count),
        "epoch": int(epoch),
        "config": cfg,
        "rng": {"torch": torch.get_rng_state()},
    }
    masters = getattr(model, "master_state_dict", None)
    if callable(masters):
        # bf16 models on the native path: the fp32 master weights, so a resume loses no precision
        payload["master_fp32"] = _to_cpu(masters())
    if extra:
        payload["extra"] = extra
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    tmp = f"{path}.tmp.{os.getpid()}"
    torch.save(payload, tmp)
    os.replace(tmp, path)
    return path


def load_checkpoint(path: str, model=None, optimizer=None, map_location="cpu", restore_rng: bool = False) -> Dict[str, Any]:
    """Load ``path``; restore ``model`` / ``optimizer`` in place when given.  Returns the payload."""
    payload = torch.load(path, map_location=map_location, weights_only=True)   # tensors and plain containers only
    if payload.get("format") != FORMAT:
        raise ValueError(f"{path}: unknown checkpoint format {payload.get('format')!r}")
    if model is not None:
        net = _unwrap(model)
        # copy in place: on the native path parameters are views into flat buffers
        own = net.state_dict()
        missing = [k for k in own if k not in payload["model"]]
        if missing:
            raise KeyError(f"{path}: checkpoint lacks {missing[:3]}{'...' if len(missing) > 3 else ''}")
        with torch.no_grad():
            for k, dst in own.items():
                dst.copy_(payload["model"][k])
        refresh = getattr(model, "parameters_were_reloaded", None)
        if callable(refresh):
            refresh()
        restore = getattr(model, "load_master_state_dict", None)
        if callable(restore) and payload.get("master_fp32"):
            restore(payload["master_fp32"])
    if optimizer is not None and payload.get("optimizer") is not None:
        optimizer.load_state_dict(payload["optimizer"])
    if restore_rng and "rng" in payload:
        torch.set_rng_state(payload["rng"]["torch"])
    return payload


def _to_cpu(obj):
    if torch.is_tensor(obj):
        return obj.detach().to("cpu", copy=True)
    if isinstance(obj, dict):
        return {k: _to_cpu(v) for k, v in obj.items()}
    if isinstance(obj, (list, tuple)):
        return type(obj)(_to_cpu(v) for v in obj)
    return obj
