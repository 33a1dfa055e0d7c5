"""Per-phase latency breakdown of a training step, written as the reference's CSV.

Parity: reference ``src/timing.py:1-41``.  The CSV layout is kept byte for byte
(``experiment_name, <phase>, ...`` with ``", "`` separators, one row per
``end_timing_experiment`` call, column order = first-seen order of the phases followed by
the extra fields), so existing plotting scripts keep working.

What changed: the reference reads ``time.time()`` on the host without a device
synchronisation, so on a GPU its columns are launch times.  Here a timer is
either a host timer (CPU runs) or a *device* timer: ``start``/``end`` record CUDA events on
the stream that is current at the call, no host synchronisation happens inside the step, and
the elapsed times are resolved once, in ``writeout`` / ``rows``, after a single
synchronise.  ``max_over_ranks`` folds the per-rank rows into the max over the ranks of a
process group, which is how every multi-GPU number of this project is reported.
"""
from __future__ import annotations

import time
from typing import Dict, List, Optional, Tuple

import torch

PHASES = ("get_data", "data2dev", "zero_grad", "forward", "backprop", "sync", "optimizer_step", "batch")


class _HostSpan:
    __slots__ = ("t0", "t1")

    def __init__(self):
        self.t0 = time.perf_counter()
        self.t1 = None

    def stop(self):
        self.t1 = time.perf_counter()

    def ms(self) -> float:
        return (self.t1 - self.t0) * 1000.0


class _DeviceSpan:
    __slots__ = ("e0", "e1")

    def __init__(self):
        self.e0 = torch.cuda.Event(enable_timing=True)
        self.e1 = torch.cuda.Event(enable_timing=True)
        self.e0.record()

    def stop(self):
        self.e1.record()

    def ms(self) -> float:
        return self.e0.elapsed_time(self.e1)


class PhaseTimer:
    """Collects named spans per step and emits the reference CSV.

    ``mode``: ``"host"`` (``perf_counter``), ``"device"`` (CUDA events) or ``"auto"``
    (device when ``device`` is a CUDA device).
    """

    def __init__(self, device=None, mode: str = "auto", nvtx: bool = False):
        dev = torch.device(device) if device is not None else torch.device("cpu")
        # NVTX: every phase is also a named range on the calling thread, so a profiler timeline
        # (ncu --nvtx, nsys) shows get_data / forward / backprop / sync / optimizer_step / batch
        self.nvtx = bool(nvtx) and dev.type == "cuda"
        if mode == "auto":
            mode = "device" if dev.type == "cuda" else "host"
        if mode not in ("host", "device"):
            raise ValueError(f"unknown timer mode {mode!r}")
        self.mode = mode
        self.device = dev
        self._span_cls = _DeviceSpan if mode == "device" else _HostSpan
        self._open: Dict[str, object] = {}
        self._step: Dict[str, list] = {}
        # rows: (experiment_name, {phase: [spans]}, extra_fields)
        self._rows: List[Tuple[str, Dict[str, list], Dict[str, object]]] = []

    # -- recording ---------------------------------------------------------------------
    def start(self, name: str) -> None:
        if self.nvtx:
            torch.cuda.nvtx.range_push(name)
        self._open[name] = self._span_cls()

    def end(self, name: str) -> None:
        span = self._open.pop(name, None)
        if span is None:
            raise KeyError(f"end_timer('{name}') without start_timer('{name}')")
        span.stop()
        if self.nvtx:
            torch.cuda.nvtx.range_pop()
        self._step.setdefault(name, []).append(span)

    def end_step(self, experiment_name: str, extra_fields: Optional[dict] = None) -> None:
        self._rows.append((experiment_name, self._step, dict(extra_fields or {})))
        self._step = {}
        self._open = {}

    # -- resolving ---------------------------------------------------------------------
    def rows(self) -> List[Tuple[str, Dict[str, object]]]:
        """Resolved rows: ``(experiment_name, {phase: mean ms, **extra_fields})``."""
        if self.mode == "device" and self._rows:
            torch.cuda.synchronize(self.device)
        out = []
        for name, spans, extra in self._rows:
            rec: Dict[str, object] = {k: sum(s.ms() for s in v) / len(v) for k, v in spans.items()}
            rec.update(extra)
            out.append((name, rec))
        return out

    def phase_means(self, skip: int = 0) -> Dict[str, float]:
        """Mean of every phase over the rows ``skip:`` (the reference's plots skip warm-up rows)."""
        rows = self.rows()[skip:]
        acc: Dict[str, List[float]] = {}
        for _, rec in rows:
            for k, v in rec.items():
                if isinstance(v, float):
                    acc.setdefault(k, []).append(v)
        return {k: sum(v) / len(v) for k, v in acc.items()}

    def max_over_ranks(self, group=None) -> List[Tuple[str, Dict[str, object]]]:
        """Rows with every float column replaced by its max over the ranks of ``group``."""
        import torch.distributed as dist

        rows = self.rows()
        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1 or not rows:
            return rows
        keys = [k for k, v in rows[0][1].items() if isinstance(v, float)]
        flat = torch.tensor([[rec.get(k, 0.0) for k in keys] for _, rec in rows], dtype=torch.float64)
        backend = dist.get_backend(group)
        if backend == "nccl":
            flat = flat.to(self.device)
        dist.all_reduce(flat, op=dist.ReduceOp.MAX, group=group)
        flat = flat.cpu()
        out = []
        for i, (name, rec) in enumerate(rows):
            rec = dict(rec)
            for j, k in enumerate(keys):
                rec[k] = float(flat[i, j])
            out.append((name, rec))
        return out

    # -- output ------------------------------------------------------------------------
    @staticmethod
    def format_csv(rows: List[Tuple[str, Dict[str, object]]]) -> str:
        if not rows:
            return ""
        keys = list(rows[0][1].keys())
        lines = ["experiment_name, " + ", ".join(keys)]
        for name, rec in rows:
            lines.append(name + ", " + ", ".join(str(rec[k]) for k in keys))
        return "\n".join(lines) + "\n"

    def writeout(self, filename: str, rows=None) -> None:
        rows = self.rows() if rows is None else rows
        self._rows = []
        if not rows:
            return
        with open(filename, "w") as f:
            f.write(self.format_csv(rows))

    def clear(self) -> None:
        self._rows, self._step, self._open = [], {}, {}


# ---- module-level API with the reference's function names (src/timing.py:8-40) -----------
_default = PhaseTimer(mode="host")


def set_default_timer(timer: PhaseTimer) -> PhaseTimer:
    global _default
    _default = timer
    return timer


def get_default_timer() -> PhaseTimer:
    return _default


def start_timer(name: str) -> None:
    _default.start(name)


def end_timer(name: str) -> None:
    _default.end(name)


def end_timing_experiment(experiment_name: str, extra_fields: Optional[dict] = None) -> None:
    _default.end_step(experiment_name, extra_fields)


def writeout_timer(filename: str) -> None:
    _default.writeout(filename)
