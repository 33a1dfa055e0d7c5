"""The native communicator: a peer-mapped arena per GPU and the in-kernel allreduce over it.

One process per GPU.  ``NativeComm.allocate`` is collective: every rank allocates an arena of
the same size, exports it, and maps every peer's arena (VMM allocations shared as POSIX file
descriptors over a unix socket, with a multicast object bound on top when the box supports NVLS;
legacy CUDA IPC handles as the fallback).  The bootstrap (token, barriers, IPC handles) goes
through a ``torch.distributed`` group -- gloo is enough, NCCL is never needed.  Afterwards a
fusion buffer is an ``(offset, numel, dtype)`` triple that means the same bytes in every rank's
arena, and ``allreduce`` launches ONE kernel on the caller's stream (``csrc/allreduce.cu``).

This replaces reference ``init_process`` + ``pgroups`` (``src/main.py:303-318``) and the
``reduce_fn`` injected into the reducers (``src/main.py:131,141,210``).

``emulated_allreduce`` runs N "ranks" in one cooperative launch on one GPU; it exists so the
kernel logic (barrier, slicing, epilogue) is testable where only one GPU is available.
"""
from __future__ import annotations

import ctypes
import os
import uuid
from ctypes import c_int, c_size_t, c_uint64, c_void_p
from typing import Dict, List, Optional, Sequence, Tuple

import torch

from ._lib import (ALGO_CODE, DLB_MAX_BLOCKS, DLB_MAX_RANKS, DLB_SIGNAL_BYTES, DLB_STATE_WORDS, AllreduceDesc, DlbError,
                   EmulatedDesc, FusedSgd, check, dtype_code, lib, ptr_array, stream_handle)

ALGOS = ("oneshot", "twoshot", "nvls")         # what `tune` tries; "twoshot_bulk" is opt-in (--algo)
_ALIGN = 4096


def _round_up(x: int, m: int) -> int:
    return (x + m - 1) // m * m


class _RawCudaMemory:
    """Adapter that lets torch wrap a raw device pointer (CUDA array interface v2)."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


def tensor_from_ptr(ptr: int, nbytes: int, device) -> torch.Tensor:
    """A uint8 tensor over ``[ptr, ptr+nbytes)``; the caller keeps the memory alive."""
    return torch.as_tensor(_RawCudaMemory(ptr, nbytes), device=device)


# HBM bytes per second that ONE CTA (512 threads) of a communication kernel sustains in its pack
# prologue and SGD epilogue: ncu of the fused kernels on real peers, 16 MiB with 32 CTAs, shows
# 117 MB of pack + update traffic in ~69 us on top of the bare allreduce at 2 and at 8 GPUs
# (profiles/ncu/r2_nvlink_n{2,8}_durations.csv).
FUSED_HBM_BYTES_PER_S_PER_CTA = 53e9


def default_nblocks(nbytes: int, algo: str, hbm_bytes: int = 0) -> int:
    """CTAs when nothing was measured.  From the 2/4/8-GPU sweeps (profiles/results/r2_suite_n*): the
    switch-reduced variant is fastest with few CTAs (16) at every size, the peer-to-peer variants
    want more as the message grows; a kernel that also packs and updates (``hbm_bytes`` > 0) is bound
    by how many SMs stream its HBM traffic and takes the full grid from 1 MiB up."""
    env = os.environ.get("DLB200_NBLOCKS")
    if env:
        return max(1, min(DLB_MAX_BLOCKS, int(env)))
    if hbm_bytes > 0 and nbytes >= 1 << 20:
        return DLB_MAX_BLOCKS
    if algo == "nvls":
        return 16
    if nbytes <= 256 << 10:
        return 16
    if nbytes <= 4 << 20:
        return 32
    return 64


def default_unroll(nbytes: int) -> int:
    """nvls pipeline depth when nothing was measured: 4 reductions in flight per thread won or tied
    at every size of the 4- and 8-GPU sweeps (8 is 2-30 % slower from 16 MiB up)."""
    env = os.environ.get("DLB200_UNROLL")
    if env:
        return int(env)
    return 4


class _ModeFailed(RuntimeError):
    """An arena mode (vmm / ipc) failed on at least one rank; raised on every rank at the same vote."""


class NativeComm:
    def __init__(self, group=None, device=None, prefer: str = "auto", nthreads: int = 512):
        import torch.distributed as dist

        self._dist = dist
        self.group = group
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        if self.world > DLB_MAX_RANKS:
            raise ValueError(f"the native backend spans one NVSwitch domain: at most {DLB_MAX_RANKS} ranks, got {self.world}")
        self.device = torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")
        self.prefer = os.environ.get("DLB200_ARENA", prefer)          # auto | vmm | ipc
        self.nthreads = nthreads
        self.mode: Optional[str] = None
        self.nbytes = 0
        self.data_offset = _round_up(DLB_SIGNAL_BYTES, _ALIGN)
        self.peer_ptrs: List[int] = []
        self.mc_ptr: Optional[int] = None
        self.mc_note = ""          # why multicast is unavailable, for the log
        self.local: Optional[torch.Tensor] = None
        self.state = torch.zeros(DLB_STATE_WORDS, dtype=torch.int32, device=self.device)
        self._sig_array = None
        self._buf_arrays: Dict[int, "ctypes.Array"] = {}
        self._handles: dict = {}
        self.tuned: Dict[Tuple[int, int], Tuple[str, int, int]] = {}     # (dtype code, pow2 bytes) -> (algo, nblocks, unroll)
        self.tuned_ms: Dict[Tuple[int, int], Dict[Tuple[str, int, int], float]] = {}   # ... -> every candidate's time
        self.launches = 0
        self.topology: Optional[dict] = None

    # ---------------------------------------------------------------------------------------
    # bring-up
    # ---------------------------------------------------------------------------------------
    def _gather(self, obj):
        out = [None] * self.world
        self._dist.all_gather_object(out, obj, group=self.group)
        return out

    def _barrier(self):
        if self.world > 1:
            self._dist.barrier(group=self.group)

    def _agree(self, ok: bool, what: str, exc: Optional[BaseException] = None) -> None:
        """Collective vote after a local step.  Either every rank goes on, or every rank raises
        ``_ModeFailed`` here -- nobody is left alone inside the next collective phase."""
        oks = self._gather(bool(ok)) if self.world > 1 else [bool(ok)]
        if not all(oks):
            bad = [i for i, o in enumerate(oks) if not o]
            raise _ModeFailed(f"{what} failed on rank(s) {bad}" + (f" (here: {exc})" if exc is not None else ""))

    def _check_topology(self) -> None:
        """Topology gate: the kernels address every peer directly, so every pair of the job's GPUs
        must be able to map each other's memory.  A GPU with fewer active NVLink links than its
        peers still works, at that GPU's bandwidth: warn."""
        from ..parallel.topology import all_to_all_capable, query_topology
        from ..utils.logging import Level, print_d

        try:
            topo = query_topology()
        except Exception as e:  # noqa: BLE001 - discovery is advisory when the query itself fails
            print_d(f"topology query failed ({e}); continuing without the peer-access gate", Level.WARNING)
            topo = None
        devs = self._gather(self.device.index) if self.world > 1 else [self.device.index]
        self.topology = topo
        if topo is None or self.world == 1:
            return
        capable = all_to_all_capable(topo, devs)
        if not all(self._gather(capable)):
            rows = "\n".join("  " + " ".join(str(int(x)) for x in row) for row in topo.get("peer_access", []))
            raise RuntimeError(f"the native backend needs peer access between all of GPUs {devs}; "
                               f"cudaDeviceCanAccessPeer matrix:\n{rows}")
        links = {d["index"]: d.get("nvlinks_active", -1) for d in topo.get("devices", [])}
        mine = [links.get(i, -1) for i in devs]
        known = [x for x in mine if x >= 0]
        if known and min(known) < max(known) and self.rank == 0:
            print_d(f"degraded NVLink: active links per GPU {dict(zip(devs, mine))}; collectives run at the "
                    "bandwidth of the slowest GPU", Level.WARNING)
        ranks_perf = topo.get("performance_rank")
        if ranks_perf and self.rank == 0:
            worst = max(ranks_perf[a][b] for a in devs for b in devs if a != b)
            if worst > 0:
                print_d(f"some GPU pairs are not on the best link class (performance rank up to {worst})", Level.WARNING)

    def allocate(self, data_bytes: int) -> None:
        """Collective.  Allocates signal pad + ``data_bytes`` and maps all peers."""
        if self.mode is not None:
            raise RuntimeError("arena already allocated")
        total = self.data_offset + _round_up(max(data_bytes, 16), _ALIGN)
        torch.cuda.set_device(self.device)
        self._check_topology()
        modes = {"auto": ("vmm", "ipc"), "vmm": ("vmm",), "ipc": ("ipc",)}[self.prefer]
        err: Optional[Exception] = None
        for mode in modes:
            try:
                if mode == "vmm":
                    self._allocate_vmm(total)
                else:
                    self._allocate_ipc(total)
                self.mode = mode
                break
            except _ModeFailed as e:         # raised on every rank at the same vote
                err = e
                self._release_partial()
        if self.mode is None:
            raise RuntimeError(f"could not set up peer-mapped memory ({err})") from err
        self.nbytes = total
        self.local = tensor_from_ptr(self.peer_ptrs[self.rank], total, self.device)
        self._sig_array = ptr_array(self.peer_ptrs)          # the signal pad sits at offset 0
        self._barrier()

    def _allocate_ipc(self, total: int) -> None:
        L = lib()
        p = c_void_p()
        handle = ctypes.create_string_buffer(64)
        self._handles = {"ipc_peers": []}
        ok, exc = True, None
        try:
            check(L.dlb_mem_alloc(total, ctypes.byref(p)))
            self._handles["ipc_local"] = p.value
            check(L.dlb_ipc_export(p, handle))
        except (DlbError, OSError) as e:
            ok, exc = False, e
        self._agree(ok, "ipc: allocate + export", exc)
        handles = self._gather(handle.raw) if self.world > 1 else [handle.raw]
        ptrs = []
        try:
            for r, h in enumerate(handles):
                if r == self.rank:
                    ptrs.append(p.value)
                else:
                    q = c_void_p()
                    check(L.dlb_ipc_open(ctypes.create_string_buffer(h, 64), ctypes.byref(q)))
                    self._handles["ipc_peers"].append(q.value)
                    ptrs.append(q.value)
        except (DlbError, OSError) as e:
            ok, exc = False, e
        self._agree(ok, "ipc: open peer handles", exc)
        self.peer_ptrs = ptrs

    def _allocate_vmm(self, total: int) -> None:
        L = lib()
        dev = self.device.index
        h, p, rounded = c_uint64(), c_void_p(), c_size_t()
        self._handles = {"vmm_peers": []}
        ptrs = [None] * self.world
        lfd = c_int(-1)
        # the token is exchanged before anything can fail locally, so the vote below is the first
        # collective a failing rank can miss -- and it does not miss it
        token = self._gather(uuid.uuid4().hex if self.rank == 0 else None)[0] if self.world > 1 else ""
        name = lambda r: f"dlb200-{token}-{r}".encode()     # noqa: E731
        ok, exc = True, None
        try:
            check(L.dlb_vmm_alloc(dev, total, ctypes.byref(h), ctypes.byref(p), ctypes.byref(rounded)))
            self._handles["vmm_local"] = (h.value, p.value, rounded.value)
            ptrs[self.rank] = p.value
            if self.world > 1:
                check(L.dlb_fd_listen(name(self.rank), ctypes.byref(lfd)))
        except (DlbError, OSError) as e:
            ok, exc = False, e
        try:
            self._agree(ok, "vmm: allocate + listen", exc)       # doubles as "everyone is listening"
            if self.world > 1:
                try:
                    fd = c_int()
                    check(L.dlb_vmm_export_fd(h.value, ctypes.byref(fd)))
                    try:
                        for r in range(self.world):
                            if r != self.rank:
                                check(L.dlb_fd_send(name(r), fd.value, self.rank, 30000))
                    finally:
                        L.dlb_fd_close(fd.value)
                    for _ in range(self.world - 1):
                        pfd, tag = c_int(), c_int()
                        check(L.dlb_fd_accept(lfd.value, ctypes.byref(pfd), ctypes.byref(tag), 30000))
                        ph, pp = c_uint64(), c_void_p()
                        try:
                            check(L.dlb_vmm_import_fd(pfd.value, ctypes.byref(ph)))
                        finally:
                            L.dlb_fd_close(pfd.value)
                        self._handles["vmm_peers"].append((ph.value, None, rounded.value))
                        check(L.dlb_vmm_map(dev, ph.value, rounded.value, ctypes.byref(pp)))
                        self._handles["vmm_peers"][-1] = (ph.value, pp.value, rounded.value)
                        ptrs[tag.value] = pp.value
                except (DlbError, OSError) as e:
                    ok, exc = False, e
                self._agree(ok, "vmm: exchange + map peer handles", exc)
                self._setup_multicast(L, dev, lfd.value, name, h.value, rounded.value)
        finally:
            if lfd.value >= 0:
                L.dlb_fd_close(lfd.value)
        self.peer_ptrs = ptrs

    def _setup_multicast(self, L, dev: int, lfd: int, name, mem_handle: int, rounded: int) -> None:
        """NVLS: bind every rank's arena to one multicast object.  Optional: any failure leaves
        ``mc_ptr`` unset on all ranks and the nvls variant unavailable."""
        if os.environ.get("DLB200_NO_MULTICAST"):
            sup = False
        else:
            yes = c_int()
            rc = L.dlb_mc_supported(dev, ctypes.byref(yes))
            sup = rc == 0 and bool(yes.value)
            if not sup:
                self.mc_note = (f"device attribute MULTICAST_SUPPORTED is 0" if rc == 0
                                else f"multicast query failed: {L.dlb_last_error().decode()}")
        if not all(self._gather(sup)):
            return
        mc = c_uint64()
        step_ok = True
        try:
            if self.rank == 0:
                mrounded = c_size_t()
                check(L.dlb_mc_create(self.world, rounded, ctypes.byref(mc), ctypes.byref(mrounded)))
                if mrounded.value != rounded:
                    raise RuntimeError("multicast granularity differs from the arena's")
                fd = c_int()
                check(L.dlb_vmm_export_fd(mc.value, ctypes.byref(fd)))
                for r in range(1, self.world):
                    check(L.dlb_fd_send(name(r), fd.value, 1000, 30000))
                L.dlb_fd_close(fd.value)
            else:
                pfd, tag = c_int(), c_int()
                check(L.dlb_fd_accept(lfd, ctypes.byref(pfd), ctypes.byref(tag), 30000))
                check(L.dlb_vmm_import_fd(pfd.value, ctypes.byref(mc)))
                L.dlb_fd_close(pfd.value)
            check(L.dlb_mc_add_device(mc.value, dev))
        except (DlbError, RuntimeError) as e:
            step_ok = False
            self.mc_note = f"multicast create/add failed: {e}"
        if not all(self._gather(step_ok)):      # also the "all devices added" barrier
            return
        try:
            check(L.dlb_mc_bind(mc.value, 0, mem_handle, rounded))
            mp = c_void_p()
            check(L.dlb_mc_map(dev, mc.value, rounded, ctypes.byref(mp)))
        except DlbError as e:
            step_ok = False
            self.mc_note = f"multicast bind/map failed: {e}"
        if all(self._gather(step_ok)):
            self.mc_ptr = mp.value
            self._handles["mc"] = (mc.value, mp.value, rounded)

    def _release_partial(self) -> None:
        """Give back whatever a failed mode created (best effort: the mode is being abandoned)."""
        L = lib()
        h = self._handles
        if "mc" in h:
            mc, mp, size = h["mc"]
            L.dlb_vmm_unmap(mp, size)
            L.dlb_vmm_release(mc)
        for ph, pp, size in h.get("vmm_peers", []):
            if pp is not None:
                L.dlb_vmm_unmap(pp, size)
            L.dlb_vmm_release(ph)
        if "vmm_local" in h:
            lh, lp, size = h["vmm_local"]
            L.dlb_vmm_unmap(lp, size)
            L.dlb_vmm_release(lh)
        for q in h.get("ipc_peers", []):
            L.dlb_ipc_close(q)
        self._barrier()                       # every rank is here (same vote): peers have let go of our block
        if h.get("ipc_local"):
            L.dlb_mem_free(h["ipc_local"])
        self.peer_ptrs, self.mc_ptr, self._handles = [], None, {}

    @property
    def has_multicast(self) -> bool:
        return self.mc_ptr is not None

    # ---------------------------------------------------------------------------------------
    # buffers
    # ---------------------------------------------------------------------------------------
    def view(self, offset: int, numel: int, dtype: torch.dtype) -> torch.Tensor:
        """Local tensor over ``numel`` elements at data offset ``offset`` (bytes, 16-aligned)."""
        esize = torch.empty((), dtype=dtype).element_size()
        lo = self.data_offset + offset
        if offset % 16 or lo + numel * esize > self.nbytes:
            raise ValueError("arena view out of range or misaligned")
        return self.local[lo:lo + numel * esize].view(dtype)

    def _bufs(self, offset: int):
        arr = self._buf_arrays.get(offset)
        if arr is None:
            base = self.data_offset + offset
            arr = ptr_array([p + base for p in self.peer_ptrs])
            self._buf_arrays[offset] = arr
        return arr

    # ---------------------------------------------------------------------------------------
    # collectives
    # ---------------------------------------------------------------------------------------
    def choose(self, nbytes: int, dtype: torch.dtype, algo: str = "auto", hbm_bytes: int = 0) -> Tuple[str, int, int]:
        """``(variant, CTAs, unroll)`` for a message: from the measured table if ``tune`` filled it, else
        a size heuristic (one-shot while latency dominates, NVLS when the switch can reduce).

        ``hbm_bytes``: local HBM traffic of the prologue and epilogue fused into the kernel (pack:
        2 bytes moved per buffer byte; fp32 SGD: 5).  ``tune`` times the bare allreduce, where the
        switch-reduced variant is fastest on 8-16 CTAs; a kernel that also streams several times the
        buffer through HBM on those few SMs would be far slower than the allreduce it saves, so the
        fused choice minimises  measured allreduce time + hbm_bytes / (CTAs x per-CTA bandwidth)."""
        if algo != "auto":
            unroll = 8 if algo == "nvls_deep" else default_unroll(nbytes)
            if algo in ("nvls", "nvls_deep"):
                algo = "nvls" if self.has_multicast else "twoshot"
            return algo, default_nblocks(nbytes, algo, hbm_bytes), unroll
        key = (dtype_code(dtype), max(nbytes, 1).bit_length())
        table = self.tuned_ms.get(key)
        if table and hbm_bytes > 0:
            def cost(item):
                (_, nb, _), ms = item
                return ms + hbm_bytes / (nb * FUSED_HBM_BYTES_PER_S_PER_CTA) * 1e3
            return min(table.items(), key=cost)[0]
        if key in self.tuned:
            best = self.tuned[key]
            if hbm_bytes > 0:                    # only the winner is known (cached table): keep its variant
                return best[0], max(best[1], default_nblocks(nbytes, best[0], hbm_bytes)), best[2]
            return best
        if self.has_multicast and self.world >= 4:
            if nbytes <= 512 << 10:
                return "oneshot", default_nblocks(nbytes, "oneshot", hbm_bytes), 0
            return "nvls", default_nblocks(nbytes, "nvls", hbm_bytes), default_unroll(nbytes)
        if nbytes * (self.world - 1) <= 16 << 20:
            return "oneshot", default_nblocks(nbytes, "oneshot", hbm_bytes), 0
        return "twoshot", default_nblocks(nbytes, "twoshot", hbm_bytes), 0

    def allreduce(self, offset: int, numel: int, dtype: torch.dtype, algo: str = "twoshot", scale: float = 1.0,
                  out_offset: Optional[int] = None, sgd: Optional[FusedSgd] = None, nblocks: Optional[int] = None,
                  stream=None, end_barrier: bool = False, pack=None, pack_scale: float = 1.0, unroll: int = 0,
                  trace: Optional[torch.Tensor] = None) -> None:
        """Sum ``numel`` elements at ``offset`` over all ranks (times ``scale``) on ``stream``.

        two-shot / nvls work in place; one-shot writes to ``out_offset``.  By default one-shot has
        no trailing barrier and the caller must alternate input buffers between consecutive uses;
        ``end_barrier=True`` adds the barrier so the same input buffer can be refilled right away
        (needed when a step is replayed from a CUDA graph).  ``sgd`` fuses the parameter update.
        ``pack`` (a ``PackPlan`` whose pointers have been set, see ``PackPlan.bind``) fuses the
        gather of the bucket's gradient tensors into the kernel's prologue (times ``pack_scale``).
        ``unroll`` (2/4/8) is the nvls variant's pipeline depth; ``trace`` an int64 tensor of 8
        elements that receives the phase timestamps.  Every rank must make the same sequence of
        calls with the same ``nblocks`` and ``unroll``."""
        esize = torch.empty((), dtype=dtype).element_size()
        if algo == "nvls_deep":
            algo, unroll = "nvls", 8
        if algo not in ALGO_CODE:
            raise ValueError(f"unknown algo {algo!r}")
        d = AllreduceDesc()
        d.algo, d.rank, d.nranks = ALGO_CODE[algo], self.rank, self.world
        d.bufs, d.sigs, d.state = self._bufs(offset), self._sig_array, self.state.data_ptr()
        d.numel, d.dtype, d.scale = numel, dtype_code(dtype), scale
        d.nblocks = nblocks or default_nblocks(numel * esize, algo)
        d.nthreads, d.unroll, d.end_barrier = self.nthreads, unroll, int(end_barrier)
        if sgd is not None:
            d.sgd = ctypes.pointer(sgd)
        if algo == "oneshot":
            if out_offset is None and sgd is None:
                raise ValueError("one-shot needs out_offset (it cannot work in place)")
            d.out = self.peer_ptrs[self.rank] + self.data_offset + out_offset if out_offset is not None else None
        elif algo == "nvls":
            if not self.has_multicast:
                raise RuntimeError("nvls requested but no multicast mapping exists")
            d.mc_buf = self.mc_ptr + self.data_offset + offset
        if pack is not None:
            d.tensors, d.tensor_dtypes, d.offsets, d.numels, d.ntensors = pack.ptrs, pack.dtypes, pack.offsets, pack.numels, pack.n
            d.pack_scale = pack_scale
        if trace is not None:
            d.trace = trace.data_ptr()
        check(lib().dlb_allreduce(ctypes.byref(d), stream_handle(stream)))
        self.launches += 1

    def broadcast(self, offset: int, nbytes: int, root: int = 0, stream=None, nblocks: int = 16) -> None:
        """Every rank's arena bytes ``[offset, offset+nbytes)`` become a copy of ``root``'s (one kernel)."""
        check(lib().dlb_broadcast(self._bufs(offset), self._sig_array, self.state.data_ptr(), self.rank, self.world, root,
                                  nbytes, nblocks, self.nthreads, stream_handle(stream)))
        self.launches += 1

    def broadcast_tensors(self, tensors: Sequence[torch.Tensor], root: int = 0) -> int:
        """Collective.  Overwrite every tensor in ``tensors`` (same shapes and dtypes on all ranks) with
        ``root``'s values, staging through the arena's data area: local copy in, one ``broadcast``
        kernel, local copy out.  The arena's contents are destroyed.  Returns the bytes moved."""
        if self.world == 1 or not tensors:
            return 0
        cap = (self.nbytes - self.data_offset) // 16 * 16
        stage = self.local[self.data_offset:self.data_offset + cap]
        moved = 0
        pending: List[Tuple[torch.Tensor, int, int]] = []      # (destination bytes, offset in the stage, count)

        def flush(used: int) -> None:
            self.broadcast(0, _round_up(used, 16), root=root)
            for dst, off, n in pending:
                dst.copy_(stage[off:off + n])
            pending.clear()

        used = 0
        with torch.no_grad():
            for t in tensors:
                flat = t.detach().contiguous().view(-1).view(torch.uint8)     # a copy when t is not contiguous
                done = 0
                while done < flat.numel():
                    if used == cap:
                        flush(used)
                        used = 0
                    n = min(flat.numel() - done, cap - used)
                    piece = flat[done:done + n]
                    stage[used:used + n].copy_(piece)
                    pending.append((piece, used, n))
                    used = _round_up(used + n, 16) if used + n < cap else cap
                    done += n
                    moved += n
                    if used >= cap:
                        flush(cap)
                        used = 0
                if flat.data_ptr() != t.data_ptr():
                    # the staging copy was filled above only after the flush; write back through it
                    if pending:
                        flush(used)
                        used = 0
                    t.detach().copy_(flat.view(t.dtype).view(t.shape))
            if pending:
                flush(used)
        return moved

    def barrier(self, stream=None, nblocks: int = 1) -> None:
        check(lib().dlb_barrier(self._sig_array, self.state.data_ptr(), self.rank, self.world, nblocks, stream_handle(stream)))
        self.launches += 1

    def check(self, stream=None) -> None:
        """Raise if a device-side wait timed out (synchronises ``stream``)."""
        check(lib().dlb_comm_check(self.state.data_ptr(), stream_handle(stream)))

    # ---------------------------------------------------------------------------------------
    # measured selection
    # ---------------------------------------------------------------------------------------
    def tune(self, sizes_bytes: Sequence[int], dtype: torch.dtype = torch.float32, iters: int = 10,
             scratch_offset: int = 0, block_choices: Sequence[int] = (8, 16, 32, 48, 64),
             unroll_choices: Sequence[int] = (2, 4), cache: bool = True) -> Dict[int, dict]:
        """Collective.  Time every available variant (grid, and pipeline depth for nvls) on each size,
        take the max over ranks, and remember the fastest per size class.  Uses arena bytes
        ``[scratch_offset, scratch_offset + 2*max(sizes))``; contents are destroyed.  With ``cache``
        the table is also read from / written to a JSON file (``$DLB200_TUNE_CACHE`` or a file in the
        temp directory named after the GPU model and the world size), so later jobs on the same box
        skip the measurement."""
        # (pipeline depths: 4 won or tied everywhere in the 4/8-GPU sweeps and 8 lost from 16 MiB up, so 8
        # is not tried; 2 halves the requests in flight, which is what a full 64-CTA grid -- wanted by
        # the fused pack/SGD phases -- needs to stay near the 16-CTA optimum of the switch reduction)
        esize = torch.empty((), dtype=dtype).element_size()
        stream = torch.cuda.current_stream(self.device)
        report: Dict[int, dict] = {}
        wanted = {(dtype_code(dtype), max(_round_up(n, 16), 1).bit_length()) for n in sizes_bytes}
        if cache and self._load_tuned() and wanted <= set(self.tuned):
            return {n: {"best": self.tuned[(dtype_code(dtype), max(_round_up(n, 16), 1).bit_length())], "cached": True}
                    for n in sizes_bytes}
        algos = [a for a in ALGOS if a != "nvls" or self.has_multicast]
        for nbytes in sizes_bytes:
            nbytes = _round_up(nbytes, 16)
            numel = nbytes // esize
            out_off = scratch_offset + _round_up(nbytes, _ALIGN)
            results = {}
            for algo in algos:
                if algo == "oneshot" and nbytes * self.world > 256 << 20:
                    continue
                for nb in block_choices:
                    for unroll in (unroll_choices if algo == "nvls" else (0,)):
                        def run():
                            self.allreduce(scratch_offset, numel, dtype, algo, 1.0, out_offset=out_off, nblocks=nb,
                                           stream=stream, unroll=unroll)
                        for _ in range(2):
                            run()
                        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                        e0.record(stream)
                        for _ in range(iters):
                            run()
                        e1.record(stream)
                        e1.synchronize()
                        results[(algo, nb, unroll)] = e0.elapsed_time(e1) / iters
            keys = sorted(results)
            t = torch.tensor([results[k] for k in keys], dtype=torch.float64)
            if self.world > 1:
                if self._dist.get_backend(self.group) == "nccl":
                    t = t.to(self.device)
                self._dist.all_reduce(t, op=self._dist.ReduceOp.MAX, group=self.group)
                t = t.cpu()
            best = int(torch.argmin(t))
            self.tuned[(dtype_code(dtype), max(nbytes, 1).bit_length())] = keys[best]
            self.tuned_ms[(dtype_code(dtype), max(nbytes, 1).bit_length())] = dict(zip(keys, t.tolist()))
            report[nbytes] = {"best": keys[best], "ms": {f"{a}/{b}/u{u}": float(v) for (a, b, u), v in zip(keys, t.tolist())}}
        if cache:
            self._save_tuned()
        return report

    # the measured table on disk: {"<dtype code>:<size class>": {"best": [algo, nblocks, unroll], "ms": {"algo/nblocks/unroll": ms}}}
    def _tune_cache_path(self) -> str:
        import re
        import tempfile

        explicit = os.environ.get("DLB200_TUNE_CACHE")
        if explicit:
            return explicit
        gpu = re.sub(r"[^A-Za-z0-9]+", "_", torch.cuda.get_device_name(self.device))
        mc = "mc" if self.has_multicast else "nomc"
        # the library version is part of the name: a table measured with other kernels is not reused
        return os.path.join(tempfile.gettempdir(), f"dlb200_tune_v{lib().dlb_version()}_{gpu}_n{self.world}_{mc}.json")

    def _load_tuned(self) -> bool:
        """Collective: rank 0 reads the file, every rank gets the same table (or none)."""
        import json

        table = None
        if self.rank == 0:
            try:
                with open(self._tune_cache_path()) as f:
                    table = json.load(f)
            except (OSError, ValueError):
                table = None
        table = self._gather(table)[0] if self.world > 1 else table
        if not isinstance(table, dict) or not table:
            return False
        try:
            for k, v in table.items():
                code, cls = (int(x) for x in k.split(":"))
                best = v["best"]
                if best[0] == "nvls" and not self.has_multicast:
                    continue
                self.tuned.setdefault((code, cls), (best[0], int(best[1]), int(best[2])))
                ms = {}
                for name, t in v.get("ms", {}).items():
                    a, nb, un = name.split("/")
                    ms[(a, int(nb), int(un))] = float(t)
                if ms:
                    self.tuned_ms.setdefault((code, cls), ms)
        except (KeyError, ValueError, TypeError, IndexError):
            return False                          # a file of another version: measure again
        return True

    def _save_tuned(self) -> None:
        import json

        if self.rank != 0:
            return
        path = self._tune_cache_path()
        try:
            try:
                with open(path) as f:
                    table = json.load(f)
                if not isinstance(table, dict):
                    table = {}
            except (OSError, ValueError):
                table = {}
            for (c, cls), best in self.tuned.items():
                ms = self.tuned_ms.get((c, cls), {})
                table[f"{c}:{cls}"] = {"best": list(best), "ms": {f"{a}/{nb}/{un}": t for (a, nb, un), t in ms.items()}}
            with open(path + f".{os.getpid()}", "w") as f:
                json.dump(table, f)
            os.replace(path + f".{os.getpid()}", path)
        except OSError:
            pass                                   # a read-only temp directory only costs the next job a re-tune

    def close(self) -> None:
        if self.mode is None:
            return
        torch.cuda.synchronize(self.device)
        self._barrier()
        L = lib()
        self.local = None
        h = self._handles
        if self.mode == "ipc":
            for p in h.get("ipc_peers", []):
                L.dlb_ipc_close(p)
            self._barrier()
            L.dlb_mem_free(h["ipc_local"])
        else:
            if "mc" in h:
                mc, mp, size = h["mc"]
                L.dlb_vmm_unmap(mp, size)
                L.dlb_vmm_release(mc)
            for ph, pp, size in h.get("vmm_peers", []):
                L.dlb_vmm_unmap(pp, size)
                L.dlb_vmm_release(ph)
            self._barrier()
            lh, lp, size = h["vmm_local"]
            L.dlb_vmm_unmap(lp, size)
            L.dlb_vmm_release(lh)
        self.mode = None
        self._handles = {}


# ---------------------------------------------------------------------------------------------
# single-GPU emulation (tests)
# ---------------------------------------------------------------------------------------------
def emulated_flags(nranks: int, device) -> torch.Tensor:
    """Zeroed scratch for ``emulated_allreduce``; reuse it across calls (epochs only grow)."""
    return torch.zeros(lib().dlb_emulated_flags_bytes(nranks), dtype=torch.uint8, device=device)


def emulated_allreduce(bufs: Sequence[torch.Tensor], algo: str = "twoshot", scale: float = 1.0,
                       outs: Optional[Sequence[torch.Tensor]] = None, sgds: Optional[Sequence[FusedSgd]] = None,
                       flags: Optional[torch.Tensor] = None, nblocks: int = 4, nthreads: int = 256, stream=None,
                       pack=None, pack_scale: float = 1.0, trace: Optional[torch.Tensor] = None) -> None:
    """All ``len(bufs)`` ranks in one cooperative launch on one GPU (see ``dlb200.h``).  ``pack``: a
    ``PackPlan`` bound to *stacked* tensors (tensor i stored as ``[nranks, numel_i]``, row r belonging
    to emulated rank r) to run the fused pack prologue."""
    n = len(bufs)
    if flags is None:
        flags = emulated_flags(n, bufs[0].device)
    d = EmulatedDesc()
    d.algo, d.nranks = ALGO_CODE[algo], n
    d.bufs = ptr_array([b.data_ptr() for b in bufs])
    if outs is not None:
        d.outs = ptr_array([o.data_ptr() for o in outs])
    keep = None
    if sgds is not None:
        keep = (FusedSgd * n)(*sgds)
        d.sgds = keep
    d.flags = flags.data_ptr()
    d.numel, d.dtype, d.scale = bufs[0].numel(), dtype_code(bufs[0].dtype), scale
    d.nblocks, d.nthreads = nblocks, nthreads
    if pack is not None:
        d.tensors, d.tensor_dtypes, d.offsets, d.numels, d.ntensors = pack.ptrs, pack.dtypes, pack.offsets, pack.numels, pack.n
        d.pack_scale = pack_scale
    if trace is not None:
        d.trace = trace.data_ptr()
    check(lib().dlb_allreduce_emulated_ex(ctypes.byref(d), stream_handle(stream)))
