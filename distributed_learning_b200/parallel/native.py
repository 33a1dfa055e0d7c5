"""``B200Dist``: the native data-parallel wrapper, and ``FusedSGD``, its optimizer.

Parity: the wrapper surface of reference ``OurDist`` / ``SeqMergeDist`` (``src/ourdist.py:53-203``)
-- ``wrapper(data)``, ``sync_gradients()``, ``cleanup()``, attribute passthrough, the same bucket
rule -- with the machinery underneath replaced:

===========================================  =================================================
reference                                    here
===========================================  =================================================
tensor hook + spin on ``grad is None``       ``register_post_accumulate_grad_hook``
sender thread: ``fuse`` (187 copies, D2H)    ONE kernel per bucket on a side stream, ordered after the
``reducer.put`` -> mp.Queue -> host sum      bucket's last gradient by an event: its prologue gathers
-> gloo ring -> mp.Queue                     the gradient tensors into the peer-mapped buffer, then it
                                             all-reduces over NVLink peer memory
                                             (``csrc/allreduce.cu``), variant chosen per size
receiver thread: ``unfuse`` (187 copies)     ``.grad`` become views of the reduced buffer, or one
                                             ``dlb_unpack`` launch
``optimizer.step()`` (foreach SGD)           ``dlb_sgd_flat`` on the reduced buffer, or fused into
                                             the allreduce kernel's epilogue
``sync_gradients``: host waits on an Event   the compute stream waits on a CUDA event
===========================================  =================================================

At the first forward the wrapper probes the graph once to find parameters the loss never reaches
(GoogLeNet's aux heads) and leaves them out, lays the buckets out in the
peer-mapped arena (two regions per bucket, used alternately), and re-homes the parameters of each
bucket into one flat fp32 buffer (plus a bf16 copy for bf16 models) so the SGD update is a single
flat pass with no tensor table.

All ranks launch buckets in bucket order: a bucket that fills early waits for its
predecessors, so every rank issues the same sequence of communication kernels.
"""
from __future__ import annotations

import os
from typing import Dict, List, Optional

import torch
import torch.nn as nn

from ..ops import fused as F_ops
from ..ops._lib import DLB_MAX_TENSORS
from ..ops.comm import NativeComm, _round_up
from ..utils.logging import Level, print_d
from .buckets import Bucket, find_unused_parameters, fusion_grouping
from .wrappers import _Wrapper

_NORM_TYPES = (nn.modules.batchnorm._BatchNorm, nn.LayerNorm, nn.GroupNorm)
_REGION_ALIGN = 4096


class _NativeBucket:
    """A fusion bucket placed in the arena: two regions, flat master/momentum/shadow."""

    def __init__(self, bucket: Bucket, offsets: List[int], comm_dtype: torch.dtype):
        self.b = bucket
        self.id = bucket.bucket_id
        self.tensors = bucket.tensors
        self.numel = bucket.numel
        self.comm_dtype = comm_dtype
        self.region_offsets = offsets            # two byte offsets in the arena's data area
        self.regions: List[torch.Tensor] = []    # two flat views (comm dtype)
        self.cur = 0
        self.plan = F_ops.PackPlan.for_tensors(self.tensors, bucket.offsets)
        self.master: Optional[torch.Tensor] = None
        self.momentum: Optional[torch.Tensor] = None
        self.shadow: Optional[torch.Tensor] = None
        self.ready = 0
        self.launched = False
        self.reduced: Optional[torch.Tensor] = None   # where this step's averaged gradient lives
        self.algo = "twoshot"
        self.nblocks = 16
        self.unroll = 0
        self.sgd_done_in_kernel = False

    @property
    def nbytes(self) -> int:
        return self.numel * self.regions[0].element_size()


class B200Dist(_Wrapper):
    handles_dtype = True      # the trainer must not wrap forward in autocast

    def __init__(self, model, reducer=None, grouping_size: int = 25 * 1024 * 1024, grad_buff_device=None, *,
                 overlap: bool = True, algo: str = "auto", fused_sgd: bool = True, comm_dtype: str = "auto",
                 model_dtype: str = "fp32", cuda_graph: bool = False, group=None, comm: Optional[NativeComm] = None,
                 exclude_unused: bool = True, unpack: str = "view", tune: bool = False, dynamic_unused: bool = False,
                 fuse_pack: bool = True, sync_initial_state: bool = True, comm_kernel_when_alone: bool = False):
        super().__init__(model, reducer)
        d = self.__dict__
        d["grouping_size"] = int(grouping_size)
        d["overlap"] = bool(overlap)
        d["algo"] = algo
        d["fused_sgd"] = bool(fused_sgd)
        d["fuse_pack"] = bool(fuse_pack) and not os.environ.get("DLB200_NO_FUSE_PACK")
        # every replica starts from rank 0's parameters and buffers, copied by the
        # native broadcast kernel once the flat buffers exist -- no other communication library
        d["sync_initial_state"] = bool(sync_initial_state)
        # a world of one normally skips the communication kernel (pack, then the flat SGD kernel);
        # with this flag it launches the same fused kernel a multi-GPU job does, with nranks = 1, so
        # the whole gradient path (pack prologue, barrier, reduce, SGD epilogue) runs on a single GPU
        d["comm_kernel_when_alone"] = bool(comm_kernel_when_alone) or bool(os.environ.get("DLB200_COMM_KERNEL_WHEN_ALONE"))
        # dynamic_unused: the reference's behaviour (src/ourdist.py:137-172) -- walk the graph after
        # EVERY forward and hand a zero gradient to the parameters this step's loss does not reach,
        # so a model whose used-parameter set changes between steps still trains.  Every parameter
        # is bucketed (none is excluded), and the step cannot be captured in a CUDA graph.
        d["dynamic_unused"] = bool(dynamic_unused)
        if d["dynamic_unused"] and cuda_graph:
            raise ValueError("dynamic_unused walks the autograd graph every step; it cannot be combined with cuda_graph")
        d["exclude_unused"] = exclude_unused and not d["dynamic_unused"]
        d["unpack_mode"] = unpack                      # "view" (zero copy) | "copy" (dlb_unpack into own tensors)
        d["tune_at_start"] = tune
        d["model_dtype"] = torch.bfloat16 if model_dtype in ("bf16", torch.bfloat16) else torch.float32
        d["comm_dtype"] = {"auto": d["model_dtype"], "fp32": torch.float32, "bf16": torch.bfloat16}[comm_dtype] \
            if isinstance(comm_dtype, str) else comm_dtype
        d["device"] = next(model.parameters()).device
        if d["device"].type != "cuda":
            raise ValueError("B200Dist needs the model on a CUDA device")
        d["group"] = group
        d["comm"] = comm
        d["owns_comm"] = comm is None
        d["buckets"] = None
        d["unused"] = []
        d["comm_stream"] = torch.cuda.Stream(device=d["device"])
        d["_hooks"] = []
        d["_next"] = 0
        d["_armed"] = False
        d["_held"] = []
        d["_last_event"] = None
        d["optimizer"] = None
        d["_fp32_stash"] = None
        d["kernel_launches"] = 0
        # under a CUDA graph the regions cannot flip per step from Python; GraphedTrainStep captures
        # TWO graphs, one per region parity (`graph_parity`), and replays them alternately, so the
        # one-shot variant keeps the alternation that spares it a trailing barrier
        d["static_regions"] = bool(cuda_graph)
        d["graph_parity"] = 0
        d["alternate_graphs"] = bool(cuda_graph) and not os.environ.get("DLB200_SINGLE_GRAPH")
        d["nvtx"] = bool(int(os.environ.get("DLB200_NVTX", "0")))
        if d["model_dtype"] == torch.bfloat16:
            self._convert_to_bf16()

    # ------------------------------------------------------------------------------------------
    # set-up
    # ------------------------------------------------------------------------------------------
    def _convert_to_bf16(self) -> None:
        """bf16 compute with fp32 master weights: everything but the normalisation layers goes to
        bf16; the fp32 originals are kept until the flat master buffers exist."""
        stash = {}
        for mod in self.model.modules():
            if isinstance(mod, _NORM_TYPES):
                continue
            for name, p in mod.named_parameters(recurse=False):
                if p.dtype == torch.float32:
                    stash[p] = p.detach().clone()
                    p.data = p.data.to(torch.bfloat16)
        self.__dict__["_fp32_stash"] = stash

    def parameters_were_reloaded(self) -> None:
        """Called after a checkpoint was copied into the parameters in place."""
        if self.buckets is None:
            if self._fp32_stash is not None:
                for p in list(self._fp32_stash):
                    self._fp32_stash[p] = p.detach().float()
            return
        for nb in self.buckets:
            if nb.shadow is None:
                continue
            for t, off in zip(nb.tensors, nb.b.offsets):
                if t.dtype == torch.bfloat16:
                    nb.master[off:off + t.numel()].copy_(t.detach().reshape(-1))

    def master_state_dict(self) -> Dict[str, torch.Tensor]:
        """fp32 master copies of the bf16 parameters, keyed like ``model.state_dict()`` (empty for an
        fp32 model, whose parameters are their own masters).  Stored in checkpoints so a bf16 run
        resumes without losing the low bits of its master weights."""
        names = {id(p): n for n, p in self.model.named_parameters()}
        out: Dict[str, torch.Tensor] = {}
        if self.buckets is None:
            for p, master in (self._fp32_stash or {}).items():
                out[names[id(p)]] = master.detach().clone()
            return out
        for nb in self.buckets:
            if nb.shadow is None:
                continue
            for t, off in zip(nb.tensors, nb.b.offsets):
                if t.dtype == torch.bfloat16:
                    out[names[id(t)]] = nb.master[off:off + t.numel()].view(t.shape).clone()
        return out

    def load_master_state_dict(self, sd: Optional[Dict[str, torch.Tensor]]) -> None:
        """Inverse of ``master_state_dict``; call after the (bf16) parameters were restored."""
        if not sd:
            return
        by_name = dict(self.model.named_parameters())
        if self.buckets is None:
            if self._fp32_stash is not None:
                for n, v in sd.items():
                    p = by_name.get(n)
                    if p is not None and p in self._fp32_stash:
                        self._fp32_stash[p] = v.detach().to(p.device, torch.float32)
            return
        names = {id(p): n for n, p in by_name.items()}
        with torch.no_grad():
            for nb in self.buckets:
                for t, off in zip(nb.tensors, nb.b.offsets):
                    v = sd.get(names[id(t)])
                    if v is not None and t.dtype == torch.bfloat16:
                        nb.master[off:off + t.numel()].copy_(v.reshape(-1))

    def _probe_unused(self, data) -> List[torch.Tensor]:
        """One throw-away forward to find parameters the loss does not reach; RNG state and
        module buffers (BatchNorm statistics) are restored so the real step is unaffected."""
        params = [p for p in self.model.parameters() if p.requires_grad]
        if not self.exclude_unused:
            return []
        cpu_rng = torch.get_rng_state()
        cuda_rng = torch.cuda.get_rng_state(self.device)
        saved = [(b, b.detach().clone()) for b in self.model.buffers()]
        out = self.model(data)
        if isinstance(out, (tuple, list)):
            out = out[0]
        unused = find_unused_parameters(out, params)
        del out
        with torch.no_grad():
            for b, v in saved:
                b.copy_(v)
        torch.set_rng_state(cpu_rng)
        torch.cuda.set_rng_state(cuda_rng, self.device)
        return unused

    def _build(self, data) -> None:
        d = self.__dict__
        d["unused"] = self._probe_unused(data)
        dead = {id(p) for p in self.unused}
        params = [p for p in self.model.parameters() if p.requires_grad and id(p) not in dead]
        if self.unused:
            print_d(f"B200Dist: {len(self.unused)} parameter tensors never receive a gradient and are left out", Level.INFO)
        esize = torch.empty((), dtype=self.comm_dtype).element_size()
        plain = [Bucket(i, ts, dtype=self.comm_dtype, aligned=True, allocate=False)
                 for i, ts in enumerate(fusion_grouping(params, self.grouping_size))]
        # arena layout: two regions per bucket
        off, buckets = 0, []
        for b in plain:
            size = _round_up(b.numel * esize, _REGION_ALIGN)
            buckets.append(_NativeBucket(b, [off, off + size], self.comm_dtype))
            off += 2 * size
        if self.comm is None:
            d["comm"] = NativeComm(self.group, self.device)
        if self.comm.mode is None:
            self.comm.allocate(off)
        elif self.comm.nbytes - self.comm.data_offset < off:
            raise RuntimeError("the communicator's arena is too small for this model's buckets")
        world = self.comm.world
        for nb in buckets:
            nb.regions = [self.comm.view(o, nb.numel, self.comm_dtype) for o in nb.region_offsets]
            nb.b.grad_buffer = nb.regions[0]
            if world > 1 or self.comm_kernel_when_alone:
                nb.algo, nb.nblocks, nb.unroll = self.comm.choose(nb.numel * esize, self.comm_dtype, self.algo,
                                                                  hbm_bytes=self._fused_hbm_bytes(nb.numel, esize))
            self._flatten_parameters(nb)
        d["buckets"] = buckets
        if world > 1 and self.sync_initial_state:
            flats = [nb.master for nb in buckets] + [nb.shadow for nb in buckets if nb.shadow is not None]
            bucketed = {id(t) for nb in buckets for t in nb.tensors}
            rest = [p for p in self.model.parameters() if id(p) not in bucketed] + list(self.model.buffers())
            moved = self.comm.broadcast_tensors(flats + rest, root=0)
            print_d(f"B200Dist: initial state broadcast from rank 0 by the native kernel ({moved / 2**20:.1f} MiB)", Level.DEBUG)
        for nb in buckets:
            for t in nb.tensors:
                self._hooks.append(t.register_post_accumulate_grad_hook(self._make_hook(nb)))
        d["_fp32_stash"] = None
        opt = self.optimizer
        if opt is not None and opt._pending_state is not None:
            # a checkpoint was loaded before the flat buffers existed; the momentum has to be in
            # place before the first bucket is launched, because the update runs inside that kernel
            opt._apply_state(opt._pending_state)
            opt._pending_state = None
        if self.tune_at_start and world > 1:
            sizes = sorted({nb.numel * esize for nb in buckets})
            self.comm.tune(sizes, self.comm_dtype)
            for nb in buckets:
                nb.algo, nb.nblocks, nb.unroll = self.comm.choose(nb.numel * esize, self.comm_dtype, self.algo,
                                                                  hbm_bytes=self._fused_hbm_bytes(nb.numel, esize))
        print_d("B200Dist: " + ", ".join(f"bucket{nb.id}: {len(nb.tensors)} tensors {nb.numel * esize / 2**20:.2f} MiB "
                                         f"{nb.algo}/{nb.nblocks}" + (f"/u{nb.unroll}" if nb.algo == "nvls" else "") for nb in buckets)
                + f" (arena {self.comm.mode}, multicast {'yes' if self.comm.has_multicast else 'no'}"
                + (f": {self.comm.mc_note}" if self.comm.mc_note else "") + ")", Level.INFO)

    def _fused_hbm_bytes(self, numel: int, esize: int) -> int:
        """Local HBM traffic that rides in a bucket's kernel besides the allreduce: the pack prologue
        reads the gradients and writes the buffer; the SGD epilogue reads the reduced gradient, the
        fp32 parameter and momentum and writes both back (plus the bf16 copy for bf16 models)."""
        total = 0
        if self.fuse_pack:
            total += 2 * numel * esize
        if self.fused_sgd:
            total += numel * (esize + 16 + (2 if self.model_dtype == torch.bfloat16 else 0))
        return total

    def _flatten_parameters(self, nb: _NativeBucket) -> None:
        """Re-home the bucket's parameters in one flat fp32 buffer (and a bf16 copy for bf16
        parameters).  fp32 parameters become views of the master buffer, bf16 parameters views
        of the shadow buffer; values are unchanged."""
        dev = self.device
        nb.master = torch.zeros(nb.numel, dtype=torch.float32, device=dev)
        nb.momentum = torch.zeros(nb.numel, dtype=torch.float32, device=dev)
        need_shadow = any(t.dtype == torch.bfloat16 for t in nb.tensors)
        nb.shadow = torch.zeros(nb.numel, dtype=torch.bfloat16, device=dev) if need_shadow else None
        stash = self._fp32_stash or {}
        with torch.no_grad():
            for t, off in zip(nb.tensors, nb.b.offsets):
                n = t.numel()
                src = stash.get(t, t.detach())
                nb.master[off:off + n].copy_(src.reshape(-1))
                if t.dtype == torch.bfloat16:
                    nb.shadow[off:off + n].copy_(nb.master[off:off + n])
                    t.data = nb.shadow[off:off + n].view(t.shape)
                else:
                    t.data = nb.master[off:off + n].view(t.shape)

    # ------------------------------------------------------------------------------------------
    # the step
    # ------------------------------------------------------------------------------------------
    def _make_hook(self, nb: _NativeBucket):
        def hook(_param):
            if not self._armed:
                return
            nb.ready += 1
            if nb.ready == len(nb.tensors) and self.overlap:
                self._launch_ready()
        return hook

    def _launch_ready(self) -> None:
        buckets = self.buckets
        while self._next < len(buckets) and buckets[self._next].ready == len(buckets[self._next].tensors):
            self._launch(buckets[self._next])
            self.__dict__["_next"] = self._next + 1

    def _launch(self, nb: _NativeBucket) -> None:
        """Pack + allreduce (+ SGD) of one bucket on the communication stream."""
        comm, world = self.comm, self.comm.world
        cs = self.comm_stream
        if self.nvtx:
            torch.cuda.nvtx.range_push(f"bucket{nb.id} {nb.algo}/{nb.nblocks} {nb.nbytes >> 10} KiB")
        ready = torch.cuda.Event()
        ready.record(torch.cuda.current_stream(self.device))
        cs.wait_event(ready)
        grads = [t.grad for t in nb.tensors]
        self._held.append(grads)              # keep the memory alive until the compute stream has waited for us
        if self.static_regions:
            nb.cur = self.graph_parity if self.alternate_graphs else 0
        src, dst = nb.regions[nb.cur], nb.regions[1 - nb.cur]
        opt = self.optimizer
        use_comm = world > 1 or self.comm_kernel_when_alone
        in_kernel = self.fused_sgd and opt is not None and use_comm
        # one kernel per bucket: the gather of the gradient tensors is the allreduce kernel's prologue
        fuse = self.fuse_pack and use_comm and len(nb.tensors) <= DLB_MAX_TENSORS
        launches = 0
        if not fuse:
            nb.plan.pack(src, grads, 1.0 / world, stream=cs)
            launches += 1
        if use_comm:
            sgd = None
            if in_kernel:
                sgd = F_ops.fused_sgd_args(nb.master, nb.momentum, opt.hyper(), nb.shadow)
            kw = dict(sgd=sgd, nblocks=nb.nblocks, stream=cs, unroll=nb.unroll)
            if fuse:
                kw.update(pack=nb.plan.bind(grads), pack_scale=1.0 / world)
            if nb.algo == "oneshot":
                comm.allreduce(nb.region_offsets[nb.cur], nb.numel, nb.comm_dtype, "oneshot", 1.0,
                               out_offset=nb.region_offsets[1 - nb.cur],
                               end_barrier=self.static_regions and not self.alternate_graphs, **kw)
                nb.reduced = dst
            else:
                comm.allreduce(nb.region_offsets[nb.cur], nb.numel, nb.comm_dtype, nb.algo, 1.0, **kw)
                nb.reduced = src
            launches += 1
        else:
            nb.reduced = src
        nb.sgd_done_in_kernel = in_kernel
        if not self.static_regions:
            nb.cur ^= 1          # alternate regions: lets one-shot go without a trailing barrier
        if self.unpack_mode == "copy":
            fresh = [torch.empty_like(t) for t in nb.tensors]
            with torch.cuda.stream(cs):
                nb.plan.unpack(nb.reduced, fresh, 1.0, stream=cs)
            launches += 1
            views = fresh
        else:
            views = nb.b.grad_views(nb.reduced)
        for t, v in zip(nb.tensors, views):
            t.grad = v if v.dtype == t.dtype else None     # a view in another dtype cannot be a .grad
        done = torch.cuda.Event()
        done.record(cs)
        nb.launched = True
        self.__dict__["_last_event"] = done
        self.__dict__["kernel_launches"] = self.kernel_launches + launches
        if self.nvtx:
            torch.cuda.nvtx.range_pop()

    def forward(self, data):
        if self.model_dtype == torch.bfloat16 and data.is_floating_point() and data.dtype != torch.bfloat16:
            data = data.to(torch.bfloat16)
        if self.buckets is None:
            self._build(data)
        for nb in self.buckets:
            nb.ready = 0
            nb.launched = False
        d = self.__dict__
        d["_next"] = 0
        d["_armed"] = True
        d["_last_event"] = None
        out = self.model(data)
        if self.dynamic_unused:
            self._mark_unused_ready(out)
        return out

    def _mark_unused_ready(self, out) -> None:
        """Per-step walk: parameters the output does not depend on count as ready with no gradient
        (the pack writes zeros for them), so their bucket still completes and every rank launches
        the same kernels."""
        first = out[0] if isinstance(out, (tuple, list)) else out
        bucketed = [t for nb in self.buckets for t in nb.tensors]
        dead = {id(p) for p in find_unused_parameters(first, bucketed)}
        if not dead:
            return
        for nb in self.buckets:
            for t in nb.tensors:
                if id(t) in dead:
                    t.grad = None
                    nb.ready += 1

    def sync_gradients(self) -> None:
        if self.buckets is None:
            return
        if not self.overlap or self._next < len(self.buckets):
            # fusion-only mode, or buckets that autograd never completed (should not happen)
            for nb in self.buckets[self._next:]:
                if nb.ready != len(nb.tensors):
                    raise RuntimeError(f"bucket {nb.id}: {nb.ready}/{len(nb.tensors)} gradients arrived; "
                                       "did the set of used parameters change?")
                self._launch(nb)
            self.__dict__["_next"] = len(self.buckets)
        if self._last_event is not None:
            torch.cuda.current_stream(self.device).wait_event(self._last_event)
        self._held.clear()
        self.__dict__["_armed"] = False

    # ------------------------------------------------------------------------------------------
    # state snapshots (used to make the CUDA-graph warm-up free of side effects)
    # ------------------------------------------------------------------------------------------
    def build_now(self, data) -> None:
        """Do the first-forward set-up (probe, buckets, arena, flat parameters) without a step."""
        if self.model_dtype == torch.bfloat16 and data.is_floating_point() and data.dtype != torch.bfloat16:
            data = data.to(torch.bfloat16)
        if self.buckets is None:
            self._build(data)

    def snapshot_state(self) -> dict:
        """Copies of everything a training step changes: flat master / momentum / shadow buffers,
        module buffers (BatchNorm statistics), the optimizer's step count and the RNG states."""
        if self.buckets is None:
            raise RuntimeError("snapshot_state needs the buckets: call build_now(data) first")
        opt = self.optimizer
        return {
            "buckets": [(nb.master.clone(), nb.momentum.clone(), None if nb.shadow is None else nb.shadow.clone())
                        for nb in self.buckets],
            "buffers": [b.detach().clone() for b in self.model.buffers()],
            "steps": None if opt is None else opt.steps,
            "cpu_rng": torch.get_rng_state(),
            "cuda_rng": torch.cuda.get_rng_state(self.device),
        }

    def restore_state(self, snap: dict) -> None:
        """In-place inverse of ``snapshot_state`` (addresses are unchanged, so a captured graph stays valid)."""
        with torch.no_grad():
            for nb, (master, mom, shadow) in zip(self.buckets, snap["buckets"]):
                nb.master.copy_(master)
                nb.momentum.copy_(mom)
                if shadow is not None:
                    nb.shadow.copy_(shadow)
                nb.sgd_done_in_kernel = False
                nb.reduced = None
            for b, v in zip(self.model.buffers(), snap["buffers"]):
                b.copy_(v)
        if self.optimizer is not None and snap["steps"] is not None:
            self.optimizer.steps = snap["steps"]
        torch.set_rng_state(snap["cpu_rng"])
        torch.cuda.set_rng_state(snap["cuda_rng"], self.device)
        for p in self.model.parameters():
            p.grad = None

    def make_optimizer(self, lr: float = 0.01, momentum: float = 0.5, **kw) -> "FusedSGD":
        opt = FusedSGD(self, lr=lr, momentum=momentum, **kw)
        self.__dict__["optimizer"] = opt
        return opt

    def cleanup(self) -> None:
        for h in self._hooks:
            h.remove()
        self._hooks.clear()
        torch.cuda.synchronize(self.device)
        if self.comm is not None and self.comm.mode is not None:
            self.comm.check()
            if self.owns_comm:
                self.comm.close()


class FusedSGD:
    """SGD-momentum over ``B200Dist``'s flat buffers (reference: ``optim.SGD(lr=0.01,
    momentum=0.5)``, ``src/main.py:36``).  ``step`` launches one ``dlb_sgd_flat`` per bucket on the
    current stream, reading the reduced gradient where the allreduce left it -- or nothing at all
    for buckets whose update already ran inside the allreduce kernel.  ``state_dict`` has
    ``torch.optim.SGD``'s layout, so checkpoints move between backends."""

    def __init__(self, wrapper: B200Dist, lr: float = 0.01, momentum: float = 0.5, dampening: float = 0.0,
                 weight_decay: float = 0.0, nesterov: bool = False):
        if nesterov and (momentum <= 0 or dampening != 0):
            raise ValueError("Nesterov momentum requires a momentum and zero dampening")
        if dampening != 0 and getattr(wrapper, "static_regions", False):
            raise ValueError("dampening != 0 treats the first step differently (torch seeds the momentum buffer with "
                             "the raw gradient), which a captured CUDA graph cannot do: use eager mode")
        self.wrapper = wrapper
        self.param_groups = [dict(lr=lr, momentum=momentum, dampening=dampening, weight_decay=weight_decay,
                                  nesterov=nesterov, maximize=False, foreach=None, differentiable=False, fused=None)]
        self._hyper_key = None
        self._hyper = None
        self._pending_state: Optional[dict] = None
        self.steps = 0

    def hyper(self):
        """Kernel hyper-parameters of the coming step.  torch.optim.SGD seeds the momentum buffer with
        the raw gradient on the first step (no dampening); with the buffer starting at zero that is
        the ordinary update with dampening 0, so the first step is issued that way."""
        g = self.param_groups[0]
        damp = g["dampening"] if self.steps > 0 else 0.0
        key = (g["lr"], g["momentum"], damp, g["weight_decay"], g["nesterov"])
        if key != self._hyper_key:
            self._hyper = F_ops.make_hyper(*key[:4], nesterov=key[4])
            self._hyper_key = key
        return self._hyper

    def zero_grad(self, set_to_none: bool = True) -> None:
        for p in self.wrapper.model.parameters():
            if set_to_none:
                p.grad = None
            elif p.grad is not None:
                p.grad.detach_().zero_()

    def step(self) -> None:
        w = self.wrapper
        if w.buckets is None:
            return
        if self._pending_state is not None:
            self._apply_state(self._pending_state)
            self._pending_state = None
        h = self.hyper()
        for nb in w.buckets:
            if nb.sgd_done_in_kernel:
                nb.sgd_done_in_kernel = False
                continue
            if nb.reduced is None:
                continue
            F_ops.sgd_flat(nb.master, nb.momentum, nb.reduced, h, nb.shadow)
            w.__dict__["kernel_launches"] = w.kernel_launches + 1
        self.steps += 1

    # -- torch.optim.SGD-compatible state ----------------------------------------------------------
    def _param_index(self) -> Dict[int, int]:
        return {id(p): i for i, p in enumerate(self.wrapper.model.parameters())}

    def state_dict(self) -> dict:
        w = self.wrapper
        index = self._param_index()
        state = {}
        if w.buckets is not None and self.steps > 0:
            for nb in w.buckets:
                for t, off in zip(nb.tensors, nb.b.offsets):
                    state[index[id(t)]] = {"momentum_buffer": nb.momentum[off:off + t.numel()].view(t.shape).clone()}
        elif self._pending_state is not None:
            state = self._pending_state["state"]
        group = dict(self.param_groups[0])
        group["params"] = list(range(len(index)))
        return {"state": state, "param_groups": [group]}

    def load_state_dict(self, sd: dict) -> None:
        g = sd["param_groups"][0]
        for k in ("lr", "momentum", "dampening", "weight_decay", "nesterov"):
            if k in g:
                self.param_groups[0][k] = g[k]
        if self.wrapper.buckets is None:
            self._pending_state = sd          # the flat buffers do not exist before the first forward
        else:
            self._apply_state(sd)

    def _apply_state(self, sd: dict) -> None:
        index = self._param_index()
        for nb in self.wrapper.buckets:
            for t, off in zip(nb.tensors, nb.b.offsets):
                st = sd["state"].get(index[id(t)]) or sd["state"].get(str(index[id(t)]))
                if st is not None and st.get("momentum_buffer") is not None:
                    nb.momentum[off:off + t.numel()].copy_(st["momentum_buffer"].reshape(-1))
                    self.steps = max(self.steps, 1)


class NativeReducer:
    """The in-kernel NVLink allreduce behind the reference's reducer surface (``put`` / ``get`` /
    ``cleanup``, ``src/reducers.py``), so that any host-driven wrapper (``OurDist``,
    ``SeqMergeDist``, ``SeqDist``) can be combined with it -- the reference's "any wrapper with any
    reducer and any allreduce".  ``put`` stages the buffer in the peer-mapped arena, averages it
    across the ranks with one kernel and copies it back, all on the reducer's own stream; ``get``
    makes the caller's stream wait for that.  ``B200Dist`` is the fast path (no staging copies, fused
    update); this class exists for composability.

    Buffers must be fp32 or bf16 CUDA tensors of at most ``capacity_bytes``.  Every rank must
    ``put`` buffers of the same sizes in the same order."""

    def __init__(self, comm: Optional[NativeComm] = None, capacity_bytes: int = 64 << 20, group=None, device=None):
        import queue

        self.owns_comm = comm is None
        self.comm = comm if comm is not None else NativeComm(group, device)
        self.capacity = _round_up(int(capacity_bytes), _REGION_ALIGN)
        if self.comm.mode is None:
            self.comm.allocate(self.capacity)
        elif self.comm.nbytes - self.comm.data_offset < self.capacity:
            raise RuntimeError("the communicator's arena is smaller than capacity_bytes")
        self.device = self.comm.device
        self.stream = torch.cuda.Stream(device=self.device)
        self.pending = queue.Queue()     # get() blocks until the matching put() has been issued (threaded wrappers)
        self.launches = 0

    def put(self, buff: torch.Tensor) -> None:
        if not buff.is_cuda or buff.dtype not in (torch.float32, torch.bfloat16):
            raise TypeError("NativeReducer reduces fp32 or bf16 CUDA buffers")
        if not buff.is_contiguous():
            raise ValueError("NativeReducer needs a contiguous buffer (it is reduced in place)")
        flat = buff.view(-1)
        n, esize = flat.numel(), flat.element_size()
        per_vec = 16 // esize
        padded = _round_up(n, per_vec)
        if padded * esize > self.capacity:
            raise ValueError(f"buffer of {n * esize} bytes exceeds the reducer's capacity of {self.capacity}")
        comm, world = self.comm, self.comm.world
        with torch.cuda.device(self.device):
            ready = torch.cuda.Event()
            ready.record(torch.cuda.current_stream(self.device))
            self.stream.wait_event(ready)
            with torch.cuda.stream(self.stream):
                if world > 1:
                    view = comm.view(0, padded, buff.dtype)
                    view[:n].copy_(flat)
                    if padded > n:
                        view[n:].zero_()
                    algo, nblocks, unroll = comm.choose(padded * esize, buff.dtype, "auto")
                    if algo == "oneshot":        # in place, and safe to refill right away
                        algo = "twoshot"
                    comm.allreduce(0, padded, buff.dtype, algo, 1.0 / world, nblocks=nblocks, stream=self.stream, unroll=unroll)
                    flat.copy_(view[:n])
                    self.launches += 1
                done = torch.cuda.Event()
                done.record(self.stream)
        self.pending.put((buff, done))

    def get(self) -> torch.Tensor:
        buff, done = self.pending.get()
        torch.cuda.current_stream(self.device).wait_event(done)
        return buff

    def cleanup(self) -> None:
        torch.cuda.synchronize(self.device)
        if self.comm.mode is not None:
            self.comm.check()
            if self.owns_comm:
                self.comm.close()
