"""Experiment registry: every ``--experiment`` name of the reference.

Parity: reference ``src/main.py:116-301``; ``main.py`` dispatches by name (``:333``).  The 22
names, their wrapper / reducer / allreduce combination, their ``experiment_name`` file prefix
and the composite suites are the same.  ``main_seq`` gets the prefix ``seq``
(the reference forgets to set one, Q7).

How a name maps onto a backend:

``gloo`` (plumbing, CPU or GPU with host buffers)
    Reference semantics.  2-step names spawn ``d`` workers and run the node pump in the main
    process; 1-step names run in-process with one rank per device.
``nccl`` (baseline)
    One process per GPU in a worker-level group.  Fusion buffers are on the device; ring names
    use grouped send/recv (``ring_allreduce_gpu``), ``central`` names use send/recv to rank 0,
    ``main_ddp`` is torch DDP.
``b200`` (product)
    One process per GPU; ``B200Dist`` with the in-kernel NVLink allreduce.  Behind NVSwitch
    "sum in the node, then ring between nodes" and "ring among all" are the same direct
    all-to-all, so 1-step and 2-step names run the same kernels (P8).  Ring names let the
    variant be chosen from the measured table (``--algo`` overrides); ``central`` names force
    the one-shot variant, in which every GPU is the centre for its own copy.
"""
from __future__ import annotations

import functools
import os
from typing import Callable, Dict

import torch.multiprocessing as mp

from .parallel.allreduce import central_allreduce, ring_allreduce, ring_allreduce_gpu
from .parallel.reducers import NodeAggregateReducerCPU, ReduceImmediately
from .parallel.wrappers import DDPDist, OurDist, SeqDist, SeqMergeDist, SingleDist, WarmupDist
from .trainer import worker_process, worker_process_wrapper

EXPERIMENTS: Dict[str, Callable] = {}

fusion_test_sizes_k = [256, 1024, 4 * 1024, 16 * 1024, 64 * 1024]


def experiment(fn: Callable) -> Callable:
    EXPERIMENTS[fn.__name__] = fn
    return fn


class _NullReducer:
    def cleanup(self):
        pass


def _device_backend(config) -> bool:
    return config.backend in ("nccl", "b200") and config.use_gpu == 1


# ------------------------------------------------------------------------------------------
# process models
# ------------------------------------------------------------------------------------------
def _join(procs) -> None:
    bad = []
    for p in procs:
        p.join()
        if p.exitcode not in (0, None):
            bad.append(p.exitcode)
    if bad:
        raise RuntimeError(f"{len(bad)} worker process(es) failed (exit codes {bad})")


def _spawn_with_pump(config, reduce_fn, gpu_reduce: bool = False, groups=None) -> None:
    """2-step process model of the reference: ``d`` workers + the node pump here."""
    ctx = mp.get_context("spawn")
    node = NodeAggregateReducerCPU(reduce_fn, config.node_dev, groups=groups, ctx=ctx)
    procs = []
    for i in range(config.node_dev):
        p = ctx.Process(target=worker_process_wrapper, args=(config.rank, i, config, node.reducers[i], gpu_reduce))
        p.start()
        procs.append(p)
    node.pump()
    _join(procs)


def _device_worker(node_id: int, worker_id: int, config) -> None:
    """A spawned one-process-per-GPU worker of a device backend: joins the worker-level group."""
    from .launcher import init_worker_group

    init_worker_group(config, node_id, worker_id)
    try:
        _run_device_worker(node_id, worker_id, config)
    finally:
        import torch.distributed as dist

        if dist.is_initialized():
            dist.destroy_process_group()


def _run_device_worker(node_id: int, worker_id: int, config) -> None:
    reducer = config.make_reducer() if getattr(config, "make_reducer", None) else _NullReducer()
    worker_process(node_id, worker_id, config, reducer, True)


def _run_device(config) -> None:
    """One process per GPU.  ``d == 1``: this process is the worker (its group already exists);
    ``d > 1``: spawn ``d`` workers that form their own group of ``D`` ranks."""
    if config.node_dev == 1:
        _run_device_worker(config.rank, 0, config)
        return
    ctx = mp.get_context("spawn")
    procs = []
    for i in range(config.node_dev):
        p = ctx.Process(target=_device_worker_entry, args=(config.rank, i, config))
        p.start()
        procs.append(p)
    _join(procs)


def _device_worker_entry(node_id, worker_id, config):
    import sys
    import traceback

    try:
        _device_worker(node_id, worker_id, config)
    except Exception:
        traceback.print_exc(file=sys.stdout)
        sys.stdout.flush()
        os._exit(1)


# ------------------------------------------------------------------------------------------
# building blocks
# ------------------------------------------------------------------------------------------
def _native_factory(overlap: bool, algo: str, config):
    from .parallel.native import B200Dist

    return functools.partial(
        B200Dist, overlap=overlap, algo=algo if config.algo == "auto" else config.algo,
        fused_sgd=bool(config.fused_sgd), fuse_pack=bool(getattr(config, "fuse_pack", 1)), comm_dtype=config.comm_dtype,
        model_dtype=config.dtype,
        cuda_graph=bool(getattr(config, "cuda_graph", 0)), dynamic_unused=bool(getattr(config, "dynamic_unused", 0)))


def _immediate(fn):
    return ReduceImmediately(fn)


def _run(config, name: str, wrapper, algo: str, two_step: bool, overlap: bool = True, gpu_buffers: bool = False) -> None:
    """Run one named experiment on whatever backend ``config`` selects."""
    config.experiment_name = name
    if config.backend == "b200":
        config.distribute_model = _native_factory(overlap, "oneshot" if algo == "central" else "auto", config)
        config.make_reducer = None
        _run_device(config)
        return
    if getattr(config, "dynamic_unused", 0) and wrapper in (OurDist, SeqMergeDist):
        wrapper = functools.partial(wrapper, dynamic_unused=True)
    config.distribute_model = wrapper
    if _device_backend(config):                      # nccl baseline, one process per GPU
        fn = {"ring": ring_allreduce_gpu, "central": central_allreduce}[algo]
        config.make_reducer = functools.partial(_immediate, fn)
        _run_device(config)
        return
    fn = {"ring": ring_allreduce, "central": central_allreduce}[algo]
    if two_step:
        _spawn_with_pump(config, fn, gpu_reduce=gpu_buffers)
    else:
        worker_process(config.rank, 0, config, ReduceImmediately(fn), gpu_buffers)


def _with_grouping(config, grouping_size: int, fn: Callable) -> None:
    old = config.grouping_size
    config.grouping_size = grouping_size
    try:
        fn()
    finally:
        config.grouping_size = old


# ------------------------------------------------------------------------------------------
# the reference's experiments
# ------------------------------------------------------------------------------------------
@experiment
def main_warmup(config):
    config.distribute_model = WarmupDist
    config.experiment_name = "warmup"
    config.make_reducer = None
    if _device_backend(config):
        _run_device(config)
        return
    ctx = mp.get_context("spawn")
    procs = []
    for i in range(config.node_dev):
        p = ctx.Process(target=worker_process_wrapper, args=(config.rank, i, config, _NullReducer()))
        p.start()
        procs.append(p)
    _join(procs)


@experiment
def main_ourdist(config):
    _run(config, "ourdist", OurDist, "ring", two_step=True)


@experiment
def main_ourdist_nccl(config):
    if config.backend == "gloo":
        # reference: device-resident buffers summed in the node process; works for d == 1 only (Q5)
        config.experiment_name = "ourdist_nccl"
        config.distribute_model = OurDist
        _spawn_with_pump(config, ring_allreduce, gpu_reduce=True)
        return
    _run(config, "ourdist_nccl", OurDist, "ring", two_step=True, gpu_buffers=True)


@experiment
def main_seq(config):
    if config.backend == "b200":     # SeqDist = one reduction per parameter, after backward
        _with_grouping(config, 0, lambda: _run(config, "seq", SeqDist, "ring", two_step=True, overlap=False))
        return
    _run(config, "seq", SeqDist, "ring", two_step=True, overlap=False)


@experiment
def main_seq_merge(config):
    _run(config, "seq_merge", SeqMergeDist, "ring", two_step=True, overlap=False)


@experiment
def main_overlap(config):
    _with_grouping(config, 0, lambda: _run(config, "overlap", OurDist, "ring", two_step=True))


@experiment
def main_ddp(config):
    config.experiment_name = "ddp"
    config.make_reducer = None
    # on the native backend main_ddp stays what it is in the reference's suites: the library
    # baseline (it creates an NCCL sub-group for itself; the b200 job's own group is gloo only).
    if _device_backend(config):
        config.distribute_model = _ddp_cuda
        _run_device(config)
        return
    config.distribute_model = DDPDist
    worker_process(config.rank, 0, config, _NullReducer())


_NCCL_SUBGROUP = None


def _ddp_cuda(model, reducer, grouping_size, grad_buff_device="cpu"):
    import torch.distributed as dist

    global _NCCL_SUBGROUP
    dev = next(model.parameters()).device
    pg = None
    if "nccl" not in str(dist.get_backend()):
        # a b200 job has a gloo-only group; the library baseline brings its own NCCL communicator
        if _NCCL_SUBGROUP is None:
            _NCCL_SUBGROUP = dist.new_group(backend="nccl")
        pg = _NCCL_SUBGROUP
    return DDPDist(model, reducer, grouping_size, grad_buff_device, device_ids=[dev.index], process_group=pg)


@experiment
def main_central_reduce(config):
    _run(config, "central_node_reduce", OurDist, "central", two_step=True)


@experiment
def main_onestep_reduce(config):
    _run(config, "onestep_reduce", OurDist, "ring", two_step=False)


@experiment
def main_onestep_central(config):
    _run(config, "onestep_central", OurDist, "central", two_step=False)


@experiment
def main_onestep_overlap(config):
    _with_grouping(config, 0, lambda: _run(config, "onestep_overlap", OurDist, "ring", two_step=False))


@experiment
def main_onestep_seq_merge(config):
    _run(config, "onestep_seq_merge", SeqMergeDist, "ring", two_step=False, overlap=False)


@experiment
def main_single(config):
    config.distribute_model = SingleDist
    config.experiment_name = "single"
    config.make_reducer = None
    worker_process(0, 0, config, _NullReducer())


@experiment
def experiment1(config):
    for f in (main_warmup, main_ourdist, main_seq_merge, main_overlap, main_central_reduce):
        f(config)


@experiment
def experiment2(config):
    for f in (main_warmup, main_ddp, main_onestep_reduce, main_onestep_central):
        f(config)


@experiment
def experiment3(config):
    for f in (main_warmup, main_ddp, main_onestep_reduce, main_onestep_seq_merge, main_onestep_overlap):
        f(config)


@experiment
def experiment_nccl(config):
    for f in (main_warmup, main_ddp, main_ourdist_nccl):
        f(config)


@experiment
def experiment_ourdist_nccl(config):
    for f in (main_warmup, main_ourdist_nccl):
        f(config)


@experiment
def experiment_single(config):
    for f in (main_warmup, main_single):
        f(config)


@experiment
def experiment_onestep_central(config):
    for f in (main_warmup, main_onestep_central):
        f(config)


def fusion_experiment(config, main_f, sizes_k=None) -> None:
    """Warm up, then run ``main_f`` once per bucket size, each into its own sub-folder
    (reference ``src/main.py:287-295``).  ``sizes_k`` in KiB; 0 means one tensor per bucket."""
    main_warmup(config)
    folder, old = config.folder, config.grouping_size
    try:
        for size_k in (fusion_test_sizes_k if sizes_k is None else sizes_k):
            config.grouping_size = size_k * 1024
            config.folder = f"{folder}/{size_k}"
            os.makedirs(config.folder, exist_ok=True)
            main_f(config)
    finally:
        config.folder, config.grouping_size = folder, old


@experiment
def fusion_experiment_ddp(config):
    fusion_experiment(config, main_ddp)


@experiment
def fusion_experiment_ourdist(config):
    # the native backend adds the per-tensor point (north-star config 3: "per-tensor to 64 MB")
    sizes = ([0] + fusion_test_sizes_k) if config.backend == "b200" else None
    fusion_experiment(config, main_ourdist, sizes)


def run_experiment(config) -> None:
    if config.experiment not in EXPERIMENTS:
        raise SystemExit(f"unknown experiment '{config.experiment}'; choose from: {', '.join(sorted(EXPERIMENTS))}")
    EXPERIMENTS[config.experiment](config)
