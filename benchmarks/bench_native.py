"""The b200 arm of bench.py: B200Dist + FusedSGD on N GPUs, device-timed and end to end."""
from __future__ import annotations

import os

import torch
import torch.distributed as dist
import torch.nn.functional as F


def _make_step(wrapper, opt):
    def step(data, target):
        opt.zero_grad()
        out = wrapper(data)
        loss = F.nll_loss(out.float() if out.dtype != torch.float32 else out, target)
        loss.backward()
        wrapper.sync_gradients()
        opt.step()
        return loss
    return step


def _timed(batches, k, device, clocks=None, comm=None):
    """Exactly k steps between two events, barrier + synchronize on both sides; returns ms.

    The host barrier goes over gloo (no NCCL communicator exists in this arm); the device-side
    barrier kernel of the native communicator then lines the GPUs up, so the first event of every
    rank is recorded within a few microseconds of its peers'."""
    def barrier():
        torch.cuda.synchronize(device)
        if dist.is_initialized():
            dist.barrier()
            if comm is not None and comm.mode is not None and comm.world > 1:
                comm.barrier()
        torch.cuda.synchronize(device)

    barrier()
    if clocks is not None:
        clocks.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(k):
        batches(i)
    e1.record()
    barrier()
    stats = clocks.stop() if clocks is not None else None
    return e0.elapsed_time(e1), stats


def run(args, helpers) -> dict:
    from distributed_learning_b200.models import create_network
    from distributed_learning_b200.parallel.native import B200Dist
    from distributed_learning_b200.utils.data import INPUT_SPECS, NativeHostRandom, PinnedPrefetcher

    rank, local_rank, world = helpers.dist_env()
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE is {world}: launch N>1 with torch.distributed.run")
    device = torch.device(f"cuda:{local_rank}")
    torch.cuda.set_device(device)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # gloo only: it carries the bootstrap (handle exchange, host barriers, the max over ranks);
        # everything on the GPUs goes through this framework's own kernels, NCCL is never initialised
        dist.init_process_group("gloo", rank=rank, world_size=world)

    model_type = helpers.MODEL_ALIASES[args.model]
    dtype = args.dtype or ("bf16" if model_type == "resnet50" else "fp32")
    torch.manual_seed(1234)
    torch.backends.cudnn.benchmark = bool(args.cudnn_benchmark)
    model = create_network[model_type]().to(device)
    if args.channels_last:
        model = model.to(memory_format=torch.channels_last)
    model.train()
    wrapper = B200Dist(model, None, args.grouping_size, device, overlap=bool(args.overlap), algo=args.algo,
                       fused_sgd=bool(args.fused_sgd), fuse_pack=bool(args.fuse_pack), model_dtype=dtype, tune=bool(args.tune),
                       cuda_graph=bool(args.cuda_graph))
    opt = wrapper.make_optimizer(lr=0.01, momentum=0.5)
    step = _make_step(wrapper, opt)

    shape, classes = INPUT_SPECS[model_type]
    bs = args.batch_size
    in_dtype = torch.bfloat16 if dtype == "bf16" else torch.float32
    fmt = torch.channels_last if args.channels_last else torch.contiguous_format
    # device-resident pool, larger than the 126 MB L2, rotated every step
    in_bytes = bs * shape[0] * shape[1] * shape[2] * (2 if dtype == "bf16" else 4)
    pool_n = max(2, -(-(160 << 20) // in_bytes))
    pool_n = min(pool_n, 64)
    g = torch.Generator(device=device).manual_seed(1000 + rank)
    pool = [(torch.rand((bs,) + shape, device=device, generator=g).to(in_dtype).contiguous(memory_format=fmt),
             torch.randint(classes, (bs,), device=device, generator=g)) for _ in range(pool_n)]

    if args.cuda_graph:
        from distributed_learning_b200.parallel.graph_step import GraphedTrainStep

        step = GraphedTrainStep(wrapper, opt, pool[0][0], pool[0][1])      # one replay = one whole step
    for i in range(args.warmup):
        step(*pool[i % pool_n])
    launches0 = wrapper.kernel_launches
    ms, clocks = _timed(lambda i: step(*pool[i % pool_n]), args.steps, device, helpers.ClockSampler(local_rank),
                        comm=wrapper.comm)
    launches = wrapper.kernel_launches - launches0
    ms = helpers.max_over_ranks(ms)
    images = bs * world * args.steps
    value = images / (ms / 1000.0)

    e2e = None
    if not args.no_e2e:
        # End to end through the public API, doing the job the reference's step does: every step a
        # FRESH random batch is produced on the host (native multi-threaded generator writing the
        # pinned slot, csrc/hostgen.cpp), copied host->device by PinnedPrefetcher, trained on, and
        # the loss is read back to the host.
        src = NativeHostRandom((bs,) + shape, classes, seed=4321 + rank, dtype=torch.float32)
        pref = PinnedPrefetcher(src, device, depth=3, data_dtype=in_dtype, channels_last=bool(args.channels_last))
        losses = []

        def e2e_step(_i):
            data, target = next(pref)
            losses.append(step(data, target).item())

        for i in range(3):
            e2e_step(i)
        ms2, clocks2 = _timed(e2e_step, args.steps, device, helpers.ClockSampler(local_rank), comm=wrapper.comm)
        ms2 = helpers.max_over_ranks(ms2)
        pref.close()
        e2e = {"value": images / (ms2 / 1000.0), "unit": "images/s", "ms_per_step": ms2 / args.steps,
               "h2d_bytes_per_step": pref.h2d_bytes_per_step, "d2h_bytes_per_step": 4,
               "input": f"fresh batch every step: native host generator ({src.nthreads} threads) -> pinned slot -> H2D",
               "last_loss": losses[-1] if losses else None, "clocks": clocks2}

        # second, labelled number: the same loop with the batch generated on the device
        # (--random_input 1 of this framework); no H2D copy, so it is NOT the end-to-end figure
        gd = torch.Generator(device=device).manual_seed(99 + rank)

        def dev_step(_i):
            data = torch.rand((bs,) + shape, device=device, generator=gd).to(in_dtype).contiguous(memory_format=fmt)
            target = torch.randint(classes, (bs,), device=device, generator=gd)
            losses.append(step(data, target).item())

        for i in range(2):
            dev_step(i)
        ms3, _ = _timed(dev_step, args.steps, device, comm=wrapper.comm)
        ms3 = helpers.max_over_ranks(ms3)
        e2e["device_rng_variant"] = {"value": images / (ms3 / 1000.0), "unit": "images/s",
                                     "ms_per_step": ms3 / args.steps, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 4,
                                     "input": "fresh batch every step generated on the device (torch.rand), loss read back"}

    wrapper.comm.check()
    buckets = [{"mib": round(nb.numel * nb.regions[0].element_size() / 2**20, 2), "algo": nb.algo, "nblocks": nb.nblocks,
                "unroll": nb.unroll} for nb in wrapper.buckets]
    result = {
        "metric": helpers.METRIC,
        "value": value, "unit": "images/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
        "dtype": dtype, "data": "synthetic (random images of the named shape, random-init weights)",
        "impl": "b200",
        **helpers.baseline_fields(value, e2e["value"] if e2e else None, model_type, bs, world, dtype),
        "config": {"model": {"imagenet": "googlenet(aux heads)", "resnet50": "resnet50", "mnist": "basicnet-mnist",
                             "cifar": "basicnet-cifar"}[model_type],
                   "global_batch": bs * world, "per_gpu_batch": bs, "seq_len": None,
                   "input": list(shape), "parallelism": f"dp{world}", "optimizer": "sgd lr=0.01 momentum=0.5",
                   "grouping_size": args.grouping_size, "buckets": buckets, "fused_sgd": bool(args.fused_sgd), "fuse_pack": bool(wrapper.fuse_pack),
                   "overlap": bool(args.overlap), "arena": wrapper.comm.mode, "multicast": wrapper.comm.has_multicast,
                   "multicast_note": wrapper.comm.mc_note,
                   "channels_last": bool(args.channels_last), "cudnn_benchmark": bool(args.cudnn_benchmark),
                   "cuda_graph": bool(args.cuda_graph),
                   "timing": "CUDA events on the compute stream around exactly K steps, device-side + host barrier "
                             "and synchronize on both sides, max over ranks",
                   "process_group": "gloo (bootstrap and host barriers only; no NCCL communicator)",
                   "l2": f"inputs rotate over a {pool_n}-batch device pool ({pool_n * in_bytes >> 20} MiB > 126 MiB L2)"},
        "clocks": clocks, "e2e": e2e, "gpu_launches": launches,
    }
    wrapper.cleanup()
    if world > 1:
        dist.destroy_process_group()
    return result
