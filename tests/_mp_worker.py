"""Multi-GPU worker for test_multigpu.py; run under torch.distributed.run, one rank per GPU.
Exits non-zero on any mismatch.  Not collected by pytest (leading underscore)."""
import os
import sys

import torch
import torch.distributed as dist
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def gathered_sum(x, world, scale, dtype):
    parts = [torch.empty_like(x) for _ in range(world)]
    dist.all_gather(parts, x)
    acc = parts[0].float().clone()
    for p in parts[1:]:
        acc += p.float()
    return (acc * scale).to(dtype)


def check_allreduce(comm, rank, world, dev):
    from distributed_learning_b200.ops import fused as K

    algos = ["twoshot", "oneshot", "twoshot_bulk"] + (["nvls", "nvls_u2", "nvls_u8"] if comm.has_multicast else [])
    for dtype in (torch.float32, torch.bfloat16):
        esize = 4 if dtype == torch.float32 else 2
        for numel in (16, 4096, 100000 // 8 * 8, 1 << 22):
            for algo in algos:
                for rep in range(3):
                    g = torch.Generator(device=dev).manual_seed(rank * 1000 + rep + numel)
                    x = torch.randn(numel, device=dev, generator=g).to(dtype)
                    want = gathered_sum(x, world, 1.0 / world, dtype)
                    buf = comm.view(0, numel, dtype)
                    buf.copy_(x)
                    out_off = (numel * esize + 4095) // 4096 * 4096
                    name, _, un = algo.partition("_u")
                    comm.allreduce(0, numel, dtype, name, scale=1.0 / world, out_offset=out_off, unroll=int(un or 0))
                    res = comm.view(out_off, numel, dtype) if algo == "oneshot" else buf
                    torch.cuda.synchronize()
                    if algo.startswith("nvls"):
                        # the switch sums in fp32 in an order of its own; for bf16 it rounds the sum to
                        # bf16 before the kernel scales and rounds again: one bf16 ulp (2^-8 relative)
                        ok = torch.allclose(res.float(), want.float(), rtol=8e-3 if esize == 2 else 1e-5, atol=1e-5)
                    else:
                        ok = torch.equal(res, want)
                    if not ok:
                        err = (res.float() - want.float()).abs().max().item()
                        raise AssertionError(f"rank {rank}: {algo} {dtype} numel {numel} rep {rep}: max err {err}")
                    dist.barrier()      # nobody refills the buffer while a peer may still read it
    # fused SGD epilogue
    numel = 1 << 18
    h = K.make_hyper(0.01, 0.5)
    algos = [a for a in algos if "_u" not in a]
    for algo in algos:
        p = torch.ones(numel, device=dev)
        m = torch.zeros(numel, device=dev)
        ref_p = torch.nn.Parameter(p.clone())
        opt = torch.optim.SGD([ref_p], lr=0.01, momentum=0.5)
        for step in range(3):
            g = torch.Generator(device=dev).manual_seed(rank * 77 + step)
            x = torch.randn(numel, device=dev, generator=g)
            ref_p.grad = gathered_sum(x, world, 1.0 / world, torch.float32)
            opt.step()
            comm.view(0, numel, torch.float32).copy_(x)
            comm.allreduce(0, numel, torch.float32, algo, scale=1.0 / world, out_offset=numel * 4,
                           sgd=K.fused_sgd_args(p, m, h))
            torch.cuda.synchronize()
            dist.barrier()
        tol = 1e-5 if algo.startswith("nvls") else 1e-6
        if not torch.allclose(p, ref_p.detach(), rtol=tol, atol=tol):
            raise AssertionError(f"rank {rank}: fused sgd via {algo} diverged: {(p - ref_p).abs().max().item()}")


def check_fused_pack(comm, rank, world, dev):
    """The pack prologue inside the allreduce kernel, on real peers: equal to dlb_pack + allreduce."""
    from distributed_learning_b200.ops import fused as K

    sizes = [1, 3, 7, 8, 64, 1000, 8191, 8192, 8193, 40003, 1 << 20]
    variants = [("twoshot", 0), ("oneshot", 0), ("twoshot_bulk", 0)] + ([("nvls", 2), ("nvls", 4), ("nvls", 8)] if comm.has_multicast else [])
    for flat_dtype, grad_dtype in ((torch.float32, torch.float32), (torch.bfloat16, torch.bfloat16),
                                   (torch.bfloat16, torch.float32)):
        esize = 4 if flat_dtype == torch.float32 else 2
        step = 16 // esize
        offs, cur = [], 0
        for n in sizes:
            offs.append(cur)
            cur = (cur + n + step - 1) // step * step
        total = cur
        g = torch.Generator(device=dev).manual_seed(4000 + rank)
        grads = [torch.randn(n, device=dev, generator=g).to(grad_dtype) for n in sizes]
        grads[3] = None
        plan = K.PackPlan(sizes, offs, [grad_dtype] * len(sizes))
        out_off = (total * esize + 4095) // 4096 * 4096
        for algo, un in variants:
            for nblocks in (4, 16):
                # reference path: separate pack launch, then the allreduce
                buf = comm.view(0, total, flat_dtype)
                buf.fill_(float("nan"))
                plan.pack(buf, grads, 1.0 / world)
                comm.allreduce(0, total, flat_dtype, algo, scale=1.0, out_offset=out_off, nblocks=nblocks, unroll=un)
                torch.cuda.synchronize()
                want = (comm.view(out_off, total, flat_dtype) if algo == "oneshot" else buf).clone()
                dist.barrier()
                buf.fill_(float("nan"))
                dist.barrier()
                comm.allreduce(0, total, flat_dtype, algo, scale=1.0, out_offset=out_off, nblocks=nblocks, unroll=un,
                               pack=plan.bind(grads), pack_scale=1.0 / world)
                torch.cuda.synchronize()
                got = comm.view(out_off, total, flat_dtype) if algo == "oneshot" else buf
                if algo == "nvls":
                    ok = torch.allclose(got.float(), want.float(), rtol=8e-3 if esize == 2 else 1e-5, atol=1e-5)
                else:
                    ok = torch.equal(got, want)
                if not ok or bool(torch.isnan(got.float()).any()):
                    raise AssertionError(f"rank {rank}: fused pack {algo}/u{un}/{nblocks} {flat_dtype}<-{grad_dtype} differs "
                                         f"(max err {(got.float() - want.float()).abs().max().item()})")
                dist.barrier()


def check_broadcast(comm, rank, world, dev):
    n = (3 << 20) + 16
    buf = comm.view(0, n, torch.uint8)
    for root in (0, world - 1):
        g = torch.Generator(device=dev).manual_seed(70 + rank)
        buf.copy_(torch.randint(0, 256, (n,), device=dev, generator=g, dtype=torch.uint8))
        want = torch.randint(0, 256, (n,), device=dev, generator=torch.Generator(device=dev).manual_seed(70 + root),
                             dtype=torch.uint8)
        dist.barrier()
        comm.broadcast(0, n, root=root)
        torch.cuda.synchronize()
        if not torch.equal(buf, want):
            raise AssertionError(f"rank {rank}: broadcast from {root} differs")
        dist.barrier()


def check_training(rank, world, dev, comm, model_dtype="fp32", fused=True, overlap=True, algo="auto", graph=False,
                   fuse_pack=True):
    from distributed_learning_b200.models import CifarNet
    from distributed_learning_b200.parallel.graph_step import GraphedTrainStep
    from distributed_learning_b200.parallel.native import B200Dist

    torch.manual_seed(11)
    ref = CifarNet().to(dev).train()
    mine = CifarNet().to(dev).train()
    mine.load_state_dict(ref.state_dict())
    for mdl in (ref, mine):
        mdl.dropout1.p = 0.0
        mdl.dropout2.p = 0.0
    w = B200Dist(mine, None, 256 * 1024, dev, fused_sgd=fused, overlap=overlap, algo=algo, model_dtype=model_dtype,
                 cuda_graph=graph, fuse_pack=fuse_pack)
    opt = w.make_optimizer(lr=0.01, momentum=0.5)
    ref_opt = torch.optim.SGD(ref.parameters(), lr=0.01, momentum=0.5)
    g = torch.Generator(device=dev).manual_seed(500 + rank)
    batches = [(torch.rand(8, 3, 32, 32, device=dev, generator=g), torch.randint(10, (8,), device=dev, generator=g))
               for _ in range(5)]
    gstep = None
    for i, (x, y) in enumerate(batches):
        if graph:
            if gstep is None:       # warm-up and capture leave no trace: every batch, the first included, is one replay
                gstep = GraphedTrainStep(w, opt, x, y, warmup=1)
            gstep(x, y)
        else:
            opt.zero_grad()
            F.nll_loss(w(x).float(), y).backward()
            w.sync_gradients()
            opt.step()
        ref_opt.zero_grad()
        F.nll_loss(ref(x), y).backward()
        for p in ref.parameters():
            dist.all_reduce(p.grad)
            p.grad /= world
        ref_opt.step()
    torch.cuda.synchronize()
    tol = 3e-2 if model_dtype == "bf16" else 1e-4
    for (n, p), q in zip(mine.named_parameters(), ref.parameters()):
        if not torch.allclose(p.float(), q, rtol=tol, atol=tol):
            raise AssertionError(f"rank {rank}: parameter {n} diverged ({(p.float() - q).abs().max().item()}) "
                                 f"[{model_dtype} fused={fused} overlap={overlap} algo={algo} graph={graph} fuse_pack={fuse_pack}]")
    # replicas bit-identical across ranks
    flat = torch.cat([p.detach().float().reshape(-1) for p in mine.parameters()])
    parts = [torch.empty_like(flat) for _ in range(world)]
    dist.all_gather(parts, flat)
    if not all(torch.equal(parts[0], q) for q in parts):
        raise AssertionError("replicas differ between ranks")
    w.comm.check()
    w.cleanup()


def check_native_reducer(rank, world, dev):
    """OurDist (threads, host-driven) over NativeReducer equals manual allreduce + SGD."""
    from distributed_learning_b200.models import CifarNet
    from distributed_learning_b200.parallel.native import NativeReducer
    from distributed_learning_b200.parallel.wrappers import OurDist

    torch.manual_seed(13)
    ref = CifarNet().to(dev).train()
    mine = CifarNet().to(dev).train()
    mine.load_state_dict(ref.state_dict())
    for mdl in (ref, mine):
        mdl.dropout1.p = mdl.dropout2.p = 0.0
    reducer = NativeReducer(capacity_bytes=8 << 20, device=dev)
    w = OurDist(mine, reducer, 256 * 1024, dev)
    opt = torch.optim.SGD(mine.parameters(), lr=0.01, momentum=0.5)
    ref_opt = torch.optim.SGD(ref.parameters(), lr=0.01, momentum=0.5)
    g = torch.Generator(device=dev).manual_seed(900 + rank)
    for _ in range(3):
        x = torch.rand(8, 3, 32, 32, device=dev, generator=g)
        y = torch.randint(10, (8,), device=dev, generator=g)
        opt.zero_grad()
        F.nll_loss(w(x), y).backward()
        w.sync_gradients()
        opt.step()
        ref_opt.zero_grad()
        F.nll_loss(ref(x), y).backward()
        for p in ref.parameters():
            dist.all_reduce(p.grad)
            p.grad /= world
        ref_opt.step()
    torch.cuda.synchronize()
    for (n, p), q in zip(mine.named_parameters(), ref.parameters()):
        if not torch.allclose(p, q, rtol=1e-4, atol=1e-5):
            raise AssertionError(f"rank {rank}: OurDist+NativeReducer: parameter {n} diverged")
    w.cleanup()
    reducer.cleanup()


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    dev = torch.device(f"cuda:{local}")
    torch.cuda.set_device(dev)
    dist.init_process_group("cpu:gloo,cuda:nccl", rank=rank, world_size=world)
    from distributed_learning_b200.ops.comm import NativeComm

    what = sys.argv[1] if len(sys.argv) > 1 else "all"
    for prefer in (["vmm", "ipc"] if what in ("all", "comm") else []):
        os.environ["DLB200_ARENA"] = prefer
        comm = NativeComm(device=dev)
        try:
            comm.allocate(64 << 20)
        except RuntimeError as e:
            if rank == 0:
                print(f"arena mode {prefer} unavailable: {e}")
            continue
        if rank == 0:
            print(f"arena mode {comm.mode}, multicast {comm.has_multicast} {comm.mc_note}")
        check_allreduce(comm, rank, world, dev)
        check_fused_pack(comm, rank, world, dev)
        check_broadcast(comm, rank, world, dev)
        comm.check()
        comm.close()
    os.environ.pop("DLB200_ARENA", None)
    if what in ("all", "train"):
        for kw in (dict(), dict(fused=False), dict(overlap=False), dict(algo="oneshot"), dict(algo="twoshot"),
                   dict(model_dtype="bf16"), dict(graph=True), dict(graph=True, algo="oneshot"),
                   dict(graph=True, algo="twoshot", fused=False), dict(fuse_pack=False), dict(algo="nvls"),
                   dict(model_dtype="bf16", algo="nvls", graph=True)):
            check_training(rank, world, dev, None, **kw)
    if what in ("all", "train", "reducer"):
        check_native_reducer(rank, world, dev)
    dist.barrier()
    if rank == 0:
        print("MP_WORKER_OK")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
