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

    algos = ["twoshot", "oneshot"] + (["nvls"] if comm.has_multicast else [])
    if comm.has_multicast and os.environ.get("DLB200_EXPERIMENTAL"):
        algos.append("nvls_deep")
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
                    comm.allreduce(0, numel, dtype, algo, scale=1.0 / world, out_offset=out_off)
                    res = comm.view(out_off, numel, dtype) if algo == "oneshot" else buf
                    torch.cuda.synchronize()
                    if algo.startswith("nvls"):
                        ok = torch.allclose(res.float(), want.float(), rtol=2e-2 if esize == 2 else 1e-5, atol=1e-2 if esize == 2 else 1e-5)
                    else:
                        ok = torch.equal(res, want)
                    if not ok:
                        err = (res.float() - want.float()).abs().max().item()
                        raise AssertionError(f"rank {rank}: {algo} {dtype} numel {numel} rep {rep}: max err {err}")
                    dist.barrier()      # nobody refills the buffer while a peer may still read it
    # fused SGD epilogue
    numel = 1 << 18
    h = K.make_hyper(0.01, 0.5)
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


def check_training(rank, world, dev, comm, model_dtype="fp32", fused=True, overlap=True, algo="auto", graph=False):
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
                 cuda_graph=graph)
    opt = w.make_optimizer(lr=0.01, momentum=0.5)
    ref_opt = torch.optim.SGD(ref.parameters(), lr=0.01, momentum=0.5)
    g = torch.Generator(device=dev).manual_seed(500 + rank)
    batches = [(torch.rand(8, 3, 32, 32, device=dev, generator=g), torch.randint(10, (8,), device=dev, generator=g))
               for _ in range(5)]
    gstep = None
    for i, (x, y) in enumerate(batches):
        if graph:
            if gstep is None:       # one eager warm-up step on batch 0, then capture; later batches are replays
                gstep = GraphedTrainStep(w, opt, x, y, warmup=1)
            else:
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
                                 f"[{model_dtype} fused={fused} overlap={overlap} algo={algo} graph={graph}]")
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
        comm.check()
        comm.close()
    os.environ.pop("DLB200_ARENA", None)
    if what in ("all", "train"):
        for kw in (dict(), dict(fused=False), dict(overlap=False), dict(algo="oneshot"), dict(algo="twoshot"),
                   dict(model_dtype="bf16"), dict(graph=True), dict(graph=True, algo="oneshot"),
                   dict(graph=True, algo="twoshot", fused=False)):
            check_training(rank, world, dev, None, **kw)
    if what in ("all", "train", "reducer"):
        check_native_reducer(rank, world, dev)
    dist.barrier()
    if rank == 0:
        print("MP_WORKER_OK")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
