"""Spawned gloo workers for test_distributed_cpu.py (not collected by pytest)."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def _init(rank, world, port):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)


def reference_ring(parts, rank, size):
    """Plain restatement of the reference ring's arithmetic (src/allreduce.py:45-98) on all ranks'
    data at once: chunk c is accumulated around the ring starting at rank c, then divided."""
    n = parts[0].numel()
    step = -(-n // size)
    out = torch.empty_like(parts[0])
    for c in range(size):
        lo, hi = min(c * step, n), min((c + 1) * step, n)
        acc = parts[c][lo:hi].clone()
        for k in range(1, size):
            acc = parts[(c + k) % size][lo:hi] + acc      # receiver adds the incoming partial sum to its own
        out[lo:hi] = acc
    return out / float(size)


def allreduce_worker(rank, world, port, result_dir):
    from distributed_learning_b200.parallel.allreduce import (built_in_allreduce, central_allreduce, chunk_bounds,
                                                               ring_allreduce)
    _init(rank, world, port)
    assert chunk_bounds(10, 4) == [(0, 3), (3, 6), (6, 9), (9, 10)]
    assert chunk_bounds(2, 4) == [(0, 1), (1, 2), (2, 2), (2, 2)]
    for numel in (1, 2, 5, 1000, 100003):
        torch.manual_seed(numel)
        parts = [torch.randn(numel) for _ in range(world)]
        mine = parts[rank]
        for name, fn in (("ring", ring_allreduce), ("central", central_allreduce), ("builtin", built_in_allreduce)):
            x = mine.clone()
            fn(x)
            if name == "ring":
                assert torch.equal(x, reference_ring(parts, rank, world)), (name, numel)
            elif name == "central":
                acc = parts[0].clone()
                for p in parts[1:]:
                    acc += p
                assert torch.equal(x, acc / float(world)), (name, numel)
            else:
                assert torch.allclose(x, torch.stack(parts).mean(0), atol=1e-6), (name, numel)
    open(os.path.join(result_dir, f"ok{rank}"), "w").close()
    dist.destroy_process_group()


def wrapper_worker(rank, world, port, result_dir):
    """All wrapper x reducer combinations give the same parameters as averaged-gradient SGD."""
    import torch.nn.functional as F

    from distributed_learning_b200.models import CifarNet
    from distributed_learning_b200.parallel.allreduce import central_allreduce, ring_allreduce
    from distributed_learning_b200.parallel.reducers import ReduceImmediately, StreamReducer
    from distributed_learning_b200.parallel.wrappers import DDPDist, OurDist, SeqDist, SeqMergeDist

    _init(rank, world, port)

    def train(make):
        torch.manual_seed(3)
        net = CifarNet()
        net.dropout1.p = net.dropout2.p = 0.0
        w = make(net)
        params = list(net.parameters())
        opt = torch.optim.SGD(params, lr=0.01, momentum=0.5)
        g = torch.Generator().manual_seed(50 + rank)
        for _ in range(3):
            x, y = torch.rand(4, 3, 32, 32, generator=g), torch.randint(10, (4,), generator=g)
            opt.zero_grad()
            F.nll_loss(w(x), y).backward()
            w.sync_gradients()
            opt.step()
        w.cleanup()
        return torch.cat([p.detach().reshape(-1) for p in params])

    class Manual:
        def __init__(self, net):
            self.net = net

        def __call__(self, x):
            return self.net(x)

        def sync_gradients(self):
            for p in self.net.parameters():
                dist.all_reduce(p.grad)
                p.grad /= world

        def cleanup(self):
            pass

    want = train(Manual)
    combos = {
        "ourdist_ring": lambda n: OurDist(n, ReduceImmediately(ring_allreduce), 256 * 1024),
        "ourdist_per_tensor": lambda n: OurDist(n, ReduceImmediately(ring_allreduce), 0),
        "ourdist_central": lambda n: OurDist(n, ReduceImmediately(central_allreduce), 25 << 20),
        "seqmerge": lambda n: SeqMergeDist(n, ReduceImmediately(ring_allreduce), 1 << 20),
        "seq": lambda n: SeqDist(n, ReduceImmediately(ring_allreduce)),
        "ourdist_stream": lambda n: OurDist(n, StreamReducer(), 1 << 20),
        "ddp": lambda n: DDPDist(n, None, 1 << 20),
    }
    for name, make in combos.items():
        got = train(make)
        assert torch.allclose(got, want, rtol=1e-5, atol=1e-6), name
    open(os.path.join(result_dir, f"ok{rank}"), "w").close()
    dist.destroy_process_group()


def pump_worker_child(reducer, rank, values):
    """A 'device worker' of the 2-step reducer: sends buffers, checks the averaged results."""
    for v in values:
        reducer.put(torch.full((6,), float(v + rank)))
        out = reducer.get()
        assert torch.allclose(out, torch.full((6,), v + 0.5)), (out, v)
    reducer.cleanup()
