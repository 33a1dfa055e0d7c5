"""Process-group bring-up and the ``main.py`` entry point.

Parity: reference ``init_process`` (``src/main.py:303-318``) and ``__main__`` (``:321-333``):
results folder ``{results_root}/{experiment}_{D}_{job_id}``, ``spawn`` start method, one
``torch.distributed`` rank per *node* on ``MASTER_ADDR:29501``, ``GLOO_SOCKET_IFNAME`` from the
``i`` argument, device pinning ``rank % device_count`` when ``d == 1``, dispatch by
experiment name.

Differences:

* the port can be overridden (``--master_port`` or ``$MASTER_PORT``; Q17), so two jobs can
  share a host and ``torchrun``'s port is honoured;
* the device backends (``nccl``, ``b200``) run one process per GPU.  With ``d == 1`` the node
  group is that group: ``cpu:gloo,cuda:nccl`` for the ``nccl`` baseline, plain ``gloo`` for ``b200``
  (bootstrap, handle exchange and host barriers; no NCCL communicator exists on the product path).
  With ``d > 1`` each spawned worker joins a second, worker-level group of ``D`` ranks on
  ``port + 1`` (``init_worker_group``);
* on the ``b200`` backend bring-up ends with GPU/NVLink *topology discovery* instead of a NIC
  name: peer-access matrix, NVLink link states and multicast capability are queried and
  logged (``distributed_learning_b200.parallel.topology``);
* the reference's pairwise ``pgroups`` (two-rank groups used to fake NCCL point-to-point,
  ``:312-316``) are not needed: ``ring_allreduce_gpu`` uses real send/recv.
"""
from __future__ import annotations

import datetime
import os
from typing import List, Optional

import torch
import torch.distributed as dist

from .utils.config import DEFAULT_MASTER_PORT, parse_args
from .utils.logging import Level, print_d


def master_port(config) -> int:
    if getattr(config, "master_port", None):
        return int(config.master_port)
    return int(os.environ.get("DLB200_MASTER_PORT", os.environ.get("MASTER_PORT", DEFAULT_MASTER_PORT)))


def _group_backend(config, one_process_per_gpu: bool) -> str:
    if config.backend == "b200":
        # bootstrap, handle exchange and host barriers only: the product path never creates an NCCL
        # communicator (main_ddp inside a b200 suite makes its own NCCL sub-group when it runs)
        return "gloo"
    if config.backend == "nccl" and config.use_gpu == 1 and one_process_per_gpu:
        return "cpu:gloo,cuda:nccl"
    return config.backend if config.backend != "nccl" or config.use_gpu == 1 else "gloo"


def init_process(config) -> None:
    """Initialise the node-level process group (one rank per node)."""
    print_d(config, Level.DEBUG)
    port = master_port(config)
    if os.environ.get("TORCHELASTIC_USE_AGENT_STORE") and str(port) != os.environ.get("MASTER_PORT"):
        # under torchrun the env:// rendezvous expects the agent's store on torchrun's port; on any
        # other port rank 0 has to host the store itself, otherwise every rank waits for a server
        os.environ.pop("TORCHELASTIC_USE_AGENT_STORE")
    os.environ["MASTER_ADDR"] = config.master_addr
    os.environ["MASTER_PORT"] = str(port)
    os.environ.setdefault("GLOO_SOCKET_IFNAME", config.ifname)
    backend = _group_backend(config, one_process_per_gpu=config.node_dev == 1)
    dist.init_process_group(backend, rank=config.rank, world_size=config.size,
                            timeout=datetime.timedelta(seconds=int(os.environ.get("DLB200_PG_TIMEOUT", 600))))
    print_d(f"Connection initialised (backend {backend}, rank {config.rank}/{config.size})", Level.INFO)


def init_worker_group(config, node_id: int, worker_id: int) -> None:
    """Join the worker-level group (``D`` ranks, one per GPU) from a spawned worker."""
    rank = node_id * config.node_dev + worker_id
    os.environ["MASTER_ADDR"] = config.master_addr
    os.environ["MASTER_PORT"] = str(master_port(config) + 1)
    os.environ.setdefault("GLOO_SOCKET_IFNAME", config.ifname)
    if config.use_gpu == 1:
        torch.cuda.set_device(worker_id)
    dist.init_process_group(_group_backend(config, True), rank=rank, world_size=config.total_dev,
                            timeout=datetime.timedelta(seconds=int(os.environ.get("DLB200_PG_TIMEOUT", 600))))


def main(argv: Optional[List[str]] = None) -> int:
    from .experiments import run_experiment

    config = parse_args(argv)
    print_d(f"Number of available devices {torch.cuda.device_count()}", Level.INFO)
    config.folder = f"{config.results_root}/{config.experiment}_{config.total_dev}_{config.job_id}"
    os.makedirs(config.folder, exist_ok=True)
    try:
        torch.multiprocessing.set_start_method("spawn")
    except RuntimeError:
        pass
    init_process(config)

    if config.node_dev == 1 and config.devices[0].startswith("cuda"):
        config.device_id = config.rank % torch.cuda.device_count()
        config.devices = [f"cuda:{config.device_id}" for _ in config.devices]
        torch.cuda.set_device(config.device_id)

    if config.backend == "b200" and config.use_gpu == 1 and config.rank == 0:
        from .parallel.topology import describe_topology

        print_d(describe_topology(), Level.INFO)
    try:
        run_experiment(config)
    finally:
        if dist.is_initialized():
            dist.destroy_process_group()
    return 0
