"""Command line of ``main.py``: ``n r d D a i M R g`` plus options.

Parity: reference ``src/config.py:10-58``.  The nine positionals, their order and meaning,
the six reference options and their defaults, the ``envarg://`` indirection on ``n``/``r``,
the hard-coded ``epoch_count=100`` / ``grouping_size=25 MiB`` defaults and the ``devices``
list are kept.  Additions (all optional, so every reference command line still parses):

* ``M`` also accepts ``cifar`` (the same CNN on 3x32x32), ``googlenet`` (alias of
  ``imagenet``) and ``resnet50``;
* ``--backend`` also accepts ``b200`` (the native NVLink kernels of this project);
* ``--grouping_size``, ``--epoch_count``, ``--lr``, ``--momentum``, ``--dtype``,
  ``--master_port`` expose what the reference hard-codes;
* ``--checkpoint``, ``--checkpoint_every``, ``--resume`` drive the state_dict checkpoint;
* ``--results_root``, ``--algo``, ``--cuda_graph``, ``--fused_sgd``, ``--channels_last``,
  ``--comm_dtype``, ``--timer`` tune the native path.

``i`` (the NIC name) is accepted and exported as ``GLOO_SOCKET_IFNAME`` for the gloo
plumbing; the NVLink path ignores it and discovers the GPU topology instead.
"""
from __future__ import annotations

import argparse
from typing import List, Optional

from .logging import eval_arg

DEFAULT_GROUPING_SIZE = 25 * 1024 * 1024
DEFAULT_EPOCH_COUNT = 100
DEFAULT_MASTER_PORT = 29501

MODEL_TYPES = ("mnist", "cifar", "imagenet", "googlenet", "resnet50")
BACKENDS = ("gloo", "nccl", "b200")


def build_parser() -> argparse.ArgumentParser:
    p = argparse.ArgumentParser(description="Distributed machine learning benchmark tool (B200-native)")
    p.add_argument("size", metavar="n", type=str, help="number of nodes")
    p.add_argument("rank", metavar="r", type=str, help="rank (id) of this node")
    p.add_argument("node_dev", metavar="d", type=int, help="number of devices managed by this node")
    p.add_argument("total_dev", metavar="D", type=int, help="total number of devices")
    p.add_argument("master_addr", metavar="a", type=str, help="address of the master node")
    p.add_argument("ifname", metavar="i", type=str, help="network card name (eg. lo, ib0); ignored on the NVLink path")
    p.add_argument("model_type", metavar="M", type=str, help="model/dataset: " + ", ".join(MODEL_TYPES))
    p.add_argument("dataset_root", metavar="R", type=str, help="root folder for the datasets")
    p.add_argument("use_gpu", metavar="g", type=int, help="whether to use the gpu (1) or just cpu (0)")
    # reference options (src/config.py:30-42)
    p.add_argument("--job_id", dest="job_id", type=str, default="job", help="unique string used for results folder name")
    p.add_argument("--experiment", dest="experiment", type=str, default="main_ourdist", help="which experiment function to run")
    p.add_argument("--limit_batches", dest="limit_batches", type=int, default=3, help="stop after this many batches")
    p.add_argument("--backend", dest="backend", type=str, default="gloo", help="gloo | nccl | b200")
    p.add_argument("--batch_size", dest="batch_size", type=int, default=128, help="per-device batch size")
    p.add_argument("--random_input", dest="random_input", type=int, default=0,
                   help="generate random training data instead of loading the dataset")
    # additions
    p.add_argument("--grouping_size", type=int, default=DEFAULT_GROUPING_SIZE, help="fusion bucket size in bytes (0: one tensor per bucket)")
    p.add_argument("--epoch_count", type=int, default=DEFAULT_EPOCH_COUNT)
    p.add_argument("--lr", type=float, default=0.01)
    p.add_argument("--momentum", type=float, default=0.5)
    p.add_argument("--dtype", type=str, default="fp32", choices=("fp32", "bf16"), help="compute dtype of the model (bf16 keeps fp32 master weights)")
    p.add_argument("--comm_dtype", type=str, default="auto", choices=("auto", "fp32", "bf16"), help="dtype of the fusion buffers on the wire")
    p.add_argument("--master_port", type=int, default=None, help="rendezvous port (default: $MASTER_PORT or 29501)")
    p.add_argument("--results_root", type=str, default="results")
    p.add_argument("--checkpoint", type=str, default=None, help="checkpoint file (torch state_dict format)")
    p.add_argument("--checkpoint_every", type=int, default=0, help="save every N batches (0: only at the end)")
    p.add_argument("--resume", type=int, default=0, help="resume from --checkpoint if it exists")
    p.add_argument("--algo", type=str, default="auto", choices=("auto", "oneshot", "twoshot", "twoshot_bulk", "nvls", "nvls_deep"),
                   help="native allreduce variant (auto: chosen per message size from measured bus bandwidth)")
    p.add_argument("--fused_sgd", type=int, default=1, help="b200 backend: apply SGD inside/after the allreduce kernel on the flat buffers")
    p.add_argument("--fuse_pack", type=int, default=1, help="b200 backend: gather the gradients inside the allreduce kernel (one launch per bucket) instead of a separate pack launch")
    p.add_argument("--cuda_graph", type=int, default=0, help="b200 backend: capture the whole step in a CUDA graph")
    p.add_argument("--channels_last", type=int, default=0, help="NHWC weights and activations on the GPU")
    p.add_argument("--cudnn_benchmark", type=int, default=0, help="let cuDNN time its algorithms at the first step")
    p.add_argument("--timer", type=str, default="auto", choices=("auto", "host", "device"))
    p.add_argument("--sync_loss", type=int, default=None,
                   help="read the loss back every batch like the reference (default: 1 on cpu, 0 on gpu where losses are flushed at the end)")
    p.add_argument("--seed", type=int, default=1234)
    p.add_argument("--dynamic_unused", type=int, default=0,
                   help="re-detect parameters without gradient after every forward, like the reference "
                        "(default: detect once at the first step and leave them out)")
    p.add_argument("--nvtx", type=int, default=0, help="emit NVTX ranges around the phases and the per-bucket launches")
    return p


def finalize_config(config: argparse.Namespace) -> argparse.Namespace:
    """Everything ``parse_args`` derives after argparse (reference ``src/config.py:44-56``)."""
    config.size = int(eval_arg(str(config.size)))
    config.rank = int(eval_arg(str(config.rank)))
    if config.model_type not in MODEL_TYPES:
        raise SystemExit(f"unknown model type '{config.model_type}' (choose from {', '.join(MODEL_TYPES)})")
    if config.backend not in BACKENDS:
        raise SystemExit(f"unknown backend '{config.backend}' (choose from {', '.join(BACKENDS)})")
    # late imports keep `--help` fast and avoid importing torch for argument errors
    from ..models import create_network
    from .data import get_partition_dataset

    config.create_network = create_network[config.model_type]
    config.get_partition_dataset = get_partition_dataset[config.model_type]
    if config.use_gpu == 1:
        config.devices = [f"cuda:{i}" for i in range(config.node_dev)]
    else:
        config.devices = ["cpu"] * config.node_dev
    if config.sync_loss is None:
        config.sync_loss = 0 if config.use_gpu == 1 else 1
    if config.backend == "b200" and config.use_gpu != 1:
        raise SystemExit("--backend b200 is the NVLink path: it needs g=1 (use --backend gloo on the CPU)")
    if getattr(config, "cuda_graph", 0) and int(config.random_input) in (3, 4):
        # the prefetcher's staging thread pins host memory while the step is being captured, which
        # a global-mode capture does not allow; bench.py shows the supported order (capture first)
        raise SystemExit("--cuda_graph 1 cannot be combined with --random_input 3 in main.py; use --random_input 1 or 2")
    if getattr(config, "cuda_graph", 0) and getattr(config, "dynamic_unused", 0):
        raise SystemExit("--dynamic_unused 1 walks the autograd graph every step and cannot be captured: drop --cuda_graph")
    if config.backend == "b200" and config.total_dev > 8:
        raise SystemExit("--backend b200 spans one NVSwitch domain (at most 8 GPUs); use --backend nccl beyond that")
    if config.size * config.node_dev != config.total_dev:
        # the reference does not check this and then mis-partitions the data silently
        raise SystemExit(f"n*d must equal D (got {config.size}*{config.node_dev} != {config.total_dev})")
    return config


def parse_args(argv: Optional[List[str]] = None) -> argparse.Namespace:
    return finalize_config(build_parser().parse_args(argv))


def picklable_config(config: argparse.Namespace) -> argparse.Namespace:
    """A copy that can cross a ``spawn`` boundary and be stored in a checkpoint."""
    skip = {"create_network", "get_partition_dataset", "distribute_model"}
    return argparse.Namespace(**{k: v for k, v in vars(config).items() if k not in skip and not callable(v)})
