"""The reference arm of bench.py: the UNMODIFIED reference (baseline/_ref/src/main.py) run through
its own command line, one process per GPU, for the same model, batch and step count.

Nothing of this framework is on that path: this module only fabricates the dataset the reference
insists on opening (it takes one real batch for shape even with --random_input 1,
src/config.py:41-42), points torch.hub at an offline stand-in for the pytorch/vision checkout
(baseline/hub_cache), launches `python main.py n r 1 N 127.0.0.1 lo M R 1 --experiment main_ddp
--backend nccl ...` and reads the reference's own per-batch timing CSV (src/timing.py).  The
`batch` column is host wall-clock around the whole step; the reference formats the loss into a
string every step, which synchronises the device inside that span, so it is a true step time.
"""
from __future__ import annotations

import csv
import os
import shutil
import subprocess
import sys
import tempfile

REF_MODEL = {"imagenet": "imagenet", "mnist": "mnist"}


def _batch_bytes(model_type: str, bs: int) -> int:
    """Bytes the reference copies host->device per step: fp32 images + int64 labels."""
    c, h, w = (3, 224, 224) if model_type == "imagenet" else (1, 28, 28)
    return bs * c * h * w * 4 + bs * 8


def _unavailable(why: str) -> dict:
    return {"impl": "reference", "unavailable": why}


def run(args, helpers) -> dict:
    root = helpers.ROOT
    main_py = os.path.join(root, "baseline", "_ref", "src", "main.py")
    if not os.path.exists(main_py):
        src = "/root/reference/src"
        if os.path.isdir(src):          # the reference has no setup.py: "installing" it is a verbatim copy
            os.makedirs(os.path.dirname(os.path.dirname(main_py)), exist_ok=True)
            shutil.copytree(src, os.path.dirname(main_py))
        else:
            return _unavailable("baseline/_ref/src/main.py is missing and /root/reference is not mounted")
    model_type = helpers.MODEL_ALIASES[args.model]
    if model_type not in REF_MODEL:
        return _unavailable(f"the reference has no '{args.model}' model (only mnist and imagenet/GoogLeNet)")
    if args.dtype == "bf16":
        return _unavailable("the reference trains in fp32 only")

    rank, local_rank, world = helpers.dist_env()
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE is {world}: launch N>1 with torch.distributed.run")
    sys.path.insert(0, root)
    from distributed_learning_b200.utils.data import write_fake_image_folder, write_fake_mnist

    work = tempfile.mkdtemp(prefix=f"dlb200_ref_{rank}_")
    data_root = os.path.join(work, "data")
    bs = args.batch_size
    if model_type == "imagenet":
        write_fake_image_folder(data_root, n=bs * world, classes=8, size=32)      # one batch per device is enough
    else:
        write_fake_mnist(data_root, n=max(2048, bs * world))
    total = args.warmup + args.steps
    env = dict(os.environ)
    env["TORCH_HOME"] = os.path.join(root, "baseline", "hub_cache")
    env.pop("MASTER_PORT", None)        # the reference sets its own (29501)
    # The reference does its own env:// rendezvous on that port.  Under torchrun the environment
    # says "an agent already runs the store" (TORCHELASTIC_USE_AGENT_STORE), which would make rank 0
    # connect to port 29501 as a client instead of serving it, and every rank would wait forever.
    for k in [k for k in env if k.startswith(("TORCHELASTIC_", "TORCH_NCCL_ASYNC", "ROLE_", "GROUP_"))] + \
            ["RANK", "WORLD_SIZE", "LOCAL_WORLD_SIZE", "OMP_NUM_THREADS"]:
        env.pop(k, None)
    cmd = [sys.executable, main_py, str(world), str(rank), "1", str(world), "127.0.0.1", "lo", REF_MODEL[model_type],
           data_root, str(int(getattr(args, "ref_use_gpu", 1))), "--experiment", args.ref_experiment, "--backend", args.ref_backend, "--random_input", "1",
           "--limit_batches", str(total), "--batch_size", str(bs), "--job_id", "bench"]
    clocks = helpers.ClockSampler(local_rank).start()
    log = os.path.join(work, "ref.log")
    with open(log, "w") as lf:
        rc = subprocess.run(cmd, cwd=work, env=env, stdout=lf, stderr=subprocess.STDOUT).returncode
    clock_stats = clocks.stop()
    if rc != 0:
        tail = open(log).read()[-600:].replace("\n", " | ")
        return _unavailable(f"reference exited {rc}: {tail}")
    name = {"main_ddp": "ddp", "main_onestep_reduce": "onestep_reduce", "main_single": "single",
            "main_onestep_central": "onestep_central"}.get(args.ref_experiment, args.ref_experiment.replace("main_", ""))
    node = 0 if args.ref_experiment == "main_single" else rank
    path = os.path.join(work, "results", f"{args.ref_experiment}_{world}_bench", f"{name}_{node}_0_times.csv")
    if not os.path.exists(path):
        return _unavailable(f"reference wrote no timing file {path}")
    with open(path) as f:
        rows = list(csv.reader(f, skipinitialspace=True))
    header, body = rows[0], rows[1:]
    col = header.index("batch")
    phases = {h: [float(r[i]) for r in body[args.warmup:]] for i, h in enumerate(header)
              if h not in ("experiment_name", "batch_count", "data_len")}
    timed = [float(r[col]) for r in body[args.warmup:]]
    if len(timed) != args.steps:
        return _unavailable(f"reference produced {len(timed)} timed rows, expected {args.steps}")
    ms_total = sum(timed)

    # max over ranks: exchange through a file-free gloo group on torchrun's own port
    if world > 1:
        import torch
        import torch.distributed as dist

        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("gloo", rank=rank, world_size=world)
        t = torch.tensor([ms_total], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total = float(t[0])
        dist.destroy_process_group()
    images = bs * world * args.steps
    value = images / (ms_total / 1000.0)
    shutil.rmtree(work, ignore_errors=True)
    return {
        "metric": helpers.METRIC,
        "value": value, "unit": "images/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": (value / helpers.PUBLISHED_DDP[world]) if (model_type == "imagenet" and bs == 128 and world in helpers.PUBLISHED_DDP) else None,
        "vs_baseline_is": "value / BASELINE.md section 2 (reference DDP on P100 + gloo, published CSVs)",
        "dtype": "fp32", "data": "synthetic (the reference's --random_input 1: host RNG every step)",
        "impl": "reference",
        "config": {"model": "googlenet(aux heads)" if model_type == "imagenet" else "basicnet-mnist",
                   "global_batch": bs * world, "per_gpu_batch": bs, "seq_len": None, "parallelism": f"dp{world}",
                   "experiment": args.ref_experiment, "backend": args.ref_backend,
                   "timing": "reference src/timing.py 'batch' column (host wall clock incl. its per-step loss read-back)",
                   "phase_means_ms": {k: sum(v) / len(v) for k, v in phases.items() if v}},
        "clocks": clock_stats,
        # the reference's step IS end to end: its timed span holds the host RNG (src/data.py:132), the
        # data.to(device) / target.to(device) copies (src/main.py:65) and the loss formatted into a
        # string, which reads the scalar back (src/main.py:94)
        "e2e": {"value": value, "unit": "images/s", "ms_per_step": ms_total / args.steps,
                "h2d_bytes_per_step": _batch_bytes(model_type, bs), "d2h_bytes_per_step": 4,
                "input": "fresh batch every step: torch.rand_like / randint_like on the host (pageable) -> .to(device)",
                "note": "same span as `value`: the reference has no device-only timer"},
        "gpu_launches": None,
    }
