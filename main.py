#!/usr/bin/env python3
"""``main.py n r d D a i M R g [options]`` -- same command line as the reference's ``src/main.py``.

Examples (single host):

    # CPU plumbing, two "nodes" on one machine (run both lines)
    python main.py 2 0 1 2 127.0.0.1 lo mnist /tmp/data 0 --experiment experiment3 --limit_batches 6 --batch_size 32
    python main.py 2 1 1 2 127.0.0.1 lo mnist /tmp/data 0 --experiment experiment3 --limit_batches 6 --batch_size 32

    # 8 x B200, native NVLink kernels, one process per GPU spawned from here
    python main.py 1 0 8 8 127.0.0.1 lo imagenet /data 1 --backend b200 --experiment main_ourdist --random_input 1 --limit_batches 50

    # the same under torchrun (one rank per GPU)
    torchrun --nproc-per-node 8 main.py envarg://WORLD_SIZE envarg://RANK 1 8 127.0.0.1 lo imagenet /data 1 --backend b200
"""
import sys

from distributed_learning_b200.launcher import main

if __name__ == "__main__":
    sys.exit(main())
