"""Data-parallel machinery: buckets, wrappers, reducers, allreduce algorithms, topology and
the native NVLink communicator."""
from .buckets import Bucket, find_unused_parameters, fusion_grouping, make_buckets  # noqa: F401
from .reducers import (NodeAggregateReducer, NodeAggregateReducerCPU, ReduceImmediately,  # noqa: F401
                       StreamReducer)
from .wrappers import DDPDist, OurDist, SeqDist, SeqMergeDist, SingleDist, WarmupDist  # noqa: F401
