"""CLI/config, logging, per-phase timing, data pipeline and checkpointing."""
from .logging import Level, eval_arg, print_d  # noqa: F401
