"""Levelled printing and ``envarg://`` argument indirection.

Parity: reference ``src/utils.py:1-24`` (``VERBOSE_LEVEL``, ``Level``, ``print_d``,
``eval_arg``).  Differences: the verbosity can be set from the environment
(``DLB200_VERBOSE``) so that benchmark runs are not dominated by stdout traffic, and
``eval_arg`` reports which variable is missing instead of raising a bare ``KeyError``.
"""
from __future__ import annotations

import enum
import os
import sys

ENVARG_PREFIX = "envarg://"


class Level(enum.IntEnum):
    WARNING = 1
    INFO = 2
    DEBUG = 3


def _initial_level() -> int:
    raw = os.environ.get("DLB200_VERBOSE")
    if raw is None:
        return int(Level.INFO)
    try:
        return int(raw)
    except ValueError:
        return int(Level[raw.upper()])


VERBOSE_LEVEL = _initial_level()


def set_verbosity(level: int) -> None:
    global VERBOSE_LEVEL
    VERBOSE_LEVEL = int(level)


def print_d(message, level: int = Level.INFO) -> None:
    """Print ``message`` when ``level`` is at or below the current verbosity; flushes."""
    if int(level) <= VERBOSE_LEVEL:
        print(message)
        sys.stdout.flush()


def eval_arg(arg: str) -> str:
    """Resolve ``envarg://NAME`` to ``os.environ['NAME']``; anything else is returned as is.

    This is how an MPI launcher hands rank and size to the CLI
    (``envarg://OMPI_COMM_WORLD_RANK``); under ``torchrun`` use ``envarg://RANK``.
    """
    if isinstance(arg, str) and arg.startswith(ENVARG_PREFIX):
        name = arg[len(ENVARG_PREFIX):]
        if name not in os.environ:
            raise KeyError(f"argument '{arg}' refers to environment variable '{name}', which is not set")
        return os.environ[name]
    return arg
