"""Build ``libdlb200.so`` in-tree with nvcc for sm_100a.

``python -m distributed_learning_b200.ops.build`` or ``build_library()``.  The library is a plain
C-ABI shared object (``include/dlb200.h``) loaded with ctypes, so it does not depend on the
torch C++ ABI and the same file serves Python, C and C++ callers.  It is built next to this
file (``ops/libdlb200.so``), which is git-ignored but travels to the GPU box with the tree.
cudart is linked statically and the driver API is resolved at run time, so the library loads
on a machine without a GPU or driver.
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
from typing import List, Optional

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), "..", ".."))
CSRC = os.path.join(ROOT, "csrc")
INCLUDE = os.path.join(ROOT, "include")
LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "libdlb200.so")
STAMP_PATH = LIB_PATH + ".stamp"

SOURCES = ["runtime.cpp", "hostgen.cpp", "pack.cu", "sgd.cu", "allreduce.cu"]
ARCH_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a"]


def nvcc_path() -> str:
    cand = os.environ.get("NVCC") or shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(cand):
        raise RuntimeError("nvcc not found (set $NVCC)")
    return cand


def _sources() -> List[str]:
    extra = sorted(f for f in os.listdir(CSRC) if f.endswith((".cu", ".cpp")) and f not in SOURCES)
    return [os.path.join(CSRC, f) for f in SOURCES + extra]


def _fingerprint() -> str:
    h = hashlib.sha256()
    for d in (CSRC, INCLUDE):
        for f in sorted(os.listdir(d)):
            if f.endswith((".cu", ".cuh", ".cpp", ".h")):
                with open(os.path.join(d, f), "rb") as fh:
                    h.update(f.encode() + b"\0" + fh.read())
    return h.hexdigest()


def is_current() -> bool:
    try:
        with open(STAMP_PATH) as f:
            return os.path.exists(LIB_PATH) and f.read().strip() == _fingerprint()
    except OSError:
        return False


def build_library(force: bool = False, verbose: bool = False, extra_flags: Optional[List[str]] = None) -> str:
    """Compile every source under ``csrc/`` into ``ops/libdlb200.so``; returns its path."""
    if not force and is_current():
        return LIB_PATH
    nvcc = nvcc_path()
    os.makedirs(os.path.join(ROOT, "build"), exist_ok=True)
    # several ranks may find the library stale at the same moment: one builds, the others wait on
    # the lock and then see a current library
    import fcntl

    with open(os.path.join(ROOT, "build", ".lock"), "w") as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)
        try:
            if not force and is_current():
                return LIB_PATH
            return _build_locked(nvcc, verbose, extra_flags)
        finally:
            fcntl.flock(lock, fcntl.LOCK_UN)


def _build_locked(nvcc: str, verbose: bool, extra_flags: Optional[List[str]]) -> str:
    objdir = os.path.join(ROOT, "build", f"obj.{os.getpid()}")
    os.makedirs(objdir, exist_ok=True)
    common = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "-I", INCLUDE] + ARCH_FLAGS
    if verbose:
        common += ["-Xptxas", "-v"]
    common += extra_flags or []
    objs, procs = [], []
    for src in _sources():
        obj = os.path.join(objdir, os.path.basename(src) + ".o")
        objs.append(obj)
        procs.append((src, subprocess.Popen([nvcc, "-c", src, "-o", obj] + common,
                                            stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    failed = False
    for src, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0 or verbose:
            sys.stderr.write(f"--- nvcc {os.path.basename(src)} (exit {p.returncode})\n{out}\n")
        failed |= p.returncode != 0
    if failed:
        raise RuntimeError("nvcc failed; see output above")
    tmp = LIB_PATH + f".tmp{os.getpid()}"
    link = [nvcc, "-shared", "-o", tmp] + objs + ARCH_FLAGS + ["-cudart", "static", "-Xlinker", "--no-undefined", "-ldl", "-lpthread", "-lrt"]
    r = subprocess.run(link, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout)
        raise RuntimeError("link failed")
    os.replace(tmp, LIB_PATH)
    with open(STAMP_PATH, "w") as f:
        f.write(_fingerprint())
    shutil.rmtree(objdir, ignore_errors=True)
    return LIB_PATH


if __name__ == "__main__":
    print(build_library(force="--force" in sys.argv, verbose="-v" in sys.argv))
