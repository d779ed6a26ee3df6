"""Builds the sm_100a C-ABI library ``_lib/liblifter_b200.so`` from ``csrc/*.cu`` with nvcc (in-tree).

Staleness is decided from the BINARY: the build passes the sha256 of its sources to the compiler, which stores it in the
library as ``LIFTER_SRC_DIGEST=<hex>`` (``lifter_source_digest()``); ``is_stale()`` scans the file for that marker, so a
library left over from another checkout is never loaded, whatever side files say.  Concurrent builders (N torchrun ranks
on a fresh checkout) serialise on a file lock and compile into per-process temporaries.
"""
from __future__ import annotations

import fcntl
import hashlib
import os
import re
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIBDIR = os.path.join(HERE, "_lib")
LIBNAME = "liblifter_b200.so"
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "-shared"]
_MARKER = re.compile(rb"LIFTER_SRC_DIGEST=([0-9a-f]{64})")


def lib_path():
    return os.path.join(LIBDIR, LIBNAME)


def _sources():
    srcs = sorted(f for f in os.listdir(CSRC) if f.endswith(".cu"))
    deps = sorted(f for f in os.listdir(CSRC) if f.endswith((".cu", ".cuh")))
    deps_paths = [os.path.join(CSRC, f) for f in deps] + [os.path.join(HERE, "..", "include", "lifter_b200.h")]
    return [os.path.join(CSRC, f) for f in srcs], deps_paths


def source_digest():
    h = hashlib.sha256()
    for p in _sources()[1]:
        h.update(os.path.basename(p).encode())
        with open(p, "rb") as f:
            h.update(f.read())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def built_digest(path=None):
    """The source digest recorded INSIDE the library file (None: no library, or one that predates the marker)."""
    path = path or lib_path()
    if not os.path.exists(path):
        return None
    with open(path, "rb") as f:
        m = _MARKER.search(f.read())
    return m.group(1).decode() if m else None


def is_stale():
    return built_digest() != source_digest()


def build(force=False, verbose=False):
    """Compile every CUDA source for sm_100a into one shared library.  Returns its path."""
    if not force and not is_stale():
        return lib_path()
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        raise RuntimeError("nvcc not found: cannot build the sm_100a lifter library")
    os.makedirs(LIBDIR, exist_ok=True)
    with open(os.path.join(LIBDIR, ".build.lock"), "w") as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)                     # one builder at a time; the others find a fresh library
        try:
            if not force and not is_stale():
                return lib_path()
            digest = source_digest()
            srcs, _ = _sources()
            tmp = f"{lib_path()}.{os.getpid()}.tmp"
            jobs = str(min(len(srcs), os.cpu_count() or 1))
            cmd = [nvcc] + NVCC_FLAGS + ["--threads", jobs, f"-DLIFTER_SOURCE_DIGEST=\"{digest}\""] + \
                  (["-Xptxas", "-v"] if verbose else []) + ["-o", tmp] + srcs
            r = subprocess.run(cmd, capture_output=True, text=True)
            if r.returncode != 0:
                if os.path.exists(tmp):
                    os.remove(tmp)
                raise RuntimeError("nvcc failed:\n" + " ".join(cmd) + "\n" + r.stdout + r.stderr)
            if verbose:
                print(r.stderr)
            os.replace(tmp, lib_path())
        finally:
            fcntl.flock(lock, fcntl.LOCK_UN)
    return lib_path()


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
