"""Builds the sm_100a C-ABI library ``_lib/liblifter_b200.so`` from ``csrc/*.cu`` with nvcc (in-tree)."""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIBDIR = os.path.join(HERE, "_lib")
LIBNAME = "liblifter_b200.so"
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "-shared"]


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


def is_stale():
    stamp = lib_path() + ".sha256"
    if not (os.path.exists(lib_path()) and os.path.exists(stamp)):
        return True
    return open(stamp).read().strip() != source_digest()


def build(force=False, verbose=False):
    """Compile every CUDA source for sm_100a into one shared library.  Returns its path."""
    if not force and not is_stale():
        return lib_path()
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        raise RuntimeError("nvcc not found: cannot build the sm_100a lifter library")
    os.makedirs(LIBDIR, exist_ok=True)
    srcs, _ = _sources()
    tmp = lib_path() + ".tmp"
    cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-o", tmp] + srcs
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + " ".join(cmd) + "\n" + r.stdout + r.stderr)
    if verbose:
        print(r.stderr)
    os.replace(tmp, lib_path())
    with open(lib_path() + ".sha256", "w") as f:
        f.write(source_digest())
    return lib_path()


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
