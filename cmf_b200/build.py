"""Build ``cmf_b200/lib/libcmf_b200.so`` (the C-ABI CUDA library, sm_100a only) with nvcc.

    python -m cmf_b200.build [--force] [--verbose]

nvcc cross-compiles without a GPU; the .so is git-ignored but travels to the GPU box with the
repo snapshot.  ``-lineinfo`` keeps the ncu source page mapped to the .cu files.
"""
from __future__ import annotations

import glob
import os
import shutil
import subprocess
import sys

PKG = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG, "csrc")
LIB = os.path.join(PKG, "lib", "libcmf_b200.so")
HEADER = os.path.join(os.path.dirname(PKG), "include", "cmf_b200.h")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "-Xcompiler", "-fPIC",
]
LINK_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-cudart", "static", "-Xcompiler", "-fPIC"]
OBJ_DIR = os.path.join(PKG, "lib", "obj")


def sources():
    return sorted(glob.glob(os.path.join(CSRC, "*.cu")))


def _headers():
    return glob.glob(os.path.join(CSRC, "*.cuh")) + [HEADER, os.path.abspath(__file__)]


def _flags_now() -> str:
    return " ".join(NVCC_FLAGS + os.environ.get("CMF_NVCC_EXTRA", "").split())


def needs_build() -> bool:
    if not os.path.exists(LIB):
        return True
    tag = os.path.join(OBJ_DIR, ".flags")
    if os.path.exists(tag) and open(tag).read() != _flags_now():      # built with other -D switches
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(p) > t for p in sources() + _headers())


def build(force: bool = False, verbose: bool = False) -> str:
    """One object per .cu (compiled in parallel, rebuilt only when the source or a header is newer),
    then one link.  Objects live in cmf_b200/lib/obj (git-ignored, gpurun-ignored)."""
    if not force and not needs_build():
        return LIB
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        raise RuntimeError("nvcc not found: cannot build libcmf_b200.so")
    os.makedirs(OBJ_DIR, exist_ok=True)
    extra = os.environ.get("CMF_NVCC_EXTRA", "").split()      # e.g. -DCMF_CONV_PROF for the clock64 trace
    tag = os.path.join(OBJ_DIR, ".flags")
    flags_now = " ".join(NVCC_FLAGS + extra)
    if not os.path.exists(tag) or open(tag).read() != flags_now:
        force = True
    hdr_t = max(os.path.getmtime(p) for p in _headers())
    jobs = []
    for src in sources():
        obj = os.path.join(OBJ_DIR, os.path.basename(src)[:-3] + ".o")
        if force or not os.path.exists(obj) or os.path.getmtime(obj) < max(os.path.getmtime(src), hdr_t):
            cmd = [nvcc, *NVCC_FLAGS, *extra, *(["-Xptxas", "-v"] if verbose else []), "-c", "-o", obj, src]
            jobs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    failed = []
    for src, proc in jobs:
        out, _ = proc.communicate()
        if verbose or proc.returncode != 0:
            sys.stderr.write(out)
        if proc.returncode != 0:
            failed.append(os.path.basename(src))
    if failed:
        raise RuntimeError("nvcc failed building libcmf_b200.so: " + ", ".join(failed))
    with open(tag, "w") as f:
        f.write(flags_now)
    objs = [os.path.join(OBJ_DIR, os.path.basename(s)[:-3] + ".o") for s in sources()]
    res = subprocess.run([nvcc, *LINK_FLAGS, "-o", LIB + ".tmp", *objs], capture_output=True, text=True)
    if res.returncode != 0:
        sys.stderr.write(res.stdout + res.stderr)
        raise RuntimeError("nvcc failed linking libcmf_b200.so")
    os.replace(LIB + ".tmp", LIB)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv))
