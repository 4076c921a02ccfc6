"""Build the native kernel library in-tree:  ``python -m rt1b200.build``  →  ``rt1b200/lib/librt1b200.so``.

Every ``rt1b200/csrc/*.cu`` is compiled by ``nvcc`` for **sm_100a only**
(``-gencode arch=compute_100a,code=sm_100a -lineinfo``; nvcc cross-compiles, no GPU needed) and linked into one
C-ABI shared object (declarations: ``include/rt1b200.h``).  No PyTorch headers are involved — the Python side
binds with ``ctypes`` — so a full rebuild takes seconds, and the .so travels with a repository snapshot.
Objects are cached by source hash under ``rt1b200/lib/obj``.
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from typing import List

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIBDIR = os.path.join(HERE, "lib")
OBJDIR = os.path.join(LIBDIR, "obj")
LIB = os.path.join(LIBDIR, "librt1b200.so")

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC,-fvisibility=hidden", "--expt-relaxed-constexpr", "-Xptxas", "-v"]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def _hash(path: str, extra: bytes) -> str:
    h = hashlib.sha1(extra)
    with open(path, "rb") as f:
        h.update(f.read())
    return h.hexdigest()[:16]


def sources() -> List[str]:
    return sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".cu"))


def build(verbose: bool = False, force: bool = False) -> str:
    os.makedirs(OBJDIR, exist_ok=True)
    nvcc = _nvcc()
    hdr = b"".join(open(os.path.join(CSRC, f), "rb").read() for f in sorted(os.listdir(CSRC))
                   if f.endswith((".cuh", ".h")))
    flags = list(NVCC_FLAGS)
    if os.environ.get("RT1B200_PDL_MODE"):      # experiment switch: where kernels wait / trigger (see common.cuh)
        flags.append("-DRT1_PDL_MODE=" + str(int(os.environ["RT1B200_PDL_MODE"])))
    hdr += " ".join(flags).encode()
    jobs = []
    objs = []
    for src in sources():
        obj = os.path.join(OBJDIR, os.path.basename(src)[:-3] + "." + _hash(src, hdr) + ".o")
        objs.append(obj)
        if force or not os.path.exists(obj):
            jobs.append((src, obj))

    def compile_one(job):
        src, obj = job
        log = obj[:-2] + ".log"
        r = subprocess.run([nvcc, *flags, "-I", CSRC, "-c", src, "-o", obj], capture_output=True, text=True)
        with open(log, "w") as f:
            f.write(r.stdout + r.stderr)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
        if verbose:
            print(f"[rt1b200.build] compiled {os.path.basename(src)}", flush=True)
        return obj

    if jobs:
        with ThreadPoolExecutor(max_workers=min(8, len(jobs))) as ex:
            list(ex.map(compile_one, jobs))
    keep = set(objs)
    for f in os.listdir(OBJDIR):
        p = os.path.join(OBJDIR, f)
        if f.endswith(".o") and p not in keep:
            os.remove(p)
            if os.path.exists(p[:-2] + ".log"):
                os.remove(p[:-2] + ".log")
    if jobs or force or not os.path.exists(LIB):
        tmp = LIB + ".tmp"
        r = subprocess.run([nvcc, "-shared", "-o", tmp, *objs, "-gencode", "arch=compute_100a,code=sm_100a"], capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
        os.replace(tmp, LIB)
        if verbose:
            print(f"[rt1b200.build] linked {LIB}", flush=True)
    return LIB


if __name__ == "__main__":
    print(build(verbose=True, force="--force" in sys.argv))
