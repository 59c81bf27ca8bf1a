"""Builds libmuscle_b200.so (sm_100a only) in-tree with nvcc.

    python -m muscle_b200.build [--force] [--verbose]

The library exposes the C ABI of include/muscle_b200.h; nothing in it depends on torch.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIBDIR = os.path.join(HERE, "lib")
OBJDIR = os.path.join(LIBDIR, "obj")
LIB = os.path.join(LIBDIR, "libmuscle_b200.so")

DIAG_LIB = os.path.join(LIBDIR, "libmuscle_b200_diag.so")    # test fixture (include/muscle_b200_diag.h), not the product

SOURCES = ["infonce.cu", "small_losses.cu", "pcm.cu", "fieldloss.cu", "emd.cu", "infonce_qk.cu", "pool.cu", "cls_losses.cu", "er.cu", "irn.cu"]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-std=c++17", "-O3", "-lineinfo",
    "-Xcompiler", "-fPIC",
    "-Xptxas", "-v",
]


def _nvcc():
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        raise RuntimeError("nvcc not found; muscle_b200 has no CPU fallback and cannot be built without CUDA")
    return nvcc


def _newest_dep():
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)]
    deps.append(os.path.join(os.path.dirname(HERE), "include", "muscle_b200.h"))
    deps.append(os.path.abspath(__file__))
    return max(os.path.getmtime(d) for d in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    os.makedirs(OBJDIR, exist_ok=True)
    if not force and os.path.exists(LIB) and os.path.exists(DIAG_LIB) and min(os.path.getmtime(LIB), os.path.getmtime(DIAG_LIB)) >= _newest_dep():
        return LIB
    nvcc = _nvcc()

    def compile_one(src):
        obj = os.path.join(OBJDIR, src.replace(".cu", ".o"))
        cmd = [nvcc, *NVCC_FLAGS, "-c", os.path.join(CSRC, src), "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        log = r.stdout + r.stderr
        with open(obj + ".log", "w") as f:
            f.write(log)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed for %s:\n%s" % (src, log))
        if verbose:
            print(log)
        return obj

    with ThreadPoolExecutor(max_workers=4) as ex:
        objs = list(ex.map(compile_one, SOURCES + ["debug_tile.cu"]))
    for lib, members in ((LIB, objs[:-1]), (DIAG_LIB, objs[-1:])):
        cmd = [nvcc, "-shared", "-o", lib, *members, "-gencode", "arch=compute_100a,code=sm_100a"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("link failed:\n" + r.stdout + r.stderr)
    return LIB


if __name__ == "__main__":
    path = build(force="--force" in sys.argv, verbose="--verbose" in sys.argv)
    print(path)
