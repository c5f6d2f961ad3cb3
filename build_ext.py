"""Builds libdcgan_b200.so (all hand-written sm_100a kernels) in-tree with nvcc.

``python build_ext.py`` cross-compiles without a GPU.  The .so is git-ignored but travels to the GPU
box with the snapshot.  Objects are rebuilt only when their source (or a header) is newer.
"""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

ROOT = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(ROOT, "csrc")
PKG = os.path.join(ROOT, "distributed_tensorflow_for_dcgan_b200")
LIB = os.path.join(PKG, "lib", "libdcgan_b200.so")
SOURCES = ["igemm_sm100.cu", "igemm2_sm100.cu", "runtime.cu", "bn_act.cu", "heads_optim.cu", "conv_c3.cu", "conv_c3_tc.cu", "comm_sm100.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC"]
HEADERS = [os.path.join(CSRC, "common.cuh"), os.path.join(CSRC, "igemm2.h"),
           os.path.join(ROOT, "include", "dcgan_b200.h")]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.isabs(cand) and os.path.exists(cand) or not os.path.isabs(cand)):
            return cand
    return "nvcc"


def _stale(target: str, deps) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def build(verbose: bool = False, force: bool = False) -> str:
    objdir = os.path.join(ROOT, "build", "obj")
    os.makedirs(objdir, exist_ok=True)
    os.makedirs(os.path.dirname(LIB), exist_ok=True)
    srcs = [s for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]
    jobs = []
    for s in srcs:
        src = os.path.join(CSRC, s)
        obj = os.path.join(objdir, s.replace(".cu", ".o"))
        if force or _stale(obj, [src] + HEADERS):
            jobs.append((src, obj))

    def compile_one(job):
        src, obj = job
        cmd = [_nvcc()] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-c", "-o", obj, src]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
        if verbose:
            print(r.stderr)
        return obj

    with ThreadPoolExecutor(max_workers=min(8, max(1, len(jobs)))) as ex:
        list(ex.map(compile_one, jobs))
    objs = [os.path.join(objdir, s.replace(".cu", ".o")) for s in srcs]
    if force or jobs or _stale(LIB, objs):
        cmd = [_nvcc(), "-shared", "-o", LIB] + objs + ["-lcudart"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    return LIB


def build_harness(name: str = "test_igemm2") -> str:
    """Stand-alone kernel harness (csrc/<name>.cu) linked against the library objects; runs without torch."""
    build()
    objdir = os.path.join(ROOT, "build", "obj")
    exe = os.path.join(ROOT, "build", name)
    objs = [os.path.join(objdir, s.replace(".cu", ".o")) for s in SOURCES]
    src = os.path.join(CSRC, name + ".cu")
    if _stale(exe, [src] + objs):
        cmd = [_nvcc()] + [f for f in NVCC_FLAGS if f not in ("-Xcompiler", "-fPIC")] + ["-o", exe, src] + objs + ["-lcuda"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"harness build failed:\n{r.stdout}\n{r.stderr}")
    return exe


if __name__ == "__main__":
    print(build(verbose="-v" in sys.argv, force="-f" in sys.argv))
    for arg in sys.argv[1:]:
        if arg.startswith("harness="):
            print(build_harness(arg.split("=", 1)[1]))
