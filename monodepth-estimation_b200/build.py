"""Builds ``libmde.so`` (the sm_100a kernels + C-ABI) in-tree with nvcc.

    python monodepth-estimation_b200/build.py [--force]

The shared object lands next to this file so it travels to the GPU box with the repo
snapshot; it is git-ignored.  nvcc cross-compiles without a GPU.
"""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "libmde.so")
OBJ_DIR = os.path.join(HERE, "build")
SOURCES = ["api.cu", "tc_gemm.cu", "tc_gemm2.cu", "tc_mlp.cu", "conv_halo.cu", "conv_small.cu", "decoder_tail.cu", "attention.cu", "pointwise.cu", "refine.cu", "terrain.cu", "depth_filters.cu", "backward.cu", "train_ops.cu"]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr",
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.isabs(cand) and os.path.exists(cand) or not os.path.isabs(cand)):
            return cand
    raise RuntimeError("nvcc not found")


def _stale(target: str, deps: list[str]) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def _inputs():
    srcs = [os.path.join(CSRC, s) for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".h", ".cuh"))]
    headers.append(os.path.join(os.path.dirname(HERE), "include", "mde.h"))
    return srcs, headers


def is_stale() -> bool:
    """True when libmde.so is missing or older than any source / header it is built from."""
    srcs, headers = _inputs()
    return _stale(OUT, srcs + headers)


def build(force: bool = False, verbose: bool = False) -> str:
    srcs, headers = _inputs()
    if not force and not _stale(OUT, srcs + headers):
        return OUT
    os.makedirs(OBJ_DIR, exist_ok=True)
    nvcc = _nvcc()

    def compile_one(src: str) -> str:
        obj = os.path.join(OBJ_DIR, os.path.basename(src)[:-3] + ".o")
        if force or _stale(obj, [src] + headers):
            cmd = [nvcc, *NVCC_FLAGS, "-c", src, "-o", obj]
            if os.environ.get("MDE_BUILD_PROF") == "1":  # in-kernel GEMM counters (tools/gemm_bench.py, MDE_GEMM_PROF=1)
                cmd.insert(1, "-DMDE_GEMM_PROF")
                cmd.insert(1, "-DMDE_ATTN_PROF")
            if verbose:
                cmd.insert(1, "-Xptxas=-v")
            r = subprocess.run(cmd, capture_output=True, text=True)
            if r.returncode != 0:
                raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
            if verbose:
                print(r.stderr)
        return obj

    with ThreadPoolExecutor(max_workers=len(srcs)) as ex:
        objs = list(ex.map(compile_one, srcs))
    r = subprocess.run([nvcc, "-shared", "-o", OUT, *objs, "-gencode", "arch=compute_100a,code=sm_100a"],
                       capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
