"""Build libgxa_b200.so in-tree with nvcc for sm_100a (cross-compiles without a GPU).

    python -m mocov2_whisper_flamingo_b200.build [--force]

The shared object is a plain C-ABI library (include/gxa.h): no torch, no pybind; the Python
shell binds it with ctypes.  It is git-ignored but travels to the GPU box with the snapshot.
"""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
# GXA_BUILD_VARIANT=name (with GXA_NVCC_EXTRA=-D...) builds libgxa_b200.<name>.so beside the product library, for A/B runs
# selected with GXA_LIB=<path>; the product library and its object directory are untouched.
_VARIANT = os.environ.get("GXA_BUILD_VARIANT", "")
OBJ = os.path.join(HERE, "build" + ("_" + _VARIANT if _VARIANT else ""))
LIB = os.path.join(HERE, "libgxa_b200." + _VARIANT + ".so" if _VARIANT else "libgxa_b200.so")
SOURCES = ["gxa_gemm.cu", "gxa_gemm2.cu", "gxa_rowwise.cu", "gxa_attn.cu", "gxa_attn_tc.cu", "gxa_attn_bwd_tc.cu", "gxa_attn_f32.cu", "gxa_api.cu", "gxa_noalloc.cu", "gxa_skinny.cu"]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
         "-Xcompiler", "-fPIC"] + os.environ.get("GXA_NVCC_EXTRA", "").split()     # e.g. -DGXA_BWD_STAMPS (tools/)


def _deps_mtime() -> float:
    files = [os.path.join(CSRC, f) for f in os.listdir(CSRC)]
    files.append(os.path.join(os.path.dirname(HERE), "include", "gxa.h"))
    return max(os.path.getmtime(f) for f in files)


def up_to_date() -> bool:
    return os.path.exists(LIB) and os.path.getmtime(LIB) >= _deps_mtime()


def _compile(src: str) -> str:
    obj = os.path.join(OBJ, src.replace(".cu", ".o"))
    cmd = [NVCC, *FLAGS, "-c", os.path.join(CSRC, src), "-o", obj]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
    return obj


def build(force: bool = False, verbose: bool = True) -> str:
    if not force and up_to_date():
        return LIB
    if not os.path.exists(NVCC):
        raise RuntimeError(f"nvcc not found at {NVCC}; libgxa_b200.so cannot be built here")
    os.makedirs(OBJ, exist_ok=True)
    with ThreadPoolExecutor(max_workers=len(SOURCES)) as ex:
        objs = list(ex.map(_compile, SOURCES))
    cmd = [NVCC, "-shared", "-o", LIB, *objs, "-gencode", "arch=compute_100a,code=sm_100a"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    if verbose:
        print(f"built {LIB}")
    return LIB


if __name__ == "__main__":
    build(force="--force" in sys.argv)
