"""In-tree build of libb200ops.so (sm_100a only) with plain nvcc.

    python -m b200ops.build            # from micro-vllm_b200/
    python micro-vllm_b200/b200ops/build.py [--force]

The shared library is written next to this file so that it travels with the
source tree (it is git-ignored, not gpurun-ignored).  nvcc cross-compiles
sm_100a without a GPU, so this also runs in the CPU-only build container.
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)                       # micro-vllm_b200/
CSRC = os.path.join(ROOT, "csrc")
INCLUDE = os.path.join(os.path.dirname(ROOT), "include")
OBJ_DIR = os.path.join(HERE, "build")
LIB_PATH = os.path.join(HERE, "libb200ops.so")

SOURCES = ["elementwise.cu", "decode_attn.cu", "prefill_attn.cu", "allreduce_norm.cu", "sampling.cu", "skinny_gemm.cu", "stream_gemm.cu"]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    "--expt-relaxed-constexpr",
    "-Xcompiler", "-fPIC",
    "-Xptxas", "-v",
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: libb200ops cannot be built (there is no CPU fallback)")


def _digest(paths) -> str:
    h = hashlib.sha256()
    h.update(" ".join(NVCC_FLAGS).encode())
    for p in sorted(paths):
        with open(p, "rb") as f:
            h.update(p.encode())
            h.update(f.read())
    return h.hexdigest()


def _headers() -> list[str]:
    return [os.path.join(CSRC, h) for h in sorted(os.listdir(CSRC)) if h.endswith(".cuh")] + \
           [os.path.join(INCLUDE, "b200_abi.h")]


def sources() -> list[str]:
    return [os.path.join(CSRC, s) for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]


def is_fresh() -> bool:
    stamp = os.path.join(OBJ_DIR, "stamp")
    if not (os.path.exists(LIB_PATH) and os.path.exists(stamp)):
        return False
    deps = sources() + _headers()
    with open(stamp) as f:
        return f.read().strip() == _digest(deps)


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile every .cu for sm_100a and link libb200ops.so; returns its path."""
    if not force and is_fresh():
        return LIB_PATH
    os.makedirs(OBJ_DIR, exist_ok=True)
    # one builder at a time (torchrun starts every rank at once): the others wait, then find the fresh library
    import fcntl
    with open(os.path.join(OBJ_DIR, "lock"), "w") as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)
        try:
            if not force and is_fresh():
                return LIB_PATH
            return _build_locked(verbose)
        finally:
            fcntl.flock(lock, fcntl.LOCK_UN)


def _build_locked(verbose: bool) -> str:
    nvcc = _nvcc()
    srcs = sources()

    def compile_one(src: str) -> str:
        obj = os.path.join(OBJ_DIR, os.path.basename(src).replace(".cu", ".o"))
        cmd = [nvcc, *NVCC_FLAGS, "-I", INCLUDE, "-c", src, "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        log = os.path.join(OBJ_DIR, os.path.basename(src) + ".ptxas.log")
        with open(log, "w") as f:
            f.write(r.stdout + r.stderr)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
        if verbose:
            print(r.stderr)
        return obj

    with ThreadPoolExecutor(max_workers=min(len(srcs), os.cpu_count() or 1)) as ex:
        objs = list(ex.map(compile_one, srcs))
    tmp = LIB_PATH + f".{os.getpid()}.tmp"
    cmd = [nvcc, "-shared", "-o", tmp, *objs]           # static cudart: self-contained .so
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    os.replace(tmp, LIB_PATH)                           # a reader never sees a half-written library
    deps = srcs + _headers()
    with open(os.path.join(OBJ_DIR, "stamp"), "w") as f:
        f.write(_digest(deps))
    return LIB_PATH


if __name__ == "__main__":
    path = build(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(path)
