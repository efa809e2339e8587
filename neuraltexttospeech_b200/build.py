"""Builds libhifigan_b200.so (hand-written sm_100a CUDA + the C ABI) in-tree with nvcc.

The shared library lives next to this file so that it travels with the repo snapshot to
the GPU box; it is git-ignored (built artefact).  nvcc cross-compiles without a GPU.
"""
from __future__ import annotations

import concurrent.futures
import hashlib
import os
import shutil
import subprocess

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG_DIR, "csrc")
ROOT = os.path.dirname(PKG_DIR)
BUILD_DIR = os.path.join(ROOT, "build")
LIB_PATH = os.path.join(PKG_DIR, "libhifigan_b200.so")
SOURCES = ("conv_ffma.cu", "conv_tc.cu", "resblock_tc.cu", "conv_tf32.cu", "engine.cu", "mel_frontend.cu", "noalloc.cu", "disc.cu")
HEADERS = (os.path.join(CSRC, "kernels.cuh"), os.path.join(CSRC, "tc_ptx.cuh"), os.path.join(ROOT, "include", "hifigan_b200.h"))
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC"]
if os.environ.get("HFG_BUILD_EPI_GROUPS"):       # experiment: number of 4-warp epilogue groups in conv_tc_kernel
    NVCC_FLAGS.append("-DHFG_TC_EPI_GROUPS=" + os.environ["HFG_BUILD_EPI_GROUPS"])
if os.environ.get("HFG_BUILD_TRACE") == "1":     # debug build: per-role clock stamps in conv_tc_kernel (tools/trace_layer.py)
    NVCC_FLAGS.append("-DHFG_TC_TRACE")


def _nvcc() -> str:
    exe = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(exe):
        raise RuntimeError("nvcc not found; cannot build libhifigan_b200.so")
    return exe


def _digest() -> str:
    h = hashlib.sha256()
    for p in [os.path.join(CSRC, s) for s in SOURCES] + list(HEADERS):
        with open(p, "rb") as f:
            h.update(f.read())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def have_nvcc() -> bool:
    return bool(shutil.which("nvcc")) or os.path.exists("/usr/local/cuda/bin/nvcc")


def is_current() -> bool:
    """True when the in-tree library was built from the sources as they are now."""
    stamp = os.path.join(BUILD_DIR, "libhifigan_b200.sha256")
    try:
        return os.path.exists(LIB_PATH) and open(stamp).read().strip() == _digest()
    except OSError:
        return False


def build_locked(verbose: bool = False) -> str:
    """build() under an exclusive file lock: concurrent callers (the ranks of a torchrun launch) compile once."""
    import fcntl
    os.makedirs(BUILD_DIR, exist_ok=True)
    with open(os.path.join(BUILD_DIR, ".build.lock"), "w") as lk:
        fcntl.flock(lk, fcntl.LOCK_EX)
        try:
            return build(verbose=verbose)      # a waiter finds the stamp current and returns at once
        finally:
            fcntl.flock(lk, fcntl.LOCK_UN)


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile (if stale) and return the path of the shared library."""
    os.makedirs(BUILD_DIR, exist_ok=True)
    stamp = os.path.join(BUILD_DIR, "libhifigan_b200.sha256")
    digest = _digest()
    if not force and os.path.exists(LIB_PATH) and os.path.exists(stamp) and open(stamp).read().strip() == digest:
        return LIB_PATH
    nvcc = _nvcc()

    def compile_one(src):
        obj = os.path.join(BUILD_DIR, src.replace(".cu", ".o"))
        cmd = [nvcc, *NVCC_FLAGS, "-Xptxas", "-v", "-c", os.path.join(CSRC, src), "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        with open(obj.replace(".o", ".log"), "w") as f:
            f.write(r.stdout + r.stderr)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
        return obj

    with concurrent.futures.ThreadPoolExecutor(max_workers=len(SOURCES)) as ex:
        objs = list(ex.map(compile_one, SOURCES))
    cmd = [nvcc, "-shared", "-o", LIB_PATH, *objs, "-cudart", "shared"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    with open(stamp, "w") as f:
        f.write(digest)
    if verbose:
        print(f"built {LIB_PATH}")
    return LIB_PATH


if __name__ == "__main__":
    build(force=True, verbose=True)
