"""In-tree build of libsailrecon_b200.so (hand-written sm_100a CUDA kernels + the C-ABI of include/sailrecon_b200.h).

nvcc cross-compiles for sm_100a without a GPU; the resulting .so lives in-tree (git-ignored) under
``self-supervise-sfm_b200/lib/`` so it travels with the source snapshot.  No JIT cache, no torch extension:
the library has a plain C ABI and is loaded with ctypes (see _lib.py).
"""
from __future__ import annotations

import concurrent.futures
import hashlib
import os
import shutil
import subprocess
import sys

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG_DIR, "csrc")
LIB_DIR = os.path.join(PKG_DIR, "lib")
OBJ_DIR = os.path.join(PKG_DIR, "build")
LIB_PATH = os.path.join(LIB_DIR, "libsailrecon_b200.so")
HEADER = os.path.join(os.path.dirname(PKG_DIR), "include", "sailrecon_b200.h")

SOURCES = [
    "sr_api.cu",
    "layernorm.cu",
    "relayout.cu",
    "misc.cu",
    "gemm_f32.cu",
    "attn_f32.cu",
    "gemm_tc.cu",
    "attn_tc.cu",
    "attn_tc4.cu",
    "attn_tc3.cu", "attn_bwd.cu", "train_rows.cu", "pose.cu", "dpt.cu", "loss.cu", "head_train.cu",
]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    "-Xcompiler", "-fPIC",
    "--expt-relaxed-constexpr",
    # mbarrier waits are bounded: a pipeline-protocol bug traps (CUDA error through the C-ABI) instead of hanging the GPU
    "-DSR_DEBUG_HANG_TRAP",
]


class NvccNotFound(RuntimeError):
    """No compiler on this machine (a GPU box running a prebuilt snapshot): the only build failure that may be tolerated."""


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise NvccNotFound("nvcc not found: libsailrecon_b200.so cannot be built (there is no CPU fallback)")


def _digest() -> str:
    h = hashlib.sha256()
    deps = [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC))] + [HEADER]
    for path in deps:
        with open(path, "rb") as f:
            h.update(os.path.basename(path).encode())      # not the absolute path: the snapshot lives elsewhere on the GPU box
            h.update(f.read())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def is_current() -> bool:
    stamp = LIB_PATH + ".sha256"
    if not (os.path.exists(LIB_PATH) and os.path.exists(stamp)):
        return False
    with open(stamp) as f:
        return f.read().strip() == _digest()


def _compile(src: str, verbose: bool) -> str:
    obj = os.path.join(OBJ_DIR, src.replace(".cu", ".o"))
    cmd = [_nvcc(), *NVCC_FLAGS, "-c", os.path.join(CSRC, src), "-o", obj]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed on {src}:\n{r.stdout}\n{r.stderr}")
    if verbose:
        with open(obj + ".ptxas.log", "w") as f:
            f.write(r.stderr)
    return obj


def build_library(force: bool = False, verbose: bool = False) -> str:
    """Compile every .cu for sm_100a and link the shared library.  Returns its path."""
    if not force and is_current():
        return LIB_PATH
    os.makedirs(LIB_DIR, exist_ok=True)
    os.makedirs(OBJ_DIR, exist_ok=True)
    import fcntl
    with open(os.path.join(LIB_DIR, ".build.lock"), "w") as lock:        # one builder at a time (torchrun: N ranks)
        fcntl.flock(lock, fcntl.LOCK_EX)
        if not force and is_current():                                    # another rank built it while we waited
            return LIB_PATH
        with concurrent.futures.ThreadPoolExecutor(max_workers=min(8, len(SOURCES))) as ex:
            objs = list(ex.map(lambda s: _compile(s, verbose), SOURCES))
        tmp = LIB_PATH + f".tmp{os.getpid()}"
        cmd = [_nvcc(), "-shared", "-o", tmp, *objs, "-gencode", "arch=compute_100a,code=sm_100a"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
        os.replace(tmp, LIB_PATH)                                         # atomic: a concurrent dlopen never sees a partial file
        with open(LIB_PATH + ".sha256", "w") as f:
            f.write(_digest())
    return LIB_PATH


if __name__ == "__main__":
    path = build_library(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(path)
