"""Builds the C-ABI shared library (adaface_b200/lib/libadaface_b200.so) for sm_100a with nvcc.

In-tree and explicit (no JIT cache): the built .so travels with the repo snapshot to the GPU box.
    python -m adaface_b200.build [--force]
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG_DIR, "csrc")
LIB_DIR = os.path.join(PKG_DIR, "lib")
LIB_PATH = os.path.join(LIB_DIR, "libadaface_b200.so")
SOURCES = ["common.cu", "elementwise.cu", "norms.cu", "gemm.cu", "attention.cu", "attn_explicit.cu", "ada.cu", "backward.cu", "conv_small.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr", "-diag-suppress", "550"]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found; cannot build adaface_b200 CUDA library")


def _digest() -> str:
    h = hashlib.sha256()
    files = [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC))]
    files.append(os.path.join(os.path.dirname(PKG_DIR), "include", "adaface_b200.h"))
    for f in files:
        with open(f, "rb") as fh:
            h.update(os.path.basename(f).encode())     # not the absolute path: the checkout moves (GPU box scratch dirs)
            h.update(fh.read())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def header_digest() -> str:
    """sha256 of include/adaface_b200.h: compiled into the library (afb_header_digest) so that a loader can tell a
    library built against another revision of the argument structs / signatures from a current one."""
    with open(os.path.join(os.path.dirname(PKG_DIR), "include", "adaface_b200.h"), "rb") as fh:
        return hashlib.sha256(fh.read()).hexdigest()


def is_current() -> bool:
    stamp = LIB_PATH + ".sha256"
    return os.path.exists(LIB_PATH) and os.path.exists(stamp) and open(stamp).read().strip() == _digest()


def build(force: bool = False, verbose: bool = False) -> str:
    """Compiles and links the library if the sources changed.  Safe under concurrent callers (one process per GPU
    under torchrun): an exclusive file lock serialises them, the late ones find the library current, and the .so is
    moved into place atomically so that no process can dlopen a half-written file."""
    if not force and is_current():
        return LIB_PATH
    os.makedirs(LIB_DIR, exist_ok=True)
    import fcntl
    with open(os.path.join(LIB_DIR, ".build.lock"), "w") as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)
        try:
            if not force and is_current():
                return LIB_PATH
            return _build_locked(verbose)
        finally:
            fcntl.flock(lock, fcntl.LOCK_UN)


def _build_locked(verbose: bool) -> str:
    nvcc = _nvcc()
    objs = []
    procs = []
    tag = f".{os.getpid()}"
    for src in SOURCES:
        obj = os.path.join(LIB_DIR, src.replace(".cu", tag + ".o"))
        cmd = [nvcc, *NVCC_FLAGS, "-c", os.path.join(CSRC, src), "-o", obj]
        if src == "common.cu":
            cmd.insert(1, f'-DAFB_HEADER_DIGEST="{header_digest()}"')
        if verbose:
            cmd[1:1] = ["-Xptxas", "-v"]
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
        objs.append(obj)
    failed = False
    for src, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0:
            failed = True
            sys.stderr.write(f"[adaface_b200.build] nvcc failed on {src}:\n{out}\n")
        elif verbose and out:
            sys.stderr.write(out)
    if failed:
        raise RuntimeError("adaface_b200 CUDA build failed")
    tmp_lib = LIB_PATH + tag + ".tmp"
    link = [nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", tmp_lib, *objs, "-lcudart_static",
            "-ldl", "-lrt", "-lpthread"]
    subprocess.run(link, check=True)
    for o in objs:
        os.remove(o)
    stamp = LIB_PATH + ".sha256"
    if os.path.exists(stamp):
        os.remove(stamp)                       # never a current-looking stamp next to a library being replaced
    os.replace(tmp_lib, LIB_PATH)
    with open(stamp + tag, "w") as fh:
        fh.write(_digest())
    os.replace(stamp + tag, stamp)
    return LIB_PATH


if __name__ == "__main__":
    path = build(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(path)
