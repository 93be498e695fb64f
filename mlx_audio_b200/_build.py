"""Builds libkokoro_b200.so in-tree with nvcc for sm_100a (no torch headers, pure C ABI)."""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG_DIR, "csrc")
LIB_NAME = "libkokoro_b200.so"
LIB_PATH = os.path.join(PKG_DIR, LIB_NAME)
SOURCES = ["common.cu", "weights.cu", "elementwise.cu", "conv_f32.cu", "conv_tc.cu", "conv_fused.cu", "source.cu", "istft_head.cu",
           "fft.cu", "codec.cu", "lstm.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=default"]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found; cannot build libkokoro_b200.so")


def _stale() -> bool:
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [
        os.path.join(os.path.dirname(PKG_DIR), "include", "kokoro_b200.h"), os.path.abspath(__file__)]
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile every CUDA source for sm_100a and link the shared library.  Returns its path.
    KK_TUNE=1 / KK_FUSED_DBG=1 in the environment add -DKK_TUNE (experiment knobs read from the environment) /
    -DKK_FUSED_DBG (per-barrier wait counters in the fused kernel) - lab builds, never the default."""
    extra = [f"-D{k}" for k in ("KK_TUNE", "KK_FUSED_DBG") if os.environ.get(k) == "1"]
    extra += os.environ.get("KK_DEFS", "").split()          # further -D flags of a lab build (A/B experiments)
    if not force and not extra and not _stale():
        return LIB_PATH
    nvcc = _nvcc()
    lib_path = LIB_PATH if not extra else os.path.join(PKG_DIR, f"libkokoro_b200_{os.environ.get('KK_LAB_NAME', 'lab')}.so")   # lab builds never shadow the product
    objdir = os.path.join(PKG_DIR, "build" if not extra else f"build_{os.environ.get('KK_LAB_NAME', 'lab')}")
    os.makedirs(objdir, exist_ok=True)
    procs = []
    for src in SOURCES:
        obj = os.path.join(objdir, src.replace(".cu", ".o"))
        cmd = [nvcc, *NVCC_FLAGS, *extra, "-c", os.path.join(CSRC, src), "-o", obj]
        if verbose:
            print(" ".join(cmd), file=sys.stderr)
        procs.append((src, obj, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    objs = []
    for src, obj, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0:
            raise RuntimeError(f"nvcc failed on {src}:\n{out}")
        if verbose and out.strip():
            print(out, file=sys.stderr)
        objs.append(obj)
    tmp = lib_path + ".tmp"
    cmd = [nvcc, "-shared", "-o", tmp, *objs, "-gencode", "arch=compute_100a,code=sm_100a", "-lcudart_static",
           "-Xcompiler", "-fPIC"]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}")
    os.replace(tmp, lib_path)
    return lib_path


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))
