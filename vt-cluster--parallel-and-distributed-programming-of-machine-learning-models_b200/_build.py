"""Builds liblora_b200.so in-tree with nvcc for sm_100a (no JIT cache, no torch headers)."""
from __future__ import annotations

import os
import shutil
import subprocess

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
REPO_ROOT = os.path.dirname(PKG_DIR)
CSRC = os.path.join(PKG_DIR, "csrc")
LIB_PATH = os.path.join(PKG_DIR, "liblora_b200.so")
SOURCES = ["lora_b200.cu"]
HEADERS = ["common.cuh", "colproj_sm100.cuh", "rowproj_small.cuh", "gemm_sm100.cuh", "gemm_masked_sm100.cuh", "skinny.cuh", "act.cuh", "ce.cuh", "p2p.cuh", "optim.cuh"]


def _stale() -> bool:
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    deps = [os.path.join(CSRC, f) for f in SOURCES + HEADERS] + [os.path.join(REPO_ROOT, "include", "lora_b200.h")]
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def build_variant(name: str, extra_flags) -> str:
    """Experiment builds (scripts/ only): the same sources with extra -D flags into liblora_b200_<name>.so next to
    the shipped library; a process selects one with LORA_B200_LIB=<path> before the first call (binding.lib)."""
    out = os.path.join(PKG_DIR, f"liblora_b200_{name}.so")
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    cmd = [nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-shared", "-Xcompiler",
           "-fPIC", "-I", os.path.join(REPO_ROOT, "include"), "-I", CSRC, "-o", out] + list(extra_flags) + \
          [os.path.join(CSRC, s) for s in SOURCES]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + res.stdout + res.stderr)
    return out


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile csrc/*.cu -> liblora_b200.so (sm_100a only).  Returns the library path.
    Safe when several ranks of one job call it at once: an exclusive file lock serialises the staleness check and
    the compile, and the library is written under a temporary name and renamed into place."""
    if not force and not _stale():
        return LIB_PATH
    import fcntl

    with open(LIB_PATH + ".lock", "w") as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)
        try:
            if not force and not _stale():  # another rank built it while this one waited
                return LIB_PATH
            nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
            tmp = f"{LIB_PATH}.tmp{os.getpid()}"
            cmd = [
                nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
                "-shared", "-Xcompiler", "-fPIC", "-I", os.path.join(REPO_ROOT, "include"), "-I", CSRC,
                "-o", tmp,
            ] + [os.path.join(CSRC, s) for s in SOURCES]
            if verbose:
                cmd.insert(1, "-Xptxas=-v")
            res = subprocess.run(cmd, capture_output=True, text=True)
            if res.returncode != 0:
                if os.path.exists(tmp):
                    os.remove(tmp)
                raise RuntimeError("nvcc failed:\n" + res.stdout + res.stderr)
            os.replace(tmp, LIB_PATH)
            if verbose:
                print(res.stderr)
        finally:
            fcntl.flock(lock, fcntl.LOCK_UN)
    return LIB_PATH
