"""Builds libmab200.so (hand-written sm_100a CUDA + C ABI) in-tree with nvcc."""
import glob
import os
import shutil
import subprocess

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG_DIR, "csrc")
LIB_PATH = os.path.join(PKG_DIR, "libmab200.so")
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo",
              "-shared", "-Xcompiler", "-fPIC", "-cudart", "static"]


def _sources():
    return sorted(glob.glob(os.path.join(CSRC, "*.cu")) + glob.glob(os.path.join(CSRC, "*.cuh")) +
                  glob.glob(os.path.join(PKG_DIR, "..", "include", "*.h")))


def is_stale():
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    return any(os.path.getmtime(s) > t for s in _sources())


def build(force=False, verbose=False, out=None, defines=()):
    """Compile csrc/mab_api.cu -> libmab200.so.  Returns the library path.  `out` / `defines` build a
    kernel-experiment variant next to it (loaded with MAB_LIB=<path>, see profiles/)."""
    if out is None and not force and not is_stale():
        return LIB_PATH
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        raise RuntimeError("nvcc not found: cannot build libmab200.so")
    extra = os.environ.get("MAB_NVCC_EXTRA", "").split()
    target = out or LIB_PATH
    cmd = [nvcc] + NVCC_FLAGS + extra + [f"-D{d}" for d in defines] + (["-Xptxas", "-v"] if verbose else []) + \
          ["-o", target, os.path.join(CSRC, "mab_api.cu")]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + res.stdout + res.stderr)
    if verbose:
        print(res.stderr)
    return target


if __name__ == "__main__":
    print(build(force=True, verbose=True))
