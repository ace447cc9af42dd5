"""Builds libmgb200.so in-tree with nvcc for sm_100a (python -m minigrid_dynamicprogramming_b200.build)."""
from __future__ import annotations

import glob
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libmgb200.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
         "-Xcompiler", "-fPIC", "--fmad=true", "-shared",
         "-cudart", "shared"]


def sources():
    return sorted(glob.glob(os.path.join(CSRC, "*.cu")))


def needs_build():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = sources() + glob.glob(os.path.join(CSRC, "*.cuh")) + glob.glob(os.path.join(HERE, "..", "include", "*.h"))
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False, defines=(), out=None):
    """defines / out: build an experimental variant (-DNAME=VALUE ...) into another file; select it at run time with
    MGB200_LIB=<path> (tools/ab_bench.sh)."""
    out = out or LIB
    if out == LIB and not force and not needs_build():
        return LIB
    # One builder at a time (the ranks of a torchrun job import the package together), and the library appears atomically:
    # nvcc writes a private file that is renamed over the target.
    import fcntl
    with open(out + ".lock", "w") as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)
        try:
            if out == LIB and not force and not needs_build():      # another process built it while this one waited
                return LIB
            tmp = f"{out}.tmp.{os.getpid()}"
            cmd = [NVCC] + FLAGS + [f"-D{d}" for d in defines] + (["-Xptxas", "-v"] if verbose else []) + ["-o", tmp] + sources()
            try:
                subprocess.check_call(cmd)
                os.replace(tmp, out)
            finally:
                if os.path.exists(tmp):
                    os.remove(tmp)
        finally:
            fcntl.flock(lock, fcntl.LOCK_UN)
    return out


if __name__ == "__main__":
    defs = [a[2:] for a in sys.argv[1:] if a.startswith("-D")]
    outs = [a[3:] for a in sys.argv[1:] if a.startswith("-o=")]
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv, defines=defs, out=outs[0] if outs else None))
