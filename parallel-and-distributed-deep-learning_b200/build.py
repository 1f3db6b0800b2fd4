"""Builds libb200dp.so in-tree with nvcc for sm_100a (no other target, no fallback)."""
import os
import subprocess
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "libb200dp.so")
SOURCES = ["api.cu", "noalloc.cu", "vmm.cpp", "layout.cpp", "fdxchg.cpp"]
HEADERS = ["kernels.cuh", "internal.h", os.path.join("..", "..", "include", "b200dp.h")]
NVCC_FLAGS = [
    "-std=c++17", "-O3", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
    "-Xcompiler", "-fPIC", "-shared", "-cudart", "static",
]


def _nvcc():
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.isabs(cand) and os.path.exists(cand) or not os.path.isabs(cand)):
            return cand
    return "nvcc"


def stale():
    if not os.path.exists(OUT):
        return True
    t = os.path.getmtime(OUT)
    return any(os.path.getmtime(os.path.join(CSRC, f)) > t for f in SOURCES + HEADERS)


def build(force=False, verbose=False):
    if not force and not stale():
        return OUT
    cmd = [_nvcc()] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-o", OUT] + SOURCES
    t0 = time.time()
    subprocess.check_call(cmd, cwd=CSRC)
    _record(cmd, time.time() - t0)
    return OUT


def _record(cmd, seconds):
    """Leaves a note of the compile that produced the library (command, seconds, host) where a
    GPU-box run brings it back from: gpurun_out/build_record.json.  Best effort."""
    try:
        import json
        import platform
        root = os.path.dirname(HERE)
        d = os.path.join(root, "gpurun_out")
        os.makedirs(d, exist_ok=True)
        with open(os.path.join(d, "build_record.json"), "w") as f:
            json.dump({"cmd": cmd, "cwd": CSRC, "seconds": round(seconds, 2), "output": OUT,
                       "bytes": os.path.getsize(OUT), "host": platform.node(), "when": time.strftime("%Y-%m-%dT%H:%M:%SZ", time.gmtime())}, f)
    except Exception:
        pass


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
