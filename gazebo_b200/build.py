"""Builds gazebo_b200/libgzcuda.so in-tree with nvcc for sm_100a (cross-compiles without a GPU)."""
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
OUT = os.path.join(_HERE, "libgzcuda.so")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
    # parity mode: no fused multiply-add, IEEE division and square root (the defaults for
    # double) so device arithmetic matches the reference built with -ffp-contract=off
    "-fmad=false", "-Xcompiler", "-fPIC", "-shared",
]


def _newest_source():
    t = 0.0
    for d in (CSRC, os.path.join(_HERE, "..", "include")):
        for f in os.listdir(d):
            t = max(t, os.path.getmtime(os.path.join(d, f)))
    return t


def build_lib(force=False, verbose=False):
    if not force and os.path.exists(OUT) and os.path.getmtime(OUT) >= _newest_source():
        return OUT
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    extra = os.environ.get("GZCUDA_NVCC_EXTRA", "").split()   # developer experiments, e.g. -DGZ_RING_G=2
    cmd = [nvcc] + NVCC_FLAGS + extra + (["-Xptxas", "-v"] if verbose else []) + ["-o", OUT, os.path.join(CSRC, "gzcuda.cu")]
    subprocess.check_call(cmd)
    return OUT


if __name__ == "__main__":
    print(build_lib(force=True, verbose=True))
