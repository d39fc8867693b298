"""Builds the sm_100a C-ABI library ``impala_b200/_C/libb200rl.so`` in-tree with nvcc.

The .so is git-ignored but travels to the GPU box with the repo snapshot.
"""
import glob
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
OUT_DIR = os.path.join(_HERE, "_C")
LIB_PATH = os.path.join(OUT_DIR, "libb200rl.so")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
    "--expt-extended-lambda", "-Xcompiler", "-fPIC", "-shared",
    "--threads", "0",                    # the .cu files compile in parallel (conv3x3.cu alone is ~70 s of templates)
]


def sources():
    return sorted(glob.glob(os.path.join(CSRC, "*.cu")))


def needs_build():
    if not os.path.exists(LIB_PATH):
        return True
    deps = sources() + glob.glob(os.path.join(CSRC, "*.cuh")) + \
        [os.path.join(_HERE, "..", "include", "b200rl.h")]
    t = os.path.getmtime(LIB_PATH)
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    """Compile every ``csrc/*.cu`` for sm_100a into one shared library."""
    if not force and not needs_build():
        return LIB_PATH
    os.makedirs(OUT_DIR, exist_ok=True)
    nvcc = os.environ.get("NVCC", "nvcc")
    cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-o", LIB_PATH] + sources()
    subprocess.check_call(cmd)
    return LIB_PATH


if __name__ == "__main__":
    print(build(force=True, verbose=True))
