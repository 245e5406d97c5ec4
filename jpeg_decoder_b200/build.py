"""Build libb200jpeg.so (CUDA, sm_100a only) and the harness encoder, in-tree.

`python -m jpeg_decoder_b200.build` or `build_all()`.  nvcc cross-compiles without a GPU.
-fmad=false: the pixel stage must not contract a*b+c (exactness recipe, kernels_pixel.cuh); the
entropy stage is integer work and does not care.
"""
import os
import subprocess
import sys

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
LIB_SO = os.path.join(_HERE, "libb200jpeg.so")
SOURCES = [os.path.join(CSRC, "b200jpeg.cu"), os.path.join(CSRC, "host_parse.cpp")]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo", "-fmad=false",
              "-Xcompiler", "-fPIC", "-shared"]


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build_lib(force=False, verbose=False):
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(_HERE, "..", "include", "b200jpeg.h")]
    if force or _stale(LIB_SO, deps):
        nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
        cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-o", LIB_SO] + SOURCES
        subprocess.check_call(cmd)
    return LIB_SO


CLI_SRC = os.path.join(_HERE, "..", "tools", "b200jpeg_cli.cpp")
CLI_BIN = os.path.join(_HERE, "b200jpeg")


def build_cli(force=False):
    """tools/b200jpeg_cli.cpp: the reference's main() (jpeg.cpp:916-929) over include/b200jpeg.hpp."""
    build_lib()
    inc = os.path.join(_HERE, "..", "include")
    deps = [CLI_SRC, os.path.join(inc, "b200jpeg.hpp"), os.path.join(inc, "b200jpeg.h"), LIB_SO]
    if force or _stale(CLI_BIN, deps):
        subprocess.check_call(["g++", "-std=c++17", "-O2", "-Wall", "-o", CLI_BIN, CLI_SRC, "-L" + _HERE, "-lb200jpeg",
                               "-Wl,-rpath,$ORIGIN"])
    return CLI_BIN


def build_all(force=False, verbose=False):
    from .harness.synth import build_harness
    build_lib(force, verbose)
    build_harness(force)
    build_cli(force)


if __name__ == "__main__":
    build_all(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(LIB_SO)
