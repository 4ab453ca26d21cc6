"""Build the CUDA library in-tree (nvcc, sm_100a only). Used by __graft_entry__.build()."""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
SRC_DIR = os.path.join(HERE, "csrc")
LIB_PATH = os.path.join(HERE, "libchronos_b200.so")
SOURCES = ["chronos.cu", "tracegen_tu.cu"]
HEADERS = ["common.cuh", "filter.cuh", "sort.cuh", "window.cuh", "chain.cuh", "fnscan.cuh", "tracegen.cuh", "exchange.cuh"]


def _stale():
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    deps = [os.path.join(SRC_DIR, f) for f in SOURCES + HEADERS] + [os.path.join(ROOT, "include", "chronos.h")]
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def build(force=False, verbose=False):
    """nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo ... -> libchronos_b200.so"""
    if not force and not _stale():
        return LIB_PATH
    cmd = [
        "nvcc", "-O3", "-std=c++17",
        "-gencode", "arch=compute_100a,code=sm_100a",
        "-lineinfo", "-Xcompiler", "-fPIC", "-shared",
        "-I", os.path.join(ROOT, "include"),
        "-o", LIB_PATH,
    ] + [os.path.join(SRC_DIR, s) for s in SOURCES]
    if verbose:
        cmd.insert(1, "-Xptxas")
        cmd.insert(2, "-v")
    subprocess.check_call(cmd)
    return LIB_PATH


HOSTGEN_PATH = os.path.join(ROOT, "tools", "libchronos_tracegen_host.so")


def build_hostgen(force=False):
    """g++ build of the trace generator alone (tools/tracegen_host.cpp): no CUDA inside."""
    src = os.path.join(ROOT, "tools", "tracegen_host.cpp")
    deps = [src, os.path.join(SRC_DIR, "tracegen.cuh"), os.path.join(ROOT, "include", "chronos.h")]
    if not force and os.path.exists(HOSTGEN_PATH) and all(os.path.getmtime(d) <= os.path.getmtime(HOSTGEN_PATH) for d in deps):
        return HOSTGEN_PATH
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-x", "c++", "-fPIC", "-shared", "-I", SRC_DIR,
                           "-I", os.path.join(ROOT, "include"), "-o", HOSTGEN_PATH, src])
    return HOSTGEN_PATH


if __name__ == "__main__":
    print(build(force=True, verbose=True))
