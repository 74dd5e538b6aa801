"""Build the CUDA library (sm_100a) in-tree: unige_tasi_path_planners_b200/libtasi_gpu.so.

nvcc cross-compiles without a GPU, so this runs on the CPU build box; the .so travels to the GPU box with the
repo snapshot. -fmad=false keeps the float32 local updates free of FMA contraction (bit-parity with the
reference built with -ffp-contract=off).
"""
import os
import shutil
import subprocess

PKG = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG, "csrc")
LIB = os.path.join(PKG, "libtasi_gpu.so")
HOSTLIB = os.path.join(PKG, "libtasi_hostops.so")
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-fmad=false", "-std=c++17",
              "-Xcompiler", "-fPIC", "-shared"]


def _nvcc():
    for c in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if c and os.path.exists(c):
            return c
    raise RuntimeError("nvcc not found: the CUDA library cannot be built")


def _stale(out, srcs):
    if not os.path.exists(out):
        return True
    t = os.path.getmtime(out)
    return any(os.path.getmtime(s) > t for s in srcs)


def sources():
    return [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC))] + [
        os.path.join(PKG, "..", "include", "tasi_gpu.h")]


def build(force=False, verbose=False):
    """Compile libtasi_gpu.so (CUDA, C ABI) and libtasi_hostops.so (the local updates compiled for the host)."""
    srcs = sources()
    if force or _stale(LIB, srcs):
        cmd = [_nvcc()] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + [
            "-o", LIB, os.path.join(CSRC, "tasi_capi.cu")]
        subprocess.check_call(cmd)
    build_hostops(force)
    return LIB


def build_hostops(force=False):
    """The local updates of csrc/tasi_update.cuh compiled for the host (operator-level tests only)."""
    srcs = [os.path.join(CSRC, "tasi_hostops.cpp"), os.path.join(CSRC, "tasi_update.cuh"),
            os.path.join(CSRC, "tasi_extract.cuh"), os.path.join(CSRC, "tasi_parents.cuh")]
    if force or _stale(HOSTLIB, srcs):
        subprocess.check_call(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-o", HOSTLIB,
                               os.path.join(CSRC, "tasi_hostops.cpp")])
    return HOSTLIB


HOST_DIR = os.path.join(PKG, "host")
HOST_DEMO = os.path.join(HOST_DIR, "plan_demo")


def build_host_demo(force=False):
    """The C++ host classes (host/GpuPlanners.h) exercised by host/plan_demo.cpp, linked against libtasi_gpu.so."""
    srcs = [os.path.join(HOST_DIR, "plan_demo.cpp"), os.path.join(HOST_DIR, "GpuPlanners.h"),
            os.path.join(PKG, "..", "include", "tasi_gpu.h"), LIB]
    if force or _stale(HOST_DEMO, srcs):
        subprocess.check_call(["g++", "-std=c++17", "-O2", "-Wall", "-I", os.path.join(PKG, "..", "include"), "-I", HOST_DIR,
                               os.path.join(HOST_DIR, "plan_demo.cpp"), "-L", PKG, "-ltasi_gpu",
                               "-Wl,-rpath,$ORIGIN/..", "-o", HOST_DEMO])
    return HOST_DEMO


FIFO_DEMO = os.path.join(HOST_DIR, "fifo_demo")


def build_fifo_demo(force=False):
    """host/fifo_demo.cpp: the demos' planner process (FIFO wire protocol) over the GPU planners."""
    srcs = [os.path.join(HOST_DIR, "fifo_demo.cpp"), os.path.join(HOST_DIR, "GpuPlanners.h"),
            os.path.join(PKG, "..", "include", "tasi_gpu.h"), LIB]
    if force or _stale(FIFO_DEMO, srcs):
        subprocess.check_call(["g++", "-std=c++17", "-O2", "-Wall", "-I", os.path.join(PKG, "..", "include"), "-I", HOST_DIR,
                               os.path.join(HOST_DIR, "fifo_demo.cpp"), "-L", PKG, "-ltasi_gpu",
                               "-Wl,-rpath,$ORIGIN/..", "-o", FIFO_DEMO])
    return FIFO_DEMO


if __name__ == "__main__":
    import sys
    build(force="--force" in sys.argv, verbose="-v" in sys.argv)
    build_host_demo(force="--force" in sys.argv)
    build_fifo_demo(force="--force" in sys.argv)
    print(LIB)
