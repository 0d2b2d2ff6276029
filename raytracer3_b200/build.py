"""Builds librt3.so (CUDA, sm_100a) in-tree with nvcc.  No JIT cache: the .so travels with the repo."""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
INCLUDE = os.path.join(HERE, "..", "include")
SO = os.path.join(HERE, "librt3.so")
SOURCES = ["api.cu", "build.cu", "trace.cu", "passes.cu"]
HEADERS = [os.path.join(CSRC, "common.cuh"), os.path.join(CSRC, "bvh_logic.h"), os.path.join(CSRC, "trace_kernel.cuh"), os.path.join(CSRC, "pass_kernels.cuh"), os.path.join(CSRC, "pass_types.h"), os.path.join(INCLUDE, "rt3.h"),
           os.path.join(INCLUDE, "rt3_shared_math.h")]

# -fmad=false: the determinism contract of include/rt3_shared_math.h (explicit fmaf still compiles to FFMA).
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-fmad=false",
              "-Xcompiler", "-fPIC,-fvisibility=hidden,-ffp-contract=off", "-shared"]


def stale():
    if not os.path.exists(SO):
        return True
    t = os.path.getmtime(SO)
    deps = [os.path.join(CSRC, s) for s in SOURCES] + HEADERS
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    if not force and not stale():
        return SO
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    extra = os.environ.get("RT3_NVCC_EXTRA", "").split()      # experiment switches (-DRT3_...), never set by the product build
    cmd = [nvcc] + NVCC_FLAGS + extra + (["-Xptxas", "-v"] if verbose else []) + ["-I", INCLUDE, "-I", CSRC, "-o", SO] + [os.path.join(CSRC, s) for s in SOURCES]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + r.stdout + r.stderr)
    if verbose:
        print(r.stderr)
    return SO
