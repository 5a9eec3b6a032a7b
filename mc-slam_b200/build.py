"""Builds mc-slam_b200/libmcslam_b200.so (CUDA kernels + C ABI) for sm_100a with nvcc.  In-tree, so the .so travels with
the snapshot to the GPU box."""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "csrc", "mcslam_b200.cu")
OUT = os.path.join(HERE, "libmcslam_b200.so")
DEPS = [os.path.join(HERE, "csrc", f) for f in ("mcslam_b200.cu", "extract_kernels.cuh", "match_kernels.cuh", "octree_core.h",
                                                 "introsort_emul.h", "plan.h")] + [
    os.path.join(HERE, "..", "include", "mcslam_b200.h"), os.path.join(HERE, "..", "include", "mcs_orb_pattern.inc")]

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "--fmad=false",  # the reference's f32/f64 expressions are evaluated without contraction
              "-Xcompiler", "-fPIC,-fvisibility=hidden", "-shared", "-cudart", "static"]


def needs_build():
    if not os.path.exists(OUT):
        return True
    t = os.path.getmtime(OUT)
    return any(os.path.getmtime(d) > t for d in DEPS)


def build(force=False, verbose=False):
    if not force and not needs_build():
        return OUT
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-o", OUT, SRC]
    subprocess.check_call(cmd)
    return OUT


if __name__ == "__main__":
    print(build(force=True, verbose=True))
