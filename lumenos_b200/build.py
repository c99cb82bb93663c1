"""Builds liblumenos_b200.so (sm_100a) in-tree with nvcc.  `python -m lumenos_b200.build`"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
SO = os.environ.get("LUMENOS_B200_SO") or os.path.join(HERE, "liblumenos_b200.so")   # override: A/B builds of tuning variants
SOURCES = ["api.cu", "api_client.cu", "api_comm.cu", "api_static.cu", "api_limbs.cu", "kernels_comm.cu", "kernels.cu", "kernels_blk.cu", "kernels_mac.cu", "kernels_commit.cu", "kernels_client.cu", "plan.cpp"]
HEADERS = ["device.cuh", "ntt_blk.cuh", "tma.cuh", "kernels.cuh", "commit.cuh", "client.cuh", "comm.cuh", "api_internal.hpp", "plan.hpp", "hostmath.hpp", os.path.join("..", "..", "include", "lumenos_b200.h")]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "--fmad=false",
         "-Xcompiler", "-fPIC,-O3,-Wall", "-shared", "-cudart", "static", "-ldl"]


def needs_build():
    if not os.path.exists(SO):
        return True
    t = os.path.getmtime(SO)
    return any(os.path.getmtime(os.path.join(CSRC, f)) > t for f in SOURCES + HEADERS)


def build(force=False, verbose=False):
    if not force and not needs_build():
        return SO
    cmd = [NVCC] + FLAGS + os.environ.get("LUMENOS_B200_NVCC_FLAGS", "").split() + (["-Xptxas", "-v"] if verbose else []) + ["-o", SO] + [os.path.join(CSRC, f) for f in SOURCES]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if verbose or r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed building liblumenos_b200.so")
    return SO


if __name__ == "__main__":
    build(force=True, verbose="-v" in sys.argv)
    print(SO)
