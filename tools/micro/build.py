"""Builds tools/micro/libdfn_microbench.so: the tcgen05 issue-rate / store-path probes used while bringing the MLP kernels up
(umma_bench.cu, store_bench.cu).  They are NOT part of the product library or its header.
    python tools/micro/build.py"""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "..", "..", "deepfields-nerf-_b200", "csrc")
LIB = os.path.join(HERE, "libdfn_microbench.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")


def build():
    srcs = [os.path.join(HERE, "umma_bench.cu"), os.path.join(HERE, "store_bench.cu"), os.path.join(CSRC, "dfn_error.cu")]
    cmd = [NVCC, "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-Xcompiler", "-fPIC",
           "--expt-relaxed-constexpr", "-shared", "-o", LIB, *srcs]
    subprocess.run(cmd, check=True)
    return LIB


if __name__ == "__main__":
    print(build())
