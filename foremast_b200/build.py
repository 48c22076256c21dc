"""Builds libforemast_gpu.so in-tree with nvcc for sm_100a (cross-compiles without a GPU)."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libforemast_gpu.so")
SOURCES = ["fm_api.cu", "fm_ingest.cu", "fm_synth.cu", "fm_band.cu", "fm_rank.cu", "fm_hw.cu", "fm_hw_tm.cu", "fm_scan.cu", "fm_verdict.cu", "fm_comm.cu"]
# --fmad=false: fused multiply-adds only where the source says fma() (oracle/DECISIONS.md)
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "--fmad=false",
              "-std=c++17", "-diag-suppress=186", "-Xcompiler", "-fPIC,-O2", "-shared", "--threads", "0", "-ldl"]


def _stale():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)]
    deps.append(os.path.join(HERE, "..", "include", "foremast_gpu.h"))
    return any(os.path.getmtime(d) > t for d in deps)


def build_library(force=False, verbose=False):
    if not force and not _stale():
        return LIB
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    tmp = LIB + ".building"  # renamed into place at the end: a snapshot of the tree never sees half a library
    cmd = [nvcc] + NVCC_FLAGS + ["-o", tmp] + [os.path.join(CSRC, s) for s in SOURCES]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
        print(" ".join(cmd))
    subprocess.run(cmd, check=True)
    os.replace(tmp, LIB)
    return LIB


if __name__ == "__main__":
    build_library(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(LIB)
