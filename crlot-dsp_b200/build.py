"""Build crlot-dsp_b200/libcrlot_cuda.so (the C-ABI library) for sm_100a with nvcc, in-tree.

    python crlot-dsp_b200/build.py [--force] [--verbose]

nvcc cross-compiles without a GPU.  The .so is git-ignored but travels to the GPU box with the
gpurun snapshot.  Also builds the C++ facade test driver (host/) when g++ is available.
"""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
LIB = os.path.join(HERE, "libcrlot_cuda.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
FLAGS = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr"]
SOURCES = ["k1_stft.cu", "k1_4096v2.cu", "k1_2048tm.cu", "k2_istft.cu", "k2_4096v2.cu", "k3_ola.cu", "k23_istft_ola.cu", "k23_4096.cu", "k23_2048tm.cu", "k_rt_2048tm.cu", "k_cfft.cu", "k_generic.cu", "aux_kernels.cu", "stream_kernels.cu", "framer.cu", "ring.cu", "shard.cu", "capi.cu"]


def _newer(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    os.makedirs(OBJ, exist_ok=True)
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".h", ".cuh"))]
    headers.append(os.path.join(HERE, "..", "include", "crlot_cuda.h"))
    jobs = []
    for s in SOURCES:
        src = os.path.join(CSRC, s)
        obj = os.path.join(OBJ, s.replace(".cu", ".o"))
        if force or _newer(obj, [src] + headers):
            cmd = [NVCC] + ARCH + FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-c", src, "-o", obj]
            jobs.append(cmd)

    def run(cmd):
        r = subprocess.run(cmd, capture_output=True, text=True)
        if verbose or r.returncode:
            sys.stderr.write(" ".join(cmd) + "\n" + r.stdout + r.stderr)
        if r.returncode:
            raise RuntimeError("nvcc failed: " + " ".join(cmd))

    with ThreadPoolExecutor(max_workers=4) as ex:
        list(ex.map(run, jobs))
    objs = [os.path.join(OBJ, s.replace(".cu", ".o")) for s in SOURCES]
    if force or jobs or _newer(LIB, objs):
        run([NVCC] + ARCH + ["-shared", "-Xcompiler", "-fPIC", "-o", LIB] + objs + ["-ldl"])
    return LIB


def build_host():
    """C++ facades (host/crlot_dsp.h) + their test driver, with g++ against the in-tree .so."""
    subprocess.check_call(["make", "-C", os.path.join(HERE, "host"), "facade_test"], stdout=subprocess.DEVNULL)
    return os.path.join(HERE, "host", "facade_test")


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv))
    print(build_host())
