"""In-tree build of the CUDA shared library (sm_100a only).  `python ray-trace_b200/build.py`."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.environ.get("RT_LIB") or os.path.join(HERE, "libraytrace_b200.so")  # RT_LIB: developer override (variant builds)
SOURCES = ["rt_kernels.cu", "rt_selftest.cu"]
HEADERS = ["rt_math.h", "rt_fast.cuh", "rt_physics.cuh", "rt_physics_fast.cuh", "rt_integrate.cuh", os.path.join("..", "..", "include", "raytrace_b200.h")]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-t", "0", "-lineinfo", "-std=c++17",
    "-fmad=false",  # contraction off: fused multiply-adds exist only where rt_math.h writes them
    "-shared", "-Xcompiler", "-fPIC,-ffp-contract=off,-mfma,-fvisibility=hidden",
]


def needs_build():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, f) for f in SOURCES + HEADERS] + [os.path.abspath(__file__)]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    if not force and not needs_build():
        return LIB
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    cmd = [nvcc] + NVCC_FLAGS + os.environ.get("RT_EXTRA_FLAGS", "").split() + (["-Xptxas", "-v"] if verbose else []) + ["-o", LIB] + [os.path.join(CSRC, s) for s in SOURCES]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + r.stdout + r.stderr)
    if verbose:
        print(r.stderr)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
