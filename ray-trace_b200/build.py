"""In-tree build of the CUDA shared libraries (sm_100a only).  `python ray-trace_b200/build.py [--force] [-v]`.

  libraytrace_b200.so      the product: the reference-facing C ABI of include/raytrace_b200.h, nothing else
  libraytrace_b200_dev.so  the same sources with -DRT_DEV: + the diagnostics of include/raytrace_b200_dev.h (bit-equality
                           self-test of the branch-free primitives, exact-path census, latency probes) and the
                           environment switches for scheduling experiments; used by the diagnostics tests and tools/
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.environ.get("RT_LIB") or os.path.join(HERE, "libraytrace_b200.so")  # RT_LIB: developer override (variant builds)
LIB_DEV = os.path.join(HERE, "libraytrace_b200_dev.so")
SOURCES = ["rt_kernels.cu"]
SOURCES_DEV = ["rt_kernels.cu", "rt_selftest.cu"]
HEADERS = ["rt_math.h", "rt_fast.cuh", "rt_physics.cuh", "rt_physics_fast.cuh", "rt_integrate.cuh",
           os.path.join("..", "..", "include", "raytrace_b200.h"), os.path.join("..", "..", "include", "raytrace_b200_dev.h")]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-t", "0", "-lineinfo", "-std=c++17",
    "-fmad=false",  # contraction off: fused multiply-adds exist only where rt_math.h writes them
    "-shared", "-Xcompiler", "-fPIC,-ffp-contract=off,-mfma,-fvisibility=hidden",
]


def _stale(lib, sources):
    if not os.path.exists(lib):
        return True
    t = os.path.getmtime(lib)
    deps = [os.path.join(CSRC, f) for f in sources + HEADERS] + [os.path.abspath(__file__)]
    return any(os.path.getmtime(d) > t for d in deps)


def needs_build():
    return _stale(LIB, SOURCES) or _stale(LIB_DEV, SOURCES_DEV)


def _compile(lib, sources, extra, verbose):
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    cmd = ([nvcc] + NVCC_FLAGS + extra + os.environ.get("RT_EXTRA_FLAGS", "").split() + (["-Xptxas", "-v"] if verbose else [])
           + ["-o", lib] + [os.path.join(CSRC, s) for s in sources])
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + r.stdout + r.stderr)
    return r.stderr


def build(force=False, verbose=False):
    jobs = []
    if force or _stale(LIB, SOURCES):
        jobs.append((LIB, SOURCES, []))
    if force or _stale(LIB_DEV, SOURCES_DEV):
        jobs.append((LIB_DEV, SOURCES_DEV, ["-DRT_DEV"]))
    with ThreadPoolExecutor(max_workers=2) as ex:
        for out in ex.map(lambda j: _compile(j[0], j[1], j[2], verbose), jobs):
            if verbose:
                print(out)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
