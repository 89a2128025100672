"""Build libnerf_b200.so (the C-ABI of include/nerf_b200.h) in-tree with nvcc for sm_100a.

    python -m nerf_kinematics_b200._build [--force]

nvcc cross-compiles without a GPU.  The library has no torch / Python dependency: it is loaded with
ctypes (nerf_kinematics_b200/_lib.py) and could equally be linked by any C caller.
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
LIB = os.path.join(HERE, "libnerf_b200.so")

ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
# NB_EXTRA_NVCC_FLAGS=-DNB_TIMING_EXPERIMENTS builds the variant with the kernel-side timing hooks (tools/ only)
COMMON = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC"] + os.environ.get("NB_EXTRA_NVCC_FLAGS", "").split()
# (source, extra flags).  The stage kernels must round like the reference's separate fp32 mul/add.
SOURCES = [
    ("stages.cu", ["--fmad=false"]),
    ("volume.cu", ["--fmad=false"]),
    ("model.cu", []),
    ("mlp_fp32.cu", []),
    ("mlp_tc.cu", []),
    ("mlp_tc_bwd.cu", []),
    ("render.cu", ["--fmad=false"]),
    ("optim.cu", ["--fmad=false"]),
]


def _nvcc():
    for c in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if c and (os.path.isabs(c) and os.path.exists(c) or not os.path.isabs(c)):
            return c
    return "nvcc"


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    os.makedirs(OBJ, exist_ok=True)
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".h", ".cuh"))]
    headers.append(os.path.join(os.path.dirname(HERE), "include", "nerf_b200.h"))
    jobs = []
    objs = []
    for src, extra in SOURCES:
        s = os.path.join(CSRC, src)
        if not os.path.exists(s):
            continue
        o = os.path.join(OBJ, src.replace(".cu", ".o"))
        objs.append(o)
        if force or _stale(o, [s] + headers):
            jobs.append([_nvcc()] + ARCH + COMMON + extra + ["-c", s, "-o", o])

    def run(cmd):
        r = subprocess.run(cmd, capture_output=True, text=True)
        if verbose or r.returncode != 0:
            sys.stderr.write(" ".join(cmd) + "\n" + r.stdout + r.stderr)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed for %s" % cmd[-3])
        return r

    if jobs:
        with ThreadPoolExecutor(max_workers=min(len(jobs), os.cpu_count() or 4)) as ex:
            list(ex.map(run, jobs))
    if jobs or force or _stale(LIB, objs):
        run([_nvcc()] + ARCH + ["-shared", "-o", LIB] + objs)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
