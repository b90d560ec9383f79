"""Build recipe for libmpcgpu_b200.so (sm_100a only, in-tree so it travels with gpurun)."""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libmpcgpu_b200.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
CFLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
          "-Xcompiler", "-fPIC"]
LFLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-cudart", "shared"]
SOURCES = ["capi.cu", "capi_lat.cu", "capi_f64.cu", "solver.cu", "noalloc.cu"]
OBJDIR = os.path.join(HERE, "_obj")


def _deps():
    out = [os.path.join(HERE, "..", "include", "mpcgpu_b200.h")]
    for f in os.listdir(CSRC):
        if f.endswith((".cu", ".cuh", ".h")):
            out.append(os.path.join(CSRC, f))
    return out


def stale():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(p) > t for p in _deps())


def build(force=False, verbose=False):
    if not force and not stale():
        return LIB
    if not os.path.exists(NVCC):
        raise RuntimeError("nvcc not found at %s; cannot build %s" % (NVCC, LIB))
    # one nvcc per translation unit, in parallel (capi.cu alone is minutes of ptxas), then one link
    os.makedirs(OBJDIR, exist_ok=True)
    extra = os.environ.get("MPCGPU_B200_NVCC_EXTRA", "").split()
    procs = []
    for src in SOURCES:
        obj = os.path.join(OBJDIR, src.replace(".cu", ".o"))
        cmd = [NVCC, *CFLAGS, *extra, "-c", os.path.join(CSRC, src), "-o", obj]
        if verbose:
            print(" ".join(cmd))
        procs.append((obj, cmd, subprocess.Popen(cmd)))
    for obj, cmd, pr in procs:
        if pr.wait() != 0:
            raise subprocess.CalledProcessError(pr.returncode, cmd)
    cmd = [NVCC, *LFLAGS, "-o", LIB] + [obj for obj, _, _ in procs]
    if verbose:
        print(" ".join(cmd))
    subprocess.run(cmd, check=True)
    return LIB


if __name__ == "__main__":
    print(build(force=True, verbose=True))
