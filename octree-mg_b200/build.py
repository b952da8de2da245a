"""Build the native libraries in-tree (they travel to the GPU box with the snapshot).

  libomg_b200.so  CUDA kernels + C ABI (include/omg_gpu.h), sm_100a only, -fmad=false so that
                  results are bitwise reproducible against the CPU oracle's -ffp-contract=off build
  libomg_host.so  C++ host mirror of the reference's Fortran API (host/m_octree_mg.cpp)
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
HOST = os.path.join(HERE, "host")
BUILD = os.path.join(HERE, "build")
INCLUDE = os.path.join(os.path.dirname(HERE), "include")

NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-fmad=false", "-std=c++17",
              "-Xcompiler", "-fPIC", "-ccbin", "/usr/bin/g++"]
CU_SOURCES = ["omg_core.cu", "omg_kernels.cu", "omg_fast.cu", "omg_comm.cu", "omg_fs.cu", "omg_fs_kernel.cpp", "omg_plan.cpp"]
LIB_CUDA = os.path.join(HERE, "libomg_b200.so")
LIB_HOST = os.path.join(HERE, "libomg_host.so")


def _newer(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def _headers():
    hs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".h", ".cuh"))]
    hs += [os.path.join(INCLUDE, f) for f in os.listdir(INCLUDE)]
    hs.append(os.path.abspath(__file__))
    return hs


def _run(cmd):
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        sys.stderr.write(" ".join(cmd) + "\n" + r.stdout)
        raise RuntimeError("build failed: " + " ".join(cmd[:3]))
    return r.stdout


def build(force=False, verbose=False):
    os.makedirs(BUILD, exist_ok=True)
    hdrs = _headers()
    jobs = []
    objs = []
    for src in CU_SOURCES:
        s = os.path.join(CSRC, src)
        o = os.path.join(BUILD, os.path.splitext(src)[0] + ".o")
        objs.append(o)
        if force or _newer(o, [s] + hdrs):
            jobs.append([NVCC] + NVCC_FLAGS + ["-c", s, "-o", o] + (["-Xptxas", "-v"] if verbose else []))
    with ThreadPoolExecutor(max_workers=4) as ex:
        outs = list(ex.map(_run, jobs))
    if verbose:
        for o in outs:
            print(o)
    if force or jobs or _newer(LIB_CUDA, objs):
        _run([NVCC, "-shared", "-o", LIB_CUDA] + objs + ["-lcudart", "-ldl"])
    host_src = os.path.join(HOST, "m_octree_mg.cpp")
    if force or _newer(LIB_HOST, [host_src, os.path.join(HOST, "m_octree_mg.hpp"), LIB_CUDA] + hdrs):
        _run(["/usr/bin/g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-o", LIB_HOST, host_src,
              "-L" + HERE, "-lomg_b200", "-Wl,-rpath,$ORIGIN"])
    return LIB_CUDA, LIB_HOST


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print("built", LIB_CUDA, LIB_HOST)
