"""Build the sm_100a shared library (psgradientsdf_b200/libpsg_b200.so) in-tree with nvcc.

nvcc cross-compiles without a GPU.  -fmad=false is part of the numerics contract (csrc/psg_device.cuh): index- and
visibility-deciding expressions must round exactly like the reference's x86-64 build, which has no FMA contraction.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
LIB = os.path.join(HERE, "libpsg_b200.so")
SOURCES = ["volume.cu", "surface.cu", "sweep.cu", "pcg.cu", "capi.cu", "host_loop.cu", "comm.cu", "normals.cu", "tracker.cu", "mesh.cu", "small_solve.cu", "host_solve.cpp"]
HEADERS = ["psg_device.cuh", "psg_host.h", "mc_table.h", "host_solve.h", os.path.join("..", "..", "include", "psg.h")]


def _nvcc() -> str:
    for c in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", shutil.which("nvcc")):
        if c and os.path.exists(c):
            return c
    raise RuntimeError("nvcc not found")


def _ccbin() -> list[str]:
    # the image exports CXX=/opt/gcc/bin/g++ whose wrapper links libstdc++ statically; use the system compiler
    return ["-ccbin", "/usr/bin/g++"] if os.path.exists("/usr/bin/g++") else []


FLAGS = ["-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-fmad=false",
         "-Xcompiler", "-fPIC", "-Xptxas", "-v"]


def needs_build() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, s) for s in SOURCES + HEADERS] + [os.path.abspath(__file__), os.path.join(HERE, "..", "examples", "voxelps_host.cpp")]
    if not os.path.exists(os.path.join(HERE, "voxelps_host")):
        return True
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not needs_build():
        return LIB
    os.makedirs(OBJ, exist_ok=True)
    nvcc = _nvcc()
    hdr_t = max(os.path.getmtime(os.path.join(CSRC, h)) for h in HEADERS)

    def compile_one(src: str) -> str:
        obj = os.path.join(OBJ, os.path.splitext(src)[0] + ".o")
        spath = os.path.join(CSRC, src)
        if not force and os.path.exists(obj) and os.path.getmtime(obj) > max(os.path.getmtime(spath), hdr_t):
            return obj
        cmd = [nvcc, *_ccbin(), *FLAGS, "-c", spath, "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        log = r.stdout + r.stderr
        with open(obj + ".log", "w") as f:
            f.write(" ".join(cmd) + "\n" + log)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src}:\n{log[-4000:]}")
        if verbose:
            print(log)
        return obj

    with ThreadPoolExecutor(max_workers=len(SOURCES)) as ex:
        objs = list(ex.map(compile_one, SOURCES))
    cmd = [nvcc, *_ccbin(), "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", LIB, *objs, "-ldl"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("link failed:\n" + r.stdout + r.stderr)
    build_host_example()
    return LIB


HOST_EXAMPLE = os.path.join(HERE, "voxelps_host")


def build_host_example() -> str:
    """examples/voxelps_host.cpp: a plain C++ host over the C ABI (no Python, no torch), built next to the library."""
    src = os.path.join(HERE, "..", "examples", "voxelps_host.cpp")
    cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    cmd = [cxx, "-std=c++17", "-O2", "-I" + os.path.join(HERE, "..", "include"), src, "-L" + HERE, "-lpsg_b200",
           "-Wl,-rpath,$ORIGIN", "-o", HOST_EXAMPLE]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("host example failed to build:\n" + r.stdout + r.stderr)
    return HOST_EXAMPLE


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
