"""In-tree build of the device libraries (nvcc cross-compiles sm_100a without a GPU).

  libb200sv.so ........ kernels.cu + tile3.cu + shard.cu + simulator.cpp + comm.cpp: CUDA kernels, host
                        layer, rank-to-rank plumbing (NCCL via dlopen), C ABI (include/b200sv.h)
  librtd_b200sv.so .... plugin.cpp: QuantumDevice vtable adapter for Catalyst's runtime.  Needs the
                        reference's interface headers (runtime/include, runtime/lib/backend/common),
                        so it is only (re)built where /root/reference or $CATALYST_RUNTIME_SRC is
                        present; the built .so travels with the tree.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
INCLUDE = os.path.join(ROOT, "include")
LIB_DIR = os.path.join(HERE, "lib")
LIB_SV = os.path.join(LIB_DIR, "libb200sv.so")
LIB_PLUGIN = os.path.join(LIB_DIR, "librtd_b200sv.so")

NVCC = os.environ.get("NVCC") or shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]


def _stale(target: str, sources: list[str]) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources if os.path.exists(s))


def reference_runtime_dir() -> str | None:
    for cand in (os.environ.get("CATALYST_RUNTIME_SRC"), "/root/reference/runtime"):
        if cand and os.path.exists(os.path.join(cand, "include", "QuantumDevice.hpp")):
            return cand
    return None


def build_sv(force: bool = False, verbose: bool = False) -> str:
    srcs = [os.path.join(CSRC, "kernels.cu"), os.path.join(CSRC, "tile3.cu"), os.path.join(CSRC, "shard.cu"), os.path.join(CSRC, "simulator.cpp"),
            os.path.join(CSRC, "simulator_jit.cpp"), os.path.join(CSRC, "jit.cpp"), os.path.join(CSRC, "comm.cpp")]
    deps = srcs + [os.path.join(CSRC, "simulator.hpp"), os.path.join(CSRC, "comm.hpp"), os.path.join(CSRC, "jit.hpp"),
                   os.path.join(INCLUDE, "b200sv.h"),
                   os.path.join(INCLUDE, "b200sv_kernels.h")]
    if not force and not _stale(LIB_SV, deps):
        return LIB_SV
    os.makedirs(LIB_DIR, exist_ok=True)
    cmd = [NVCC, *ARCH, "-lineinfo", "-O3", "-std=c++17", "-shared", "-Xcompiler", "-fPIC,-Wall",
           "-I", INCLUDE, "-I", CSRC, "-x", "cu", *srcs, "-o", LIB_SV, "-lcudart_static", "-ldl",
           "-lpthread", "-lrt"]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    subprocess.check_call(cmd)
    return LIB_SV


def build_plugin(force: bool = False) -> str | None:
    rt = reference_runtime_dir()
    src = os.path.join(CSRC, "plugin.cpp")
    if rt is None or not os.path.exists(src):
        return LIB_PLUGIN if os.path.exists(LIB_PLUGIN) else None
    deps = [src, os.path.join(INCLUDE, "b200sv.h"), LIB_SV]
    if not force and not _stale(LIB_PLUGIN, deps):
        return LIB_PLUGIN
    cmd = ["g++", "-std=c++20", "-O2", "-shared", "-fPIC", "-Wall", "-I", INCLUDE,
           "-I", os.path.join(rt, "include"), "-I", os.path.join(rt, "lib", "backend", "common"),
           src, "-o", LIB_PLUGIN, "-L", LIB_DIR, "-lb200sv", "-Wl,-rpath,$ORIGIN"]
    subprocess.check_call(cmd)
    return LIB_PLUGIN


def build_all(force: bool = False, verbose: bool = False):
    sv = build_sv(force, verbose)
    pl = build_plugin(force)
    return sv, pl


if __name__ == "__main__":
    print(build_all(force="--force" in sys.argv, verbose="-v" in sys.argv))
