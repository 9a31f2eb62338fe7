"""Build recipe for the native pieces (explicit nvcc / g++; nothing is JIT-compiled at import).

  libpiccolo_b200.so       CUDA kernels + C ABI        (piccolo_b200/csrc, sm_100a only)
  libpiccolo_host.so       host C++ mirror of the reference classes over the C ABI (piccolo_b200/host)
"""
from __future__ import annotations

import os
import subprocess
import sys

PKG = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG)
LIB = os.path.join(PKG, "libpiccolo_b200.so")
HOST_LIB = os.path.join(PKG, "libpiccolo_host.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "--shared",
              "-Xcompiler", "-fPIC", "-Xcompiler", "-ffp-contract=off"]


def _newer(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources)


def kernel_sha16() -> str:
    """Hash of the CUDA sources the library is built from: ties a profile (profiles/lbs_traffic.json) to the kernels it measured."""
    import hashlib
    h = hashlib.sha256()
    csrc = os.path.join(PKG, "csrc")
    for f in sorted(os.listdir(csrc)):
        h.update(open(os.path.join(csrc, f), "rb").read())
    return h.hexdigest()[:16]


def build_cuda(force: bool = False, verbose: bool = False) -> str:
    csrc = os.path.join(PKG, "csrc")
    srcs = [os.path.join(csrc, f) for f in sorted(os.listdir(csrc))] + [os.path.join(ROOT, "include", "piccolo_b200.h")]
    if force or _newer(LIB, srcs):
        cmd = [NVCC] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-o", LIB] + [s for s in srcs if s.endswith(".cu")]
        subprocess.check_call(cmd)
    return LIB


def build_host(force: bool = False) -> str:
    hdir = os.path.join(PKG, "host")
    if not os.path.isdir(hdir):
        return ""
    srcs = [os.path.join(hdir, f) for f in sorted(os.listdir(hdir)) if f.endswith((".cpp", ".h"))]  # not the built test_host: it is newer than the library by construction
    cpps = [s for s in srcs if s.endswith(".cpp") and not os.path.basename(s).startswith("test_")]
    if not cpps:
        return ""
    if force or _newer(HOST_LIB, srcs + [LIB]):
        cmd = [os.environ.get("CXX", "g++"), "-std=c++17", "-O2", "-fPIC", "-shared", "-ffp-contract=off", "-I", os.path.join(ROOT, "include"),
               "-I", hdir, "-o", HOST_LIB] + cpps + ["-L", PKG, "-lpiccolo_b200", "-Wl,-rpath,$ORIGIN"]
        subprocess.check_call(cmd)
    return HOST_LIB


def build_host_test(force: bool = False) -> str:
    """piccolo_b200/host/test_host: the driver tests/test_host_*.py run against the host C++ mirror."""
    hdir = os.path.join(PKG, "host")
    exe = os.path.join(hdir, "test_host")
    src = os.path.join(hdir, "test_host.cpp")
    if force or _newer(exe, [src, HOST_LIB, os.path.join(hdir, "piccolo_anim.h")]):
        subprocess.check_call([os.environ.get("CXX", "g++"), "-std=c++17", "-O2", "-I", os.path.join(ROOT, "include"), "-I", hdir, "-o", exe, src,
                               "-L", PKG, "-lpiccolo_host", "-lpiccolo_b200", "-Wl,-rpath,$ORIGIN/.."])
    return exe


def build_all(force: bool = False, verbose: bool = False):
    build_cuda(force, verbose)
    build_host(force)
    build_host_test(force)


if __name__ == "__main__":
    build_all(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(LIB)
