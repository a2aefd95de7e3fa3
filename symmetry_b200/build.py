"""Builds symmetry_b200/libsymrelay.so in-tree with nvcc for sm_100a.

The shared object is the product: hand-written CUDA kernels plus the C ABI of
include/symrelay.h.  It links the CUDA runtime statically so it loads (and its
symbols can be checked) on machines without a GPU; every computing entry point
still needs a B200 at run time."""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
SO = os.path.join(HERE, "libsymrelay.so")
SOURCES = ["symrelay.cu", "plan_scan_avx2.cpp", "udp_batch.cpp"]
# everything symrelay.cu includes: stale() rebuilds when any of them is newer than the .so
HEADERS = sorted(f for f in os.listdir(CSRC) if f.endswith((".cuh", ".h", ".inc"))) + [
    os.path.join("..", "..", "include", "symrelay.h")]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17", "-DSYMRELAY_HAVE_AVX2_SCAN",
    "-Xcompiler", "-fPIC,-O3,-fvisibility=hidden",
    "-shared", "-cudart", "static", "-Xlinker", "--exclude-libs,ALL",
    "-Xptxas", "-v",
]


def nvcc() -> str:
    p = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(p):
        raise RuntimeError("nvcc not found: the CUDA extension cannot be built")
    return p


def stale() -> bool:
    if not os.path.exists(SO):
        return True
    t = os.path.getmtime(SO)
    deps = [os.path.join(CSRC, s) for s in SOURCES + HEADERS] + [os.path.abspath(__file__)]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False, extra=(), out: str = SO) -> str:
    if not force and not stale() and out == SO:
        return SO
    cmd = [nvcc()] + NVCC_FLAGS + list(extra) + ["-o", out] + [os.path.join(CSRC, s) for s in SOURCES]
    res = subprocess.run(cmd, capture_output=True, text=True)
    log = res.stdout + res.stderr
    with open(os.path.join(HERE, "build.log"), "w") as f:
        f.write(" ".join(cmd) + "\n" + log)
    if res.returncode != 0:
        sys.stderr.write(log)
        raise RuntimeError("nvcc failed building libsymrelay.so")
    if verbose:
        print(log)
    return out


DRIVER_SO = os.path.join(os.path.dirname(HERE), "tools", "libe2e_driver.so")


def build_driver(force: bool = False) -> str:
    """tools/e2e_driver.c: a plain-C caller of the public ABI (no CUDA in it), linked against libsymrelay.so."""
    tools = os.path.join(os.path.dirname(HERE), "tools")
    srcs = [os.path.join(tools, "e2e_driver.c"), os.path.join(tools, "latency_driver.c")]
    if force or not os.path.exists(DRIVER_SO) or any(os.path.getmtime(s) > os.path.getmtime(DRIVER_SO) for s in srcs):
        subprocess.run(["gcc", "-O2", "-fPIC", "-shared", "-o", DRIVER_SO] + srcs + ["-L" + HERE, "-lsymrelay", "-lm",
                        "-Wl,-rpath," + HERE, "-Wl,-rpath,$ORIGIN/../symmetry_b200"], check=True)
    return DRIVER_SO


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose=True)
    build_driver(force=True)
    print(SO)
