"""In-tree build of libyoda_b200.so (sm_100a only; nvcc cross-compiles without a GPU)."""
from __future__ import annotations

import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
# YODA_B200_SO: load another build of the library instead (kernel-tuning experiments: variants compiled by hand with -D flags)
SO = os.environ.get("YODA_B200_SO") or os.path.join(HERE, "libyoda_b200.so")
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "-shared",
]


def sources():
    return sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cu", ".cuh", ".h")))


def stale() -> bool:
    if os.environ.get("YODA_B200_SO"):
        return False  # a hand-built variant: never rebuilt behind the experimenter's back
    if not os.path.exists(SO):
        return True
    t = os.path.getmtime(SO)
    deps = sources() + [os.path.join(HERE, "..", "include", "yoda_b200.h")]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not stale():
        return SO
    nvcc = os.environ.get("NVCC", "nvcc")
    cus = [s for s in sources() if s.endswith(".cu")]
    cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-o", SO] + cus + ["-ldl"]
    subprocess.check_call(cmd, cwd=CSRC)
    return SO


HOST = os.path.join(HERE, "host")
HARNESS = os.path.join(HOST, "yoda_harness")


def build_host(force: bool = False) -> str:
    """C++ host layer: plugin mirror + fake-framework benchmark harness (links libyoda_b200.so)."""
    srcs = [os.path.join(HOST, f) for f in ("harness.cpp", "yoda_plugin.cpp")]
    deps = [os.path.join(HOST, f) for f in os.listdir(HOST) if f.endswith((".cpp", ".hpp"))] + [SO]
    if not force and os.path.exists(HARNESS) and all(os.path.getmtime(d) <= os.path.getmtime(HARNESS) for d in deps):
        return HARNESS
    cmd = [os.environ.get("CXX", "g++"), "-std=c++17", "-O2", "-Wall", "-o", HARNESS] + srcs + ["-L" + HERE, "-lyoda_b200", "-Wl,-rpath,$ORIGIN/..", "-lpthread"]
    subprocess.check_call(cmd, cwd=HOST)
    return HARNESS
