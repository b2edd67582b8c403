"""In-tree build of the native code (nvcc for sm_100a; g++ for the host driver).

Outputs go to particle_simulation_b200/_build/ (git-ignored; they travel to the
GPU box with the snapshot):
  libpsim_b200.so   the C-ABI library (include/psim.h)
  simu_b200         the C++ host: `simu_b200 <parameter-file> <particle-file> <output-file>`
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

PKG = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG)
CSRC = os.path.join(PKG, "csrc")
OUT = os.path.join(PKG, "_build")
LIB = os.environ.get("PSIM_LIBRARY") or os.path.join(OUT, "libpsim_b200.so")   # PSIM_LIBRARY: a differently built library (profiling)
SIMU = os.path.join(OUT, "simu_b200")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    "-Xcompiler", "-fPIC",
]


def _nvcc() -> str:
    for cand in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def _newer(target: str, sources: list[str]) -> bool:
    if not os.path.exists(target):
        return False
    t = os.path.getmtime(target)
    return all(os.path.getmtime(s) <= t for s in sources)


def _sources(dirs, exts):
    out = []
    for d in dirs:
        for base, _, files in os.walk(d):
            for f in files:
                if f.endswith(exts):
                    out.append(os.path.join(base, f))
    return sorted(out)


def build_library(force: bool = False, verbose: bool = False) -> str:
    os.makedirs(OUT, exist_ok=True)
    deps = _sources([CSRC], (".cu", ".cuh")) + [os.path.join(ROOT, "include", "psim.h")]
    deps = [d for d in deps if os.sep + "host" + os.sep not in d]
    if not force and _newer(LIB, deps):
        return LIB
    units = sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".cu"))
    cmd = [_nvcc(), *NVCC_FLAGS, "-shared", "-I", os.path.join(ROOT, "include"), "-o", LIB, *units, "-ldl"]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
        print(" ".join(cmd), file=sys.stderr)
    subprocess.run(cmd, check=True)
    return LIB


def build_host(force: bool = False) -> str | None:
    host_dir = os.path.join(CSRC, "host")
    if not os.path.isdir(host_dir):
        return None
    srcs = _sources([host_dir], (".cpp",))
    deps = srcs + _sources([host_dir], (".hpp",)) + [os.path.join(ROOT, "include", "psim.h"), LIB]
    if not force and _newer(SIMU, deps):
        return SIMU
    cmd = ["g++", "-std=c++17", "-O2", "-Wall", "-I", os.path.join(ROOT, "include"), "-o", SIMU, *srcs,
           "-L", OUT, "-lpsim_b200", "-Wl,-rpath,$ORIGIN", "-lpthread"]
    subprocess.run(cmd, check=True)
    return SIMU


def build_all(force: bool = False, verbose: bool = False):
    lib = build_library(force, verbose)
    simu = build_host(force)
    return lib, simu


if __name__ == "__main__":
    print(build_all(force="--force" in sys.argv, verbose="-v" in sys.argv))
