"""In-tree build of the native libraries (no JIT cache: the .so files travel with the snapshot).

libfastapriori_b200.so — the sm_100a kernels + C-ABI (include/fastapriori_b200.h), nvcc.
libfa_quest.so         — the synthetic workload generator (host only), g++.
"""
from __future__ import annotations

import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libfastapriori_b200.so")
QUEST = os.path.join(HERE, "libfa_quest.so")
MAIN = os.path.join(HERE, "fa_main")
HOST_CXX = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else (shutil.which("g++") or "g++")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
    "-Xcompiler", "-fPIC,-fopenmp", "-shared", "-ccbin", HOST_CXX, "-Wno-deprecated-declarations", "-lgomp",
]


def _newer(target: str, sources: list[str]) -> bool:
    if not os.path.exists(target):
        return False
    t = os.path.getmtime(target)
    return all(os.path.getmtime(s) <= t for s in sources)


def _sources(exts) -> list[str]:
    out = [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC)) if f.endswith(exts)]
    out.append(os.path.join(HERE, "..", "include", "fastapriori_b200.h"))
    return out


def build_cuda(force: bool = False, verbose: bool = False) -> str:
    srcs = _sources((".cu", ".cuh", ".inl"))
    if not force and _newer(LIB, srcs):
        return LIB
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    cmd = [nvcc, *NVCC_FLAGS, "-o", LIB, os.path.join(CSRC, "fa_lib.cu")]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    subprocess.run(cmd, check=True, cwd=CSRC)
    return LIB


def build_quest(force: bool = False) -> str:
    src = os.path.join(CSRC, "quest_gen.cpp")
    if not force and _newer(QUEST, [src]):
        return QUEST
    subprocess.run([HOST_CXX, "-O2", "-std=c++17", "-fPIC", "-fopenmp", "-shared", "-o", QUEST, src], check=True)
    return QUEST


def build_main(force: bool = False) -> str:
    """fa_main: the C++ twin of com.hazzacheng.AR.Main, linked against the in-tree libfastapriori_b200.so."""
    src = os.path.join(HERE, "host", "fa_main.cpp")
    if not force and _newer(MAIN, [src, LIB, os.path.join(HERE, "..", "include", "fastapriori_b200.h")]):
        return MAIN
    subprocess.run([HOST_CXX, "-O2", "-std=c++17", "-o", MAIN, src, "-L" + HERE, "-lfastapriori_b200",
                    "-Wl,-rpath,$ORIGIN"], check=True)
    return MAIN


def build_all(force: bool = False) -> None:
    build_cuda(force)
    build_quest(force)
    build_main(force)


if __name__ == "__main__":
    import sys
    build_all(force="--force" in sys.argv)
    print(LIB, QUEST, sep="\n")
