"""Builds the CUDA extension in-tree for sm_100a: ``python -m msprime_b200.build``.

nvcc cross-compiles without a GPU.  ``-fmad=false`` keeps the arithmetic of the
replay harness identical to the reference's (no fused multiply-add contraction)."""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "csrc", "mspb_api.cu")
OUT = os.path.join(HERE, "_mspb.so")
DEPS = sorted(os.path.join(HERE, "csrc", f) for f in os.listdir(os.path.join(HERE, "csrc"))
              if f.endswith((".cu", ".cuh"))) + [os.path.join(HERE, "..", "include", "msprime_b200.h")]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
    "--extended-lambda", "-fmad=false", "-Xcompiler", "-fPIC", "-shared",
]


TREEGEN_SRC = os.path.join(HERE, "csrc", "treegen.c")
TREEGEN_OUT = os.path.join(HERE, "_treegen.so")


def needs_build() -> bool:
    if not os.path.exists(OUT):
        return True
    t = os.path.getmtime(OUT)
    return any(os.path.exists(d) and os.path.getmtime(d) > t for d in DEPS)


def build_treegen(force: bool = False) -> str:
    """The synthetic-input generator (plain C, benchmark / test inputs only)."""
    if force or not os.path.exists(TREEGEN_OUT) or os.path.getmtime(TREEGEN_SRC) > os.path.getmtime(TREEGEN_OUT):
        cc = os.environ.get("CC", "gcc")
        subprocess.run([cc, "-O2", "-std=c99", "-fPIC", "-shared", "-o", TREEGEN_OUT, TREEGEN_SRC, "-lm"], check=True)
    return TREEGEN_OUT


TREESIO_SRC = os.path.join(HERE, "csrc", "trees_io.c")
TREESIO_OUT = os.path.join(HERE, "_treesio.so")
TREESIO_DEPS = [TREESIO_SRC, os.path.join(HERE, "..", "include", "msprime_b200_io.h"),
                os.path.join(HERE, "..", "include", "msprime_b200.h")]


def build_treesio(force: bool = False) -> str:
    """The `.trees` (kastore) reader / writer: host-only C (include/msprime_b200_io.h)."""
    if force or not os.path.exists(TREESIO_OUT) or any(
            os.path.getmtime(d) > os.path.getmtime(TREESIO_OUT) for d in TREESIO_DEPS):
        cc = os.environ.get("CC", "gcc")
        subprocess.run([cc, "-O2", "-std=c11", "-Wall", "-Wextra", "-fPIC", "-shared", "-o", TREESIO_OUT,
                        TREESIO_SRC], check=True)
    return TREESIO_OUT


def build(force: bool = False, verbose: bool = False) -> str:
    if force or needs_build():
        nvcc = os.environ.get("NVCC", "nvcc")
        cmd = [nvcc, *NVCC_FLAGS, "-o", OUT, SRC]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
            print(" ".join(cmd), file=sys.stderr)
        subprocess.run(cmd, check=True)
    build_treegen(force)
    build_treesio(force)
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
