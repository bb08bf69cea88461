"""Builds libzebra_tppr.so in-tree with nvcc for sm_100a (cross-compiles without a GPU)."""
from __future__ import annotations

import glob
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIBDIR = os.path.join(HERE, "lib")
SO = os.path.join(LIBDIR, "libzebra_tppr.so")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
    "-fmad=false",              # the reference's fp64 arithmetic has no fused multiply-add
    "-Xcompiler", "-fPIC", "-Xcompiler", "-ffp-contract=off", "-shared",
]


def sources():
    return sorted(glob.glob(os.path.join(CSRC, "*.cu")))


def stale() -> bool:
    if not os.path.exists(SO):
        return True
    t = os.path.getmtime(SO)
    deps = glob.glob(os.path.join(CSRC, "*")) + glob.glob(os.path.join(HERE, "..", "include", "*.h"))
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False, profile: bool = False, ablate: int = 0) -> str:
    """profile=True builds the developer variant with per-phase cycle counters next to the product library."""
    out = SO.replace(".so", "_prof.so") if profile else SO
    if ablate:
        out = SO.replace(".so", f"_abl{ablate}.so")
    if not force and not profile and not ablate and not stale():
        return SO
    os.makedirs(LIBDIR, exist_ok=True)
    nvcc = os.environ.get("NVCC") or ("/usr/local/cuda/bin/nvcc" if os.path.exists("/usr/local/cuda/bin/nvcc") else "nvcc")
    cmd = ([nvcc] + NVCC_FLAGS + (["-DZEBRA_PROFILE"] if profile else []) + ([f"-DZEBRA_ABLATE={ablate}"] if ablate else []) + (["-Xptxas", "-v"] if verbose else [])
           + ["-o", out] + sources())
    subprocess.check_call(cmd)
    return out


if __name__ == "__main__":
    import sys
    abl = [int(a.split("=")[1]) for a in sys.argv if a.startswith("--ablate=")]
    print(build(force=True, verbose=True, profile="--profile" in sys.argv, ablate=abl[0] if abl else 0))
