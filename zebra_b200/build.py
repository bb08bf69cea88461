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


def source_digest() -> str:
    """SHA-256 over the contents of csrc/ and include/ (and the flags): what the shipped .so must match.
    Content, not mtimes: a snapshot copied to the GPU box keeps the bytes but not necessarily the times."""
    import hashlib
    h = hashlib.sha256(" ".join(NVCC_FLAGS).encode())
    for d in sorted(glob.glob(os.path.join(CSRC, "*")) + glob.glob(os.path.join(HERE, "..", "include", "*.h"))):
        h.update(os.path.basename(d).encode())
        h.update(open(d, "rb").read())
    return h.hexdigest()


def stale() -> bool:
    stamp = SO + ".stamp"
    if not os.path.exists(SO) or not os.path.exists(stamp):
        return True
    return open(stamp).read().strip() != source_digest()


def build(force: bool = False, verbose: bool = False, profile: bool = False, ablate: int = 0) -> str:
    """profile=True builds the developer variant with per-phase cycle counters next to the product library."""
    out = SO.replace(".so", "_prof.so") if profile else SO
    if ablate:
        out = SO.replace(".so", f"_abl{ablate}.so")
    if not force and not profile and not ablate and not stale():
        return SO
    os.makedirs(LIBDIR, exist_ok=True)
    nvcc = os.environ.get("NVCC") or ("/usr/local/cuda/bin/nvcc" if os.path.exists("/usr/local/cuda/bin/nvcc") else "nvcc")
    import fcntl
    with open(os.path.join(LIBDIR, ".build.lock"), "w") as lock:      # several ranks may import at once
        fcntl.flock(lock, fcntl.LOCK_EX)
        if not force and not profile and not ablate and not stale():
            return SO                                                 # another process built it meanwhile
        tmp = out + f".tmp{os.getpid()}"
        cmd = ([nvcc] + NVCC_FLAGS + (["-DZEBRA_PROFILE"] if profile else []) + ([f"-DZEBRA_ABLATE={ablate}"] if ablate else [])
               + (["-Xptxas", "-v"] if verbose else []) + ["-o", tmp] + sources())
        subprocess.check_call(cmd)
        os.replace(tmp, out)
        if out == SO:
            with open(SO + ".stamp", "w") as f:
                f.write(source_digest())
    return out


if __name__ == "__main__":
    import sys
    abl = [int(a.split("=")[1]) for a in sys.argv if a.startswith("--ablate=")]
    print(build(force=True, verbose=True, profile="--profile" in sys.argv, ablate=abl[0] if abl else 0))
