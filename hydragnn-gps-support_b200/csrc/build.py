"""Build libhgk.so (sm_100a only) in-tree with nvcc.

Usage: python build.py [--force] [--verbose]
The shared object lands next to the python package
(hydragnn-gps-support_b200/libhgk.so) so that it travels with the repo snapshot.
"""
import hashlib
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.dirname(HERE)
OUT = os.path.join(PKG, "libhgk.so")
STAMP = os.path.join(PKG, ".libhgk.stamp")

NVCC_FLAGS = [
    "-O3",
    "-std=c++17",
    "-lineinfo",
    "-gencode",
    "arch=compute_100a,code=sm_100a",
    "--shared",
    "-Xcompiler",
    "-fPIC",
    "--threads",
    "0",
]


def sources():
    return sorted(
        os.path.join(HERE, f) for f in os.listdir(HERE) if f.endswith(".cu")
    )


def _digest():
    h = hashlib.sha256()
    names = sources() + sorted(
        os.path.join(HERE, f) for f in os.listdir(HERE) if f.endswith((".cuh", ".h"))
    )
    names.append(os.path.join(os.path.dirname(PKG), "include", "hgk.h"))
    for n in names:
        h.update(os.path.basename(n).encode())   # location-independent: the .so travels
        with open(n, "rb") as fh:
            h.update(fh.read())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def find_nvcc():
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found; libhgk.so cannot be built")


def build(force=False, verbose=False):
    digest = _digest()
    if not force and os.path.exists(OUT) and os.path.exists(STAMP):
        with open(STAMP) as fh:
            if fh.read().strip() == digest:
                return OUT
    cmd = [find_nvcc()] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else [])
    cmd += ["-o", OUT] + sources()
    res = subprocess.run(cmd, capture_output=True, text=True)
    if verbose or res.returncode != 0:
        sys.stderr.write(res.stdout + res.stderr)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed building libhgk.so")
    with open(STAMP, "w") as fh:
        fh.write(digest)
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv))
