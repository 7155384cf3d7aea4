"""In-tree build of the native libraries (no JIT cache: the built .so travels with the repo snapshot).

* ``libdalle_b200.so``   - sm_100a CUDA kernels behind a C ABI (include/b200_api.h), built with nvcc.
* ``libdalle_b200_host.so`` - host-side C++ (block-wise 8-bit quantisation, CPU LAMB, compression, TCP transport).

``python -m dalle_b200.ops.build`` builds both; ``build_all()`` is what ``__graft_entry__.build()`` calls.
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

ROOT = Path(__file__).resolve().parents[2]
CSRC = Path(__file__).resolve().parent / "csrc"
HOST_SRC = ROOT / "dalle_b200" / "native"
OUT_DIR = Path(__file__).resolve().parent
BUILD_DIR = ROOT / "build"
CUDA_LIB = OUT_DIR / "libdalle_b200.so"
HOST_LIB = OUT_DIR / "libdalle_b200_host.so"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "--expt-relaxed-constexpr", "-Xcompiler", "-fPIC", "-I", str(ROOT / "include"), "-I", str(CSRC),
]
NVCC_FLAGS += os.environ.get("B200_NVCC_EXTRA", "").split()
CXX_FLAGS = ["-O3", "-std=c++17", "-fPIC", "-fopenmp", "-march=x86-64-v3", "-I", str(ROOT / "include")]


def _nvcc() -> str:
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        raise RuntimeError("nvcc not found; cannot build the sm_100a kernel library")
    return nvcc


def _digest(paths, flags) -> str:
    h = hashlib.sha256()
    for p in sorted(paths):
        h.update(p.name.encode())
        h.update(p.read_bytes())
    h.update(" ".join(flags).encode())
    return h.hexdigest()


def _run(cmd):
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError(f"build failed: {' '.join(map(str, cmd))}\n{res.stdout}\n{res.stderr}")
    return res.stdout + res.stderr


def build_cuda(force: bool = False, verbose: bool = False) -> Path:
    srcs = sorted(CSRC.glob("*.cu"))
    deps = srcs + sorted(CSRC.glob("*.cuh")) + sorted(CSRC.glob("*.inc")) + sorted((ROOT / "include").glob("*.h"))
    stamp = BUILD_DIR / "cuda.stamp"
    dig = _digest(deps, NVCC_FLAGS)
    if not force and CUDA_LIB.exists() and stamp.exists() and stamp.read_text() == dig:
        return CUDA_LIB
    BUILD_DIR.mkdir(exist_ok=True)
    nvcc = _nvcc()
    objs = []

    def compile_one(src: Path):
        obj = BUILD_DIR / (src.stem + ".o")
        # per-file incremental: skip when neither the file nor any header changed
        fstamp = BUILD_DIR / (src.stem + ".stamp")
        fdig = _digest([src] + sorted(CSRC.glob("*.cuh")) + sorted(CSRC.glob("*.inc")) + sorted((ROOT / "include").glob("*.h")), NVCC_FLAGS)
        if not force and obj.exists() and fstamp.exists() and fstamp.read_text() == fdig:
            return obj
        extra = ["-Xptxas", "-v"] if verbose else []
        out = _run([nvcc, *NVCC_FLAGS, *extra, "-c", str(src), "-o", str(obj)])
        if verbose:
            print(out)
        fstamp.write_text(fdig)
        return obj

    with ThreadPoolExecutor(max_workers=min(8, max(1, len(srcs)))) as ex:
        objs = list(ex.map(compile_one, srcs))
    _run([nvcc, "-shared", "-o", str(CUDA_LIB), *map(str, objs), "-lcudart"])
    stamp.write_text(dig)
    return CUDA_LIB


def build_host(force: bool = False) -> Path:
    srcs = sorted(HOST_SRC.glob("*.cpp"))
    if not srcs:
        return HOST_LIB
    deps = srcs + sorted(HOST_SRC.glob("*.h")) + sorted((ROOT / "include").glob("*.h"))
    stamp = BUILD_DIR / "host.stamp"
    dig = _digest(deps, CXX_FLAGS)
    if not force and HOST_LIB.exists() and stamp.exists() and stamp.read_text() == dig:
        return HOST_LIB
    BUILD_DIR.mkdir(exist_ok=True)
    _run(["g++", *CXX_FLAGS, "-shared", "-o", str(HOST_LIB), *map(str, srcs), "-lpthread"])
    stamp.write_text(dig)
    return HOST_LIB


def build_all(force: bool = False, verbose: bool = False):
    return build_cuda(force, verbose), build_host(force)


if __name__ == "__main__":
    libs = build_all(force="--force" in sys.argv, verbose="-v" in sys.argv)
    for lib in libs:
        print(lib)
