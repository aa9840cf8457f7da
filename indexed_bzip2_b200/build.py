"""Builds the in-tree native library indexed_bzip2_b200/libibz2_cuda.so (sm_100a only) with nvcc.

    python -m indexed_bzip2_b200.build [--force] [--verbose]

nvcc cross-compiles without a GPU, so this runs in the CPU-only build container; the resulting .so travels to the
GPU box with the repo snapshot (it is git-ignored, not gpurun-ignored).
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

PKG = Path(__file__).resolve().parent
CSRC = PKG / "csrc"
BUILD = PKG / "build"
LIB = PKG / "libibz2_cuda.so"
HOST_LIB = PKG / "libibz2_host.so"     # C++ readers + flat C entry points for the Python module
CLI = PKG / "bin" / "ibzip2"           # command-line tool
HOST_FLAGS = ["-std=c++17", "-O2", "-g", "-fPIC", "-Wall", "-Wextra", "-pthread", "-rdynamic"]

CUDA_SOURCES = ["blockfinder.cu", "huffman_header.cu", "huffman_bits.cu", "huffman_careful.cu", "mtf.cu", "bwt_prepare.cu", "bwt_traverse.cu", "rle.cu", "crc.cu", "capi.cu", "multi.cu"]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    "-Xcompiler", "-fPIC,-O2,-Wall,-Wno-unused-function",
    "-Xptxas", "-v",
]


def nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    raise RuntimeError("nvcc not found (needed to build libibz2_cuda.so; there is no CPU fallback)")


def _stale(target: Path, deps: list[Path]) -> bool:
    if not target.exists():
        return True
    t = target.stat().st_mtime
    return any(d.stat().st_mtime > t for d in deps)


def build_native(force: bool = False, verbose: bool = False) -> Path:
    BUILD.mkdir(exist_ok=True)
    headers = list(CSRC.glob("*.cuh")) + list(CSRC.glob("*.inl")) + list((PKG.parent / "include").glob("*.h")) + [Path(__file__)]
    cc = nvcc()
    # objects built with other experiment flags ($IBZ2_NVCC_EXTRA) are stale whatever their age
    stamp = BUILD / "nvcc_extra.txt"
    extra = os.environ.get("IBZ2_NVCC_EXTRA", "")
    if (stamp.read_text() if stamp.exists() else "") != extra:
        force = True
        stamp.write_text(extra)
    jobs = []
    for name in CUDA_SOURCES:
        src = CSRC / name
        obj = BUILD / (src.stem + ".o")
        if force or _stale(obj, [src] + headers):
            jobs.append((src, obj))

    def compile_one(job):
        src, obj = job
        # IBZ2_NVCC_EXTRA: extra flags for experiments (-D switches of the kernels), e.g. on the GPU box
        cmd = [cc, *NVCC_FLAGS, *os.environ.get("IBZ2_NVCC_EXTRA", "").split(), "-c", str(src), "-o", str(obj)]
        res = subprocess.run(cmd, capture_output=True, text=True)
        (BUILD / (src.stem + ".ptxas.log")).write_text(res.stderr)
        if res.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src.name}:\n{res.stdout}\n{res.stderr}")
        if verbose:
            sys.stderr.write(res.stderr)
        return obj

    if jobs:
        with ThreadPoolExecutor(max_workers=min(len(jobs), os.cpu_count() or 4)) as pool:
            list(pool.map(compile_one, jobs))
    objs = [BUILD / (Path(n).stem + ".o") for n in CUDA_SOURCES]
    if force or jobs or _stale(LIB, objs):
        cmd = [cc, "-shared", "-o", str(LIB), *map(str, objs), "-gencode", "arch=compute_100a,code=sm_100a",
               "-Xcompiler", "-fPIC", "-cudart", "shared"]
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError(f"link failed:\n{res.stdout}\n{res.stderr}")
    build_host(force=force or bool(jobs))
    return LIB


def build_host(force: bool = False) -> None:
    """The host side above the C ABI: readers (header-only C++), their C wrapper and the CLI, compiled with g++."""
    # the system g++ (dynamic libstdc++); $CXX in this image points at a toolchain that links libstdc++ statically, which
    # gives the .so a second, uninitialised copy of the iostream locale machinery inside a Python process
    cxx = os.environ.get("IBZ2_CXX") or ("/usr/bin/g++" if Path("/usr/bin/g++").exists() else shutil.which("g++") or "g++")
    host = PKG / "host"
    tools = PKG / "tools"
    deps = list(host.glob("*.hpp")) + list(host.glob("*.cpp")) + list(tools.glob("*.cpp")) + [
        PKG.parent / "include" / "ibz2_cuda.h", LIB, Path(__file__)]
    link = ["-L" + str(PKG), "-libz2_cuda", "-Wl,-rpath,$ORIGIN"]
    if force or _stale(HOST_LIB, deps):
        cmd = [cxx, *HOST_FLAGS, "-shared", "-DIBZ2_NO_MAIN", str(host / "reader_capi.cpp"), str(tools / "ibzip2.cpp"),
               "-o", str(HOST_LIB), *link]
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError(f"host library build failed:\n{res.stdout}\n{res.stderr}")
    if force or _stale(CLI, deps):
        CLI.parent.mkdir(exist_ok=True)
        cmd = [cxx, *HOST_FLAGS, str(tools / "ibzip2.cpp"), "-o", str(CLI), "-L" + str(PKG), "-libz2_cuda",
               "-Wl,-rpath,$ORIGIN/.."]
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError(f"ibzip2 CLI build failed:\n{res.stdout}\n{res.stderr}")


if __name__ == "__main__":
    lib = build_native(force="--force" in sys.argv, verbose="--verbose" in sys.argv)
    print(lib)
