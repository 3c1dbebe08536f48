"""Build recipes for the native libraries (nvcc / g++ invoked directly, outputs stay in-tree).

    python -m pytorch_cpp_rl_b200.build            # builds lib/libcpprl_b200.so (C ABI + sm_100a kernels)

The kernels are compiled for sm_100a only (`-gencode arch=compute_100a,code=sm_100a`), with -lineinfo so
ncu's source page maps back to the .cu files.  nvcc cross-compiles without a GPU.
"""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

PKG = Path(__file__).resolve().parent
ROOT = PKG.parent
CSRC = PKG / "csrc"
LIB = PKG / "lib"
OBJ = PKG / "build"
INCLUDE = ROOT / "include"

NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC,-O3,-Wall",
    "-I", str(INCLUDE), "-I", str(CSRC),
]

KERNEL_LIB = LIB / "libcpprl_b200.so"


def _newer(target: Path, deps) -> bool:
    if not target.exists():
        return False
    t = target.stat().st_mtime
    return all(Path(d).stat().st_mtime <= t for d in deps)


def _run(cmd, verbose):
    if verbose:
        print(" ".join(str(c) for c in cmd), flush=True)
    res = subprocess.run([str(c) for c in cmd], capture_output=True, text=True)
    if res.returncode != 0:
        sys.stderr.write(res.stdout + res.stderr)
        raise RuntimeError(f"command failed ({res.returncode}): {' '.join(str(c) for c in cmd)}")
    if verbose and (res.stdout or res.stderr):
        print(res.stdout + res.stderr, flush=True)
    return res


def kernel_sources():
    return sorted(CSRC.glob("*.cu"))


def build_kernels(force: bool = False, verbose: bool = False, ptxas_info: bool = False) -> Path:
    """Compile every .cu under csrc/ for sm_100a and link the C-ABI shared library."""
    LIB.mkdir(exist_ok=True)
    OBJ.mkdir(exist_ok=True)
    headers = list(CSRC.glob("*.cuh")) + list(INCLUDE.glob("*.h"))
    srcs = kernel_sources()
    if not force and _newer(KERNEL_LIB, srcs + headers + [Path(__file__)]):
        return KERNEL_LIB
    objs = []
    jobs = []
    for src in srcs:
        obj = OBJ / (src.stem + ".o")
        objs.append(obj)
        if force or not _newer(obj, [src] + headers + [Path(__file__)]):
            cmd = [NVCC, *NVCC_FLAGS, "-c", src, "-o", obj]
            if ptxas_info:
                cmd[1:1] = ["-Xptxas", "-v"]
            jobs.append(cmd)
    with ThreadPoolExecutor(max_workers=min(8, max(1, len(jobs)))) as ex:
        list(ex.map(lambda c: _run(c, verbose or ptxas_info), jobs))
    _run([NVCC, "-shared", "-o", KERNEL_LIB, *objs, "-gencode", "arch=compute_100a,code=sm_100a",
          "-Xcompiler", "-fPIC", "-cudart", "static"], verbose)
    return KERNEL_LIB


HOST_LIB = LIB / "libcpprl_host.so"
HOST_SRC = PKG / "host"
CXX = os.environ.get("CXX", "g++")


def torch_paths():
    import torch
    t = Path(torch.__file__).resolve().parent
    return t, [t / "include", t / "include" / "torch" / "csrc" / "api" / "include"], t / "lib"


def host_cxxflags():
    _, incs, _ = torch_paths()
    flags = ["-O2", "-std=c++17", "-fPIC", "-w", "-D_GLIBCXX_USE_CXX11_ABI=1", "-I", INCLUDE, "-I", HOST_SRC,
             "-I", "/usr/local/cuda/include"]
    for i in incs:
        flags += ["-I", i]
    return flags


def host_ldflags():
    _, _, tlib = torch_paths()
    return ["-L", LIB, "-lcpprl_b200", "-L", tlib, "-ltorch", "-ltorch_cpu", "-lc10", "-lc10_cuda", "-ltorch_cuda",
            f"-Wl,-rpath,{tlib}", "-Wl,-rpath,$ORIGIN", "-Wl,--no-as-needed"]


def build_host(force: bool = False, verbose: bool = False) -> Path:
    """The C++ drop-in (namespace cpprl, include/cpprl/*.h) over the C ABI: lib/libcpprl_host.so."""
    build_kernels(force=False, verbose=verbose)
    srcs = sorted(HOST_SRC.glob("*.cpp"))
    headers = list(HOST_SRC.glob("*.h")) + list(INCLUDE.rglob("*.h"))
    if not force and _newer(HOST_LIB, srcs + headers + [Path(__file__), KERNEL_LIB]):
        return HOST_LIB
    objs, jobs = [], []
    for src in srcs:
        obj = OBJ / ("host_" + src.stem + ".o")
        objs.append(obj)
        if force or not _newer(obj, [src] + headers + [Path(__file__)]):
            jobs.append([CXX, *host_cxxflags(), "-c", src, "-o", obj])
    with ThreadPoolExecutor(max_workers=min(8, max(1, len(jobs)))) as ex:
        list(ex.map(lambda c: _run(c, verbose), jobs))
    _run([CXX, "-shared", "-o", HOST_LIB, *objs, *host_ldflags()], verbose)
    return HOST_LIB


def build_all(force: bool = False, verbose: bool = False) -> None:
    build_kernels(force=force, verbose=verbose)
    build_host(force=force, verbose=verbose)


if __name__ == "__main__":
    build_kernels(force="--force" in sys.argv, verbose=True, ptxas_info="--ptxas" in sys.argv)
    print(f"built {KERNEL_LIB}")
    if "--host" in sys.argv:
        print(f"built {build_host(force='--force' in sys.argv, verbose=True)}")
