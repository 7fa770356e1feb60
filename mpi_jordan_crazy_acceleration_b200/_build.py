"""In-tree native build: nvcc (sm_100a) for the CUDA engine, g++ for the CPU-only host library, the
CLI, and the MPI shim.  Artefacts land in ``mpi_jordan_crazy_acceleration_b200/lib`` and ``bin/`` so
that they travel with the source tree (nothing is installed into site-packages).

Usage:  python -m mpi_jordan_crazy_acceleration_b200._build [--force] [--only host|cuda|cli|shim]
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
PKG = Path(__file__).resolve().parent
LIBDIR = PKG / "lib"
BINDIR = ROOT / "bin"
OBJDIR = ROOT / "build" / "obj"

NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
CXX = os.environ.get("CXX", "g++")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
NVCC_FLAGS = ["-std=c++17", "-O3", "-lineinfo", "-Xcompiler", "-fPIC", "-Xptxas", "-v", *ARCH]
CXX_FLAGS = ["-std=c++17", "-O2", "-fPIC", "-Wall", "-I" + os.environ.get("CUDA_HOME", "/usr/local/cuda") + "/include"]

HOST_SRCS = ["csrc/host/layout.cpp", "csrc/host/io_text.cpp", "csrc/host/cpu_oracle.cpp", "csrc/host/plan_enum.cpp"]
CUDA_INC = os.environ.get("CUDA_HOME", "/usr/local/cuda") + "/include"
CUDA_SRCS = [
    "csrc/block_inverse.cu",
    "csrc/step_fused.cu",
    "csrc/fill_norm.cu",
    "csrc/dist_gemm.cu",
    "csrc/engine.cu",
]
# host-only translation units that need the CUDA headers: part of libjordan_b200.so only
CUDA_HOST_SRCS = ["csrc/comm/symm_heap.cpp"]
CUDA_HDRS = ["csrc/common.cuh", "csrc/kernels.h", "csrc/kernels_formula.h", "csrc/engine_types.h", "csrc/step_plan.h",
             "include/jordan_b200.h"]


def _newer(target: Path, deps: list[Path]) -> bool:
    if not target.exists():
        return True
    t = target.stat().st_mtime
    return any(d.exists() and d.stat().st_mtime > t for d in deps)


def _run(cmd: list[str], log: Path | None = None) -> None:
    res = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True)
    if log is not None:
        log.parent.mkdir(parents=True, exist_ok=True)
        log.write_text(" ".join(cmd) + "\n" + res.stdout + res.stderr)
    if res.returncode != 0:
        sys.stderr.write(res.stdout + res.stderr)
        raise RuntimeError(f"build step failed: {' '.join(cmd)}")


def build_host(force: bool = False) -> Path:
    LIBDIR.mkdir(parents=True, exist_ok=True)
    out = LIBDIR / "libjordan_host.so"
    deps = [ROOT / s for s in HOST_SRCS] + [ROOT / "include/jordan_b200.h", ROOT / "csrc/kernels_formula.h",
                                            ROOT / "csrc/step_plan.h", ROOT / "csrc/kernels.h"]
    if force or _newer(out, deps):
        _run([CXX, *CXX_FLAGS, "-shared", "-o", str(out), *HOST_SRCS])
    return out


def build_cuda(force: bool = False) -> Path:
    """Compile every CUDA translation unit for sm_100a and link libjordan_b200.so."""
    LIBDIR.mkdir(parents=True, exist_ok=True)
    OBJDIR.mkdir(parents=True, exist_ok=True)
    out = LIBDIR / "libjordan_b200.so"
    hdrs = [ROOT / h for h in CUDA_HDRS]
    objs = []
    extra = sorted(str(p.relative_to(ROOT)) for p in (ROOT / "csrc").glob("*.cu"))
    srcs = list(dict.fromkeys(CUDA_SRCS + extra))
    for src in srcs:
        obj = OBJDIR / (Path(src).stem + ".o")
        if force or _newer(obj, [ROOT / src, *hdrs]):
            _run([NVCC, *NVCC_FLAGS, "-c", src, "-o", str(obj)], log=OBJDIR / (Path(src).stem + ".ptxas.log"))
        objs.append(str(obj))
    host_objs = []
    for src in HOST_SRCS + CUDA_HOST_SRCS:
        obj = OBJDIR / (Path(src).stem + ".host.o")
        if force or _newer(obj, [ROOT / src, *hdrs]):
            _run([CXX, *CXX_FLAGS, "-c", src, "-o", str(obj)])
        host_objs.append(str(obj))
    if force or _newer(out, [Path(o) for o in objs + host_objs]):
        _run([NVCC, "-shared", *ARCH, "-o", str(out), *objs, *host_objs])
    return out


def build_shim(force: bool = False) -> Path:
    BINDIR.mkdir(parents=True, exist_ok=True)
    out = BINDIR / "jb_mpirun"
    srcs = ["mpi_shim/mpi_shim.cpp", "mpi_shim/mpirun.cpp"]
    if not (ROOT / srcs[0]).exists():
        return out
    lib = LIBDIR / "libjbmpi.a"
    obj = OBJDIR / "mpi_shim.o"
    OBJDIR.mkdir(parents=True, exist_ok=True)
    LIBDIR.mkdir(parents=True, exist_ok=True)
    if force or _newer(lib, [ROOT / srcs[0], ROOT / "mpi_shim/mpi.h"]):
        _run([CXX, *CXX_FLAGS, "-Impi_shim", "-c", srcs[0], "-o", str(obj)])
        if lib.exists():
            lib.unlink()
        _run(["ar", "rcs", str(lib), str(obj)])
    if force or _newer(out, [ROOT / srcs[1]]):
        _run([CXX, *CXX_FLAGS, "-o", str(out), srcs[1]])
    return out


def build_cli(force: bool = False) -> Path:
    BINDIR.mkdir(parents=True, exist_ok=True)
    out = BINDIR / "jordan"
    src = ROOT / "csrc/cli/main.cpp"
    if not src.exists():
        return out
    build_shim(force)
    cuda_lib = build_cuda(force)
    if force or _newer(out, [src, cuda_lib, LIBDIR / "libjbmpi.a"]):
        _run([NVCC, "-std=c++17", "-O2", *ARCH, "-Impi_shim", "-Iinclude", "-o", str(out), str(src),
              f"-L{LIBDIR}", "-ljordan_b200", "-ljbmpi", "-lcudart",
              "-Xlinker", f"-rpath={LIBDIR}", "-Xlinker", "-rpath=$ORIGIN/../mpi_jordan_crazy_acceleration_b200/lib"])
    return out


def build_all(force: bool = False) -> dict:
    res = {"host": str(build_host(force))}
    if shutil.which(NVCC) or Path(NVCC).exists():
        res["cuda"] = str(build_cuda(force))
        res["shim"] = str(build_shim(force))
        res["cli"] = str(build_cli(force))
    else:  # pragma: no cover - the image always has nvcc
        raise RuntimeError("nvcc not found; cannot build the sm_100a engine")
    return res


if __name__ == "__main__":
    force = "--force" in sys.argv
    only = None
    if "--only" in sys.argv:
        only = sys.argv[sys.argv.index("--only") + 1]
    if only == "host":
        print(build_host(force))
    elif only == "cuda":
        print(build_cuda(force))
    elif only == "shim":
        print(build_shim(force))
    elif only == "cli":
        print(build_cli(force))
    else:
        for k, v in build_all(force).items():
            print(f"{k}: {v}")
