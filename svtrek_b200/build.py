"""In-tree build of libsvtk_b200.so (hand-written CUDA for sm_100a + the C ABI). No JIT cache: the .so lives next
to this file so it travels to the GPU box with the repo snapshot."""
from __future__ import annotations

import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.environ.get("SVTK_LIB_OUT") or os.path.join(HERE, "libsvtk_b200.so")
SOURCES = ["svtk_kernels.cu", "svtk_index.cu", "svtk_poa.cu", "svtk_bam.cu", "svtk_ins.cu", "svtk_static.cu", "svtk_capi.cpp", "svtk_host.cpp"]
HEADERS = ["svtk_device.cuh", "svtk_poa_core.cuh", "svtk_bgzf_core.cuh", "svtk_launch.h", "svtk_ctx.h", os.path.join("..", "..", "include", "svtk.h")]
NVCC_FLAGS = ["-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-Xcompiler", "-fPIC",
              "-Xcompiler", "-Wall", "-shared", "-cudart", "shared"]


def nvcc() -> str:
    for cand in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


CLI = os.path.join(HERE, "svtrek_gpu")          # the C host driver (csrc/svtrek_gpu.c): `svtrek audt` over the C ABI
CLI_SRC = os.path.join(CSRC, "svtrek_gpu.c")


def stale() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(os.path.join(CSRC, f)) > t for f in SOURCES + HEADERS)


def build_cli() -> str:
    """gcc: the driver links libsvtk_b200.so next to it (rpath $ORIGIN)."""
    if os.path.exists(CLI) and os.path.getmtime(CLI) >= max(os.path.getmtime(CLI_SRC), os.path.getmtime(os.path.join(HERE, "libsvtk_b200.so"))):
        return CLI
    cmd = ["gcc", "-O2", "-std=c99", "-D_GNU_SOURCE", "-Wall", "-pthread", "-o", CLI, CLI_SRC, "-L" + HERE, "-lsvtk_b200", "-Wl,-rpath,$ORIGIN"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("gcc failed:\n" + r.stdout + r.stderr)
    return CLI


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not stale():
        if LIB == os.path.join(HERE, "libsvtk_b200.so"):
            build_cli()
        return LIB
    cmd = [nvcc(), *NVCC_FLAGS, *os.environ.get("SVTK_NVCC_EXTRA", "").split(), "-o", LIB, *[os.path.join(CSRC, s) for s in SOURCES]]
    if verbose:
        cmd.insert(1, "-Xptxas"); cmd.insert(2, "-v")
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + r.stdout + r.stderr)
    if verbose:
        print(r.stderr)
    if LIB == os.path.join(HERE, "libsvtk_b200.so"):
        build_cli()
    return LIB


if __name__ == "__main__":
    import sys
    print(build(force=True, verbose="-v" in sys.argv))
