"""Build libdgs.so (sm_100a only) in-tree with nvcc.

The shared object is git-ignored but travels to the GPU box with the repo snapshot; nothing is JIT
compiled at run time.
"""
import concurrent.futures
import os
import pathlib
import shutil
import subprocess
import sys
from typing import List

CSRC = pathlib.Path(__file__).resolve().parent / "csrc"
LIB = CSRC / "libdgs.so"
SOURCES = ["dgs_api.cu", "dgs_flatten.cu", "dgs_norm.cu", "dgs_noalloc.cu", "dgs_step.cu", "dgs_metric.cu", "dgs_comm.cu"]
NVCC_FLAGS = [
    "-O3",
    "-std=c++17",
    "-gencode",
    "arch=compute_100a,code=sm_100a",
    "-lineinfo",
    "-Xcompiler",
    "-fPIC",
]


def _nvcc() -> str:
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        raise RuntimeError("nvcc not found; libdgs.so cannot be built")
    return nvcc


def _stale() -> bool:
    if not LIB.exists():
        return True
    newest = max(p.stat().st_mtime for p in list(CSRC.glob("*.cu")) + list(CSRC.glob("*.cuh")) + [CSRC.parent.parent / "include" / "dgs.h"])
    return newest > LIB.stat().st_mtime


def build(force: bool = False, verbose: bool = False) -> pathlib.Path:
    """Compile every CUDA source for sm_100a and link csrc/libdgs.so."""
    if not force and not _stale():
        return LIB
    nvcc = _nvcc()
    objdir = CSRC / "build"
    objdir.mkdir(exist_ok=True)
    extra = ["-Xptxas", "-v"] if verbose else []

    def compile_one(src: str) -> str:
        obj = objdir / (src[:-3] + ".o")
        cmd = [nvcc, *NVCC_FLAGS, *extra, "-c", str(CSRC / src), "-o", str(obj)]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed on {src}:\n{r.stdout}\n{r.stderr}")
        if verbose:
            sys.stderr.write(r.stderr)
        return str(obj)

    with concurrent.futures.ThreadPoolExecutor(max_workers=len(SOURCES)) as ex:
        objs: List[str] = list(ex.map(compile_one, SOURCES))
    tmp = LIB.with_suffix(".so.tmp")
    cmd = [nvcc, "-shared", "-o", str(tmp), *objs, "-ldl", "-Xlinker", "--exclude-libs,ALL"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    os.replace(tmp, LIB)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
