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
SOURCES = ["dgs_api.cu", "dgs_flatten.cu", "dgs_norm.cu", "dgs_noalloc.cu", "dgs_step.cu", "dgs_metric.cu", "dgs_comm.cu", "dgs_peer.cu", "dgs_direct.cu"]
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


def host_ext_path() -> pathlib.Path:
    import sysconfig

    return CSRC / ("_dgs_host" + (sysconfig.get_config_var("EXT_SUFFIX") or ".so"))


def build_host(force: bool = False) -> pathlib.Path:
    """g++ csrc/dgs_host.cpp -> csrc/_dgs_host.<abi>.so: the optional native host loop of the engine (a torch C++
    extension; no CUDA code in it).  About 40 s of compile time, almost all of it torch's headers."""
    import sysconfig

    import torch
    from torch.utils import cpp_extension

    out, src = host_ext_path(), CSRC / "dgs_host.cpp"
    if not force and out.exists() and out.stat().st_mtime >= src.stat().st_mtime:
        return out
    tmp = out.with_suffix(".tmp")
    libs = cpp_extension.library_paths()
    cmd = ["g++", "-O2", "-std=c++17", "-shared", "-fPIC", str(src), "-o", str(tmp), "-DTORCH_EXTENSION_NAME=_dgs_host",
           f"-D_GLIBCXX_USE_CXX11_ABI={int(torch._C._GLIBCXX_USE_CXX11_ABI)}", "-I" + sysconfig.get_paths()["include"],
           *("-I" + d for d in cpp_extension.include_paths()), *("-L" + d for d in libs),
           "-ltorch", "-ltorch_cpu", "-ltorch_python", "-lc10", *("-Wl,-rpath," + d for d in libs)]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"g++ failed on dgs_host.cpp:\n{r.stdout}\n{r.stderr}")
    os.replace(tmp, out)
    return out


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
