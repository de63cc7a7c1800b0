"""In-tree build of the B200 kernels (sm_100a only).

Produces, next to this file:
  libb200vllm.so  -- the C-ABI library (include/b200_vllm.h): nvcc, no torch dependency
  _C.so           -- the torch-library registration (csrc/torch_bindings.cpp) that exposes the
                     reference's op schemas (torch.ops._C / _C_cache_ops / _C_cuda_utils)

Both are git-ignored but travel to the GPU box with the gpurun snapshot.  nvcc cross-compiles
without a GPU.  Re-running is incremental (mtime based).
"""
from __future__ import annotations

import concurrent.futures as cf
import glob
import os
import shutil
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
INC = os.path.join(ROOT, "include")
OBJ = os.path.join(HERE, "build")
LIB = os.path.join(HERE, "libb200vllm.so")
EXT = os.path.join(HERE, "_C.so")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "--expt-relaxed-constexpr", "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden",
    f"-I{INC}", f"-I{CSRC}",
] + os.environ.get("B200_EXTRA_NVCC_FLAGS", "").split()


def _nvcc() -> str:
    for c in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", shutil.which("nvcc")):
        if c and os.path.exists(c):
            return c
    raise RuntimeError("nvcc not found")


def _newer(target: str, deps: list[str]) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def _headers() -> list[str]:
    return (glob.glob(os.path.join(CSRC, "*.cuh")) + glob.glob(os.path.join(CSRC, "*.h")) +
            glob.glob(os.path.join(CSRC, "*.inc")) + glob.glob(os.path.join(INC, "*.h")) +
            [os.path.abspath(__file__)])


def _run(cmd: list[str]) -> None:
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(" ".join(cmd) + "\n" + r.stdout + r.stderr)
        raise RuntimeError(f"build step failed: {cmd[0]} {cmd[-1]}")


def build_lib(verbose: bool = False) -> str:
    os.makedirs(OBJ, exist_ok=True)
    nvcc = _nvcc()
    hdrs = _headers()
    srcs = sorted(glob.glob(os.path.join(CSRC, "*.cu")))
    jobs, objs = [], []
    for s in srcs:
        o = os.path.join(OBJ, os.path.basename(s)[:-3] + ".o")
        objs.append(o)
        if _newer(o, [s] + hdrs):
            jobs.append([nvcc, *NVCC_FLAGS, "-c", s, "-o", o])
    if jobs:
        if verbose:
            print(f"[build] nvcc: {len(jobs)} translation unit(s)", flush=True)
        with cf.ThreadPoolExecutor(max_workers=min(len(jobs), os.cpu_count() or 4)) as ex:
            list(ex.map(_run, jobs))
    if jobs or _newer(LIB, objs):
        _run([nvcc, "-shared", "-o", LIB, *objs, "-gencode", "arch=compute_100a,code=sm_100a",
              "-Xcompiler", "-fPIC"])
    return LIB


def build_ext(verbose: bool = False) -> str:
    import torch
    from torch.utils import cpp_extension as ce
    src = os.path.join(CSRC, "torch_bindings.cpp")
    if not _newer(EXT, [src, LIB] + _headers()):
        return EXT
    if verbose:
        print("[build] g++: torch_bindings.cpp", flush=True)
    torch_lib = os.path.join(os.path.dirname(torch.__file__), "lib")
    cxx = os.environ.get("CXX") or ("/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++")
    inc = [f"-I{p}" for p in ce.include_paths(device_type="cuda")]
    inc += [f"-I{sysconfig.get_paths()['include']}", f"-I{INC}", f"-I{CSRC}",
            "-I/usr/local/cuda/include"]
    cmd = [cxx, "-O2", "-std=c++17", "-fPIC", "-shared", src, "-o", EXT,
           f"-D_GLIBCXX_USE_CXX11_ABI={int(torch._C._GLIBCXX_USE_CXX11_ABI)}",
           "-DTORCH_API_INCLUDE_EXTENSION_H", "-Wno-deprecated-declarations", *inc,
           f"-L{HERE}", "-lb200vllm", f"-L{torch_lib}", "-ltorch", "-ltorch_cpu", "-lc10",
           "-lc10_cuda", "-ltorch_cuda", "-Wl,-rpath,$ORIGIN", f"-Wl,-rpath,{torch_lib}"]
    _run(cmd)
    return EXT


def build(verbose: bool = False) -> tuple[str, str]:
    return build_lib(verbose), build_ext(verbose)


if __name__ == "__main__":
    print(*build(verbose=True), sep="\n")
