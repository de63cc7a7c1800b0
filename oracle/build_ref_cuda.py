"""Build recipe for oracle/_ref/_C_refcuda.so: the reference's OWN CUDA kernels, recompiled for sm_100.

TEST / BASELINE INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product path; only
tests/, __graft_entry__.smoke(), bench.py's baseline legs and benchmarks/*_bench.py `ref_us`
columns may load what this script produces.

What it does: compiles the reference's CUDA translation units for the hot path exactly where
they lie under /root/reference/csrc (no source is copied into this repo) with
`-gencode arch=compute_100,code=sm_100` -- i.e. the "recompiled mma.sync / cp.async kernels" that
are the same-box GPU baseline -- and links them with oracle/ref_cuda_bindings.cpp, which registers
them under the torch namespaces `_C_refcuda` / `_C_refcuda_cache_ops` so they coexist with the
product's `_C` / `_C_cache_ops` registration in one process.

Not built: csrc/quantization/cutlass_w8a8/*.cu (needs the CUTLASS v3.5.0 tree that the reference's
CMakeLists.txt:157-165 fetches at configure time; not on disk).

The GPU box has no /root/reference: there the prebuilt .so (git-ignored, shipped by gpurun) is used
as is, and this script is a no-op.
"""
from __future__ import annotations

import concurrent.futures as cf
import os
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("B200_REFERENCE_ROOT", "/root/reference")
OUT_DIR = os.path.join(HERE, "_ref")
OBJ_DIR = os.path.join(OUT_DIR, "obj_cuda")
OUT = os.path.join(OUT_DIR, "_C_refcuda.so")

SOURCES = [
    "attention/attention_kernels.cu",
    "cache_kernels.cu",
    "activation_kernels.cu",
    "layernorm_kernels.cu",
    "pos_encoding_kernels.cu",
    "quantization/gptq_marlin/gptq_marlin.cu",
    "quantization/gptq_marlin/gptq_marlin_repack.cu",
    "quantization/marlin/sparse/marlin_24_cuda_kernel.cu",
    "quantization/compressed_tensors/int8_quant_kernels.cu",
    "quantization/fp8/common.cu",
    "quantization/fp8/fp8_marlin.cu",
    "quantization/awq/gemm_kernels.cu",
]


def _run(cmd):
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(" ".join(cmd) + "\n" + r.stdout[-4000:] + r.stderr[-8000:])
        raise RuntimeError(f"reference CUDA build step failed: {cmd[-3]}")


def build(force: bool = False, verbose: bool = False) -> str | None:
    """Returns the path of the built library, or None when it cannot be built."""
    srcs = [os.path.join(REF, "csrc", s) for s in SOURCES]
    bind = os.path.join(HERE, "ref_cuda_bindings.cpp")
    if not all(os.path.exists(s) for s in srcs):
        return OUT if os.path.exists(OUT) else None
    if os.path.exists(OUT) and not force:
        newest = max(os.path.getmtime(s) for s in srcs + [bind, __file__])
        if os.path.getmtime(OUT) >= newest:
            return OUT
    import torch
    from torch.utils import cpp_extension as ce
    os.makedirs(OBJ_DIR, exist_ok=True)
    inc = [f"-I{p}" for p in ce.include_paths(device_type="cuda")]
    inc += [f"-I{sysconfig.get_paths()['include']}", f"-I{os.path.join(REF, 'csrc')}",
            "-I/usr/local/cuda/include"]
    abi = f"-D_GLIBCXX_USE_CXX11_ABI={int(torch._C._GLIBCXX_USE_CXX11_ABI)}"
    nvcc = "/usr/local/cuda/bin/nvcc"
    # flags of the reference's own CMakeLists.txt (:131-142) minus the arch list: sm_100 only
    nv_flags = ["-gencode", "arch=compute_100,code=sm_100", "-O2", "-std=c++17",
                "--expt-relaxed-constexpr", "--expt-extended-lambda", "-Xcompiler", "-fPIC",
                "-DTORCH_API_INCLUDE_EXTENSION_H", "-DENABLE_FP8", abi,
                "-U__CUDA_NO_HALF_OPERATORS__", "-U__CUDA_NO_HALF_CONVERSIONS__",
                "-U__CUDA_NO_BFLOAT16_CONVERSIONS__", "-U__CUDA_NO_HALF2_OPERATORS__",
                "-w", "--threads", "4"]
    jobs, objs = [], []
    for s in srcs:
        o = os.path.join(OBJ_DIR, os.path.basename(s)[:-3] + ".o")
        objs.append(o)
        if force or not os.path.exists(o) or os.path.getmtime(o) < os.path.getmtime(s):
            jobs.append([nvcc, *nv_flags, *inc, "-c", s, "-o", o])
    if verbose:
        print(f"[build_ref_cuda] nvcc: {len(jobs)} reference translation unit(s) for sm_100", flush=True)
    with cf.ThreadPoolExecutor(max_workers=max(1, min(len(jobs), (os.cpu_count() or 4) // 2))) as ex:
        list(ex.map(_run, jobs))
    cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    bo = os.path.join(OBJ_DIR, "ref_cuda_bindings.o")
    _run([cxx, "-O2", "-std=c++17", "-fPIC", "-c", bind, "-o", bo, abi,
          "-DTORCH_API_INCLUDE_EXTENSION_H", "-Wno-deprecated-declarations", *inc])
    torch_lib = os.path.join(os.path.dirname(torch.__file__), "lib")
    _run([nvcc, "-shared", "-o", OUT, *objs, bo, "-gencode", "arch=compute_100,code=sm_100",
          "-Xcompiler", "-fPIC", f"-L{torch_lib}", "-ltorch", "-ltorch_cpu", "-lc10", "-lc10_cuda",
          "-ltorch_cuda", "-Xlinker", f"-rpath={torch_lib}"])
    return OUT


if __name__ == "__main__":
    path = build(force="--force" in sys.argv, verbose=True)
    print(path or "reference sources not present and no prebuilt oracle/_ref/_C_refcuda.so")
