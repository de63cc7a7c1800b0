"""Build recipe for oracle/_ref: the reference's OWN CPU backend, compiled in place.

TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product path; only
tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
may load what this script produces.

What it does: compiles the six translation units of the reference's CPU extension
(/root/reference/csrc/cpu/{attention,cache,activation,layernorm,pos_encoding,
torch_bindings}.cpp) exactly where they lie (no source is copied into this repo) into
oracle/_ref/_C_ref_cpu.so.  The ops are registered under the torch namespaces
`_C_ref` / `_C_ref_cache_ops` (-DTORCH_EXTENSION_NAME=_C_ref) so that they can live in
the same process as the product's `_C` / `_C_cache_ops` registration.

Coverage of the compiled reference (csrc/cpu/attention.cpp:412-432, cache.cpp:110,
cpu_types_x86.hpp:16-18): paged_attention_v1/v2, reshape_and_cache, copy_blocks for
fp32 and bf16, block_size 16, kv_scale 1.0, kv_cache_dtype "auto".

The GPU box has no /root/reference: there the prebuilt .so (git-ignored, but shipped
by gpurun) is used as is, and this script is a no-op.
"""
import os
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("B200_REFERENCE_ROOT", "/root/reference")
OUT_DIR = os.path.join(HERE, "_ref")
OUT = os.path.join(OUT_DIR, "_C_ref_cpu.so")

SOURCES = ["attention.cpp", "cache.cpp", "activation.cpp", "layernorm.cpp",
           "pos_encoding.cpp", "torch_bindings.cpp"]


def _isa_flags():
    flags = []
    try:
        cpuinfo = open("/proc/cpuinfo").read()
    except OSError:
        cpuinfo = ""
    # same probe as cmake/cpu_extension.cmake: avx512f => the AVX-512 path.
    # The .so travels to the GPU box, whose host may differ; bf16 conversion
    # intrinsics are only enabled if this host has them AND the env does not veto.
    if "avx512f" in cpuinfo:
        flags += ["-mavx512f", "-mavx512vl", "-mavx512bw", "-mavx512dq"]
        # off by default: the prebuilt .so must also run on the GPU box's host CPU
        if "avx512_bf16" in cpuinfo and os.environ.get("B200_REF_BF16_ISA") == "1":
            flags += ["-mavx512bf16"]
    elif "avx2" in cpuinfo:
        flags += ["-mavx2"]
    return flags


def build(force: bool = False) -> str | None:
    """Returns the path of the built library, or None when it cannot be built."""
    srcs = [os.path.join(REF, "csrc", "cpu", s) for s in SOURCES]
    if not all(os.path.exists(s) for s in srcs):
        return OUT if os.path.exists(OUT) else None
    if os.path.exists(OUT) and not force:
        newest = max(os.path.getmtime(s) for s in srcs + [__file__])
        if os.path.getmtime(OUT) >= newest:
            return OUT
    import torch
    from torch.utils import cpp_extension as ce
    os.makedirs(OUT_DIR, exist_ok=True)
    inc = [f"-I{p}" for p in ce.include_paths()]
    inc += [f"-I{sysconfig.get_paths()['include']}", f"-I{os.path.join(REF, 'csrc')}"]
    torch_lib = os.path.join(os.path.dirname(torch.__file__), "lib")
    cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    cflags = ["-O2", "-std=c++17", "-fPIC", "-fopenmp", "-DVLLM_CPU_EXTENSION",
              "-DTORCH_EXTENSION_NAME=_C_ref", "-DTORCH_API_INCLUDE_EXTENSION_H",
              f"-D_GLIBCXX_USE_CXX11_ABI={int(torch._C._GLIBCXX_USE_CXX11_ABI)}",
              "-Wno-unused-variable", "-Wno-sign-compare"] + _isa_flags()
    objs = []
    procs = []
    for s in srcs:
        o = os.path.join(OUT_DIR, os.path.basename(s).replace(".cpp", ".o"))
        objs.append(o)
        procs.append(subprocess.Popen([cxx, "-c", s, "-o", o] + cflags + inc))
    for p in procs:
        if p.wait() != 0:
            raise RuntimeError("reference CPU extension failed to compile")
    link = [cxx, "-shared", "-o", OUT] + objs + [
        f"-L{torch_lib}", "-ltorch", "-ltorch_cpu", "-lc10", "-ltorch_python",
        f"-Wl,-rpath,{torch_lib}"]
    # /opt/gcc ships no libgomp.spec: link the distro libgomp explicitly
    gomp_dirs = [d for d in ("/usr/lib/gcc/x86_64-linux-gnu/13", "/usr/lib/x86_64-linux-gnu")
                 if os.path.isdir(d)]
    link += [f"-L{d}" for d in gomp_dirs] + ["-lgomp"]
    subprocess.check_call(link)
    for o in objs:
        os.remove(o)
    # sidecar: the ISA the library needs, checked by oracle/ref_cpu.py before loading
    with open(os.path.join(OUT_DIR, "isa.txt"), "w") as f:
        f.write(" ".join(fl[2:] for fl in _isa_flags()) + "\n")
    return OUT


if __name__ == "__main__":
    path = build(force="--force" in sys.argv)
    print(path or "reference sources not present and no prebuilt oracle/_ref")
