"""Per-role cycle accounting of the stream-K Marlin decode kernel (debug tool).

Needs a library built with B200_EXTRA_NVCC_FLAGS=-DB200_STREAM_PROF; prints, averaged over the
CTAs, where the MMA issuer / one dequant warp / the producer / the epilogue spend their cycles.

  python benchmarks/stream_prof.py [--m 16] [--layer gate_up]
"""
import argparse
import ctypes
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
SHAPES = {"qkv": (4096, 6144), "o": (4096, 4096), "gate_up": (4096, 28672), "down": (14336, 4096)}
ROLES = {
    "mma": (0, ["wait full_p", "wait full_a", "issue/other", "wait tmem_empty"]),
    "dequant(warp2)": (8, ["wait full_p", "lds+release", "wait empty_a", "unpack+st", "st_wait+arrive", "loop/skip"]),
    "producer": (16, ["wait empty_p", "issue/other"]),
    "epilogue": (24, ["wait tmem_full", "ld+store+publish", "spin+barrier", "reduce"]),
}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--m", type=int, default=16)
    ap.add_argument("--layer", default="gate_up")
    args = ap.parse_args()
    import neural_magic_vllm_b200 as pkg
    ops = pkg.ops
    lib = ctypes.CDLL(os.path.join(ROOT, "neural_magic_vllm_b200", "libb200vllm.so"))
    lib.b200_debug_stream_prof.argtypes = [ctypes.c_void_p, ctypes.c_int32]
    k, n = SHAPES[args.layer]
    g = torch.Generator(device="cuda").manual_seed(0)
    copies = 6
    qs = [torch.randint(-2**31, 2**31 - 1, (k // 16, n * 2), dtype=torch.int32, device="cuda", generator=g)
          for _ in range(copies)]
    ss = [(torch.rand(k // 128, n, device="cuda", generator=g) * 0.02 + 0.001).half() for _ in range(copies)]
    e = torch.empty(0, dtype=torch.int32, device="cuda")
    ws = torch.zeros(n // 64 * 16, dtype=torch.int32, device="cuda")
    a = torch.randn(args.m, k, dtype=torch.half, device="cuda")
    for i in range(copies):
        ops.gptq_marlin_gemm(a, qs[i], ss[i], e, e, ws, 4, args.m, n, k, True)
    torch.cuda.synchronize()
    buf = np.zeros(160 * 32, dtype=np.uint64)
    got = lib.b200_debug_stream_prof(buf.ctypes.data, buf.size)
    if got <= 0:
        print("library was not built with -DB200_STREAM_PROF")
        return
    prof = buf.reshape(160, 32)[:148].astype(np.float64)
    marks(prof)
    for role, (base, names) in ROLES.items():
        tot = prof[:, base:base + 6].sum(axis=1)
        print(f"{role}: total cycles avg {tot.mean():.0f} (min {tot.min():.0f} max {tot.max():.0f})")
        for j, nm in enumerate(names):
            c = prof[:, base + j]
            print(f"    {nm:24s} avg {c.mean():9.0f}  ({100 * c.mean() / max(tot.mean(), 1):5.1f} %)  max {c.max():9.0f}")


def marks(prof):
    for nm, j in (("prologue done", 6), ("MMA warp done", 7), ("epilogue role done", 30), ("kernel end", 31)):
        c = prof[:, j]
        print(f"  since kernel entry: {nm:20s} avg {c.mean():9.0f}  min {c.min():9.0f}  max {c.max():9.0f}")


if __name__ == "__main__":
    main()
