"""Per-role cycle accounting of the prefill attention kernel (debug tool).

Needs a library built with B200_EXTRA_NVCC_FLAGS=-DB200_FA_PROF.  Prints, summed over all CTAs and
normalised per kv step of one query tile, where the MMA issuer and a softmax warp spend their cycles.

  python benchmarks/fa_prof.py [--shape 1x8192]
"""
import argparse
import ctypes
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
H, KVH, D = 32, 8, 128
MMA = ["issue / other", "wait k_full / v_full", "wait p_full[A]", "wait p_full[B]"]
SOFTMAX = ["loop / other", "wait s_full", "tcgen05.ld + max", "rescale", "exp2 + pack + st", "st wait + arrive", "epilogue"]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shape", default="1x8192")
    args = ap.parse_args()
    import neural_magic_vllm_b200 as pkg
    ops = pkg.ops
    lib = ctypes.CDLL(os.path.join(ROOT, "neural_magic_vllm_b200", "libb200vllm.so"))
    lib.b200_debug_fa_prof.argtypes = [ctypes.c_void_p, ctypes.c_int32, ctypes.c_int32]
    batch, seqlen = (int(x) for x in args.shape.split("x"))
    t = batch * seqlen
    q = torch.randn(t, H, D, dtype=torch.bfloat16, device="cuda")
    k = torch.randn(t, KVH, D, dtype=torch.bfloat16, device="cuda")
    v = torch.randn(t, KVH, D, dtype=torch.bfloat16, device="cuda")
    cu = torch.arange(0, t + 1, seqlen, dtype=torch.int32, device="cuda")
    out = torch.empty_like(q)
    buf = np.zeros(32, dtype=np.uint64)
    for _ in range(2):
        ops.flash_attn_varlen_func(q, k, v, cu, cu, seqlen, seqlen, D ** -0.5, True, out)
    torch.cuda.synchronize()
    lib.b200_debug_fa_prof(buf.ctypes.data, 32, 1)
    ops.flash_attn_varlen_func(q, k, v, cu, cu, seqlen, seqlen, D ** -0.5, True, out)
    torch.cuda.synchronize()
    if lib.b200_debug_fa_prof(buf.ctypes.data, 32, 0) <= 0:
        print("library was not built with -DB200_FA_PROF")
        return
    nq = seqlen // 128
    steps_tile = batch * H * nq * (nq + 1) // 2          # kv steps summed over all 128-query tiles
    steps_a = steps_tile / 2                               # steps of the A tiles (about half)
    ctas = batch * H * (nq // 2)
    print(f"shape {args.shape}: {ctas} CTAs, {steps_tile} tile-steps")
    tot = float(buf[0:8].sum())
    print(f"MMA issuer: {tot / ctas:.0f} cycles per CTA, {tot / steps_tile:.0f} per tile-step")
    for i, nm in enumerate(MMA):
        print(f"    {nm:24s} {buf[i] / steps_tile:8.0f} / tile-step  ({100 * buf[i] / tot:5.1f} %)")
    for name, base in (("softmax warp of tile A", 8), ("softmax warp of tile B", 16)):
        tot = float(buf[base:base + 8].sum())
        print(f"{name}: {tot / ctas:.0f} cycles per CTA, {tot / steps_a:.0f} per step")
        for i, nm in enumerate(SOFTMAX):
            print(f"    {nm:24s} {buf[base + i] / steps_a:8.0f} / step  ({100 * buf[base + i] / tot:5.1f} %)")


if __name__ == "__main__":
    main()
