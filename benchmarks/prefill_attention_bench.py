"""Prefill attention microbench (BASELINE configs[2]: 2k-token prefill at the Llama-3-8B shape, plus
batched / longer variants): varlen_attention_fwd TFLOP/s against the measured bf16 tensor peak, with
torch's fused SDPA and (when importable) the flash_attn library on the same box beside it.

FLOPs counted: 4 * D * (visible query-key pairs) per head (QK^T + PV, causal pairs only).

  python benchmarks/prefill_attention_bench.py [--iters 30]
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
H, KVH, D = 32, 8, 128


def peak_tflops():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["bf16_tflops"])
    except Exception:
        return 1590.0


def timeit(fn, iters):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e-3


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=30)
    ap.add_argument("--shape", default=None, help="BATCHxSEQLEN: only this shape, only our kernel (for ncu)")
    args = ap.parse_args()
    import neural_magic_vllm_b200 as pkg
    ops = pkg.ops
    peak = peak_tflops()
    try:
        from flash_attn import flash_attn_varlen_func as fa_lib
    except Exception:
        fa_lib = None
    shapes = ((1, 2048), (8, 2048), (1, 8192), (4, 512))
    if args.shape:
        shapes = (tuple(int(x) for x in args.shape.split("x")),)
        fa_lib = None
    for dtype in (torch.half, torch.bfloat16) if not args.shape else (torch.bfloat16,):
        for batch, seqlen in shapes:
            t = batch * seqlen
            q = torch.randn(t, H, D, dtype=dtype, device="cuda")
            k = torch.randn(t, KVH, D, dtype=dtype, device="cuda")
            v = torch.randn(t, KVH, D, dtype=dtype, device="cuda")
            cu = torch.arange(0, t + 1, seqlen, dtype=torch.int32, device="cuda")
            out = torch.empty_like(q)
            flops = 4.0 * D * H * batch * (seqlen * (seqlen + 1) / 2)
            ours = timeit(lambda: ops.flash_attn_varlen_func(q, k, v, cu, cu, seqlen, seqlen, D ** -0.5, True, out), args.iters)
            r = {"op": "varlen_attention_fwd", "dtype": str(dtype).split(".")[-1], "batch": batch, "seqlen": seqlen,
                 "heads": H, "kv_heads": KVH, "head_size": D, "causal": True, "us": ours * 1e6,
                 "TFLOPs": flops / ours / 1e12, "tensor_frac": flops / ours / 1e12 / peak}
            if args.shape:
                print(json.dumps(r), flush=True)
                continue
            q4 = q.view(batch, seqlen, H, D).transpose(1, 2)
            k4 = k.view(batch, seqlen, KVH, D).transpose(1, 2)
            v4 = v.view(batch, seqlen, KVH, D).transpose(1, 2)
            try:
                sd = timeit(lambda: torch.nn.functional.scaled_dot_product_attention(q4, k4, v4, is_causal=True,
                                                                                    enable_gqa=True), args.iters)
                r["torch_sdpa_us"] = sd * 1e6
            except Exception as e:
                r["torch_sdpa_us"] = None
                r["torch_sdpa_err"] = repr(e)[:80]
            if fa_lib is not None:
                try:
                    fl = timeit(lambda: fa_lib(q, k, v, cu, cu, seqlen, seqlen, softmax_scale=D ** -0.5, causal=True),
                                args.iters)
                    r["flash_attn_lib_us"] = fl * 1e6
                except Exception as e:
                    r["flash_attn_lib_us"] = None
                    r["flash_attn_lib_err"] = repr(e)[:80]
            print(json.dumps(r), flush=True)


if __name__ == "__main__":
    main()
