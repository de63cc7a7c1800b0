"""paged_attention_v1 / v2 at one decode-step shape (default: the bench step's batch 64 x ctx 1024,
32 q / 8 kv heads, D 128, block 16), standalone, CUDA-graph replay of 8 back-to-back calls.
KV bytes per call exceed L2, so every call streams from HBM.

  python benchmarks/attn_shape_bench.py [--batch 64] [--ctx 1024]
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--ctx", type=int, default=1024)
    ap.add_argument("--ragged", action="store_true", help="sequence lengths uniform in [ctx/2, ctx]")
    args = ap.parse_args()
    import neural_magic_vllm_b200 as pkg
    ops = pkg.ops
    h, kvh, d, bs = 32, 8, 128, 16
    b, ctx = args.batch, args.ctx
    nblk = b * ((ctx + bs - 1) // bs)
    g = torch.Generator(device="cuda").manual_seed(0)
    kc = torch.randn(nblk, kvh, d // 8, bs, 8, dtype=torch.half, device="cuda", generator=g)
    vc = torch.randn(nblk, kvh, d, bs, dtype=torch.half, device="cuda", generator=g)
    q = torch.randn(b, h, d, dtype=torch.half, device="cuda", generator=g)
    bt = torch.randperm(nblk, device="cuda", generator=g).to(torch.int32).view(b, -1)
    if args.ragged:
        sl = torch.randint(ctx // 2, ctx + 1, (b,), device="cuda", generator=g).to(torch.int32)
    else:
        sl = torch.full((b,), ctx, dtype=torch.int32, device="cuda")
    out = torch.empty_like(q)
    parts = (ctx + 511) // 512
    tmp = torch.empty(b, h, parts, d, dtype=torch.half, device="cuda")
    es = torch.empty(b, h, parts, device="cuda")
    ml = torch.empty_like(es)
    kv_bytes = float((sl.sum().item()) * kvh * d * 2 * 2)
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs", 6551.7) if os.path.exists(
        os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6551.7

    def v1():
        ops.paged_attention_v1(out, q, kc, vc, kvh, d ** -0.5, bt, sl, bs, ctx, None, "auto", 1.0)

    def v2():
        ops.paged_attention_v2(out, es, ml, tmp, q, kc, vc, kvh, d ** -0.5, bt, sl, bs, ctx, None, "auto", 1.0)

    for name, fn in (("paged_attention_v1", v1), ("paged_attention_v2", v2)):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr):
            for _ in range(8):
                fn()
        gr.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            gr.replay()
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) / 80 * 1e3
        print(json.dumps({"op": name, "batch": b, "ctx": ctx, "ragged": args.ragged, "us": us,
                          "GBps": kv_bytes / us / 1e3, "hbm_frac": kv_bytes / us / 1e3 / peak}), flush=True)


if __name__ == "__main__":
    main()
