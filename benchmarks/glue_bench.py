"""Microbench of the decode-side glue kernels: rms_norm,
fused_add_rms_norm, rotary_embedding, rotary_embedding_and_cache, silu_and_mul at Llama-3-8B shapes.

All of them are HBM-bound element-wise work.  Every line reports the time per call (CUDA events on
the current stream around a CUDA graph holding one call per rotated buffer set; the sets together
exceed the 126 MB L2 at the prefill sizes) and the algorithmic HBM GB/s against the measured copy
bandwidth.  Small-token lines are latency, not bandwidth: they are what a decode step pays.

  python benchmarks/glue_bench.py [--tokens 1,64,4096,16384] [--iters 200]
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
L2_BYTES = 126 * 2**20
HIDDEN, HEADS, KV_HEADS, HEAD, INTER, BLOCK = 4096, 32, 8, 128, 14336, 16


def hbm_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"])
    except Exception:
        return 6545.6


def time_graph(fns, iters):
    for f in fns:
        f()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for f in fns:
            f()
    g.replay()
    torch.cuda.synchronize()
    reps = max(1, iters // len(fns))
    st, en = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    st.record()
    for _ in range(reps):
        g.replay()
    en.record()
    torch.cuda.synchronize()
    return st.elapsed_time(en) * 1e3 / (reps * len(fns))  # us per call


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tokens", default="1,64,4096,16384")
    ap.add_argument("--iters", type=int, default=200)
    args = ap.parse_args()
    import neural_magic_vllm_b200 as pkg
    ops = pkg.ops
    peak = hbm_peak()
    dev, dt = "cuda", torch.half
    for t in [int(x) for x in args.tokens.split(",")]:
        cases = []
        # (name, bytes per call, builder -> closure)
        def rms():
            x = torch.randn(t, HIDDEN, device=dev, dtype=dt); o = torch.empty_like(x)
            w = torch.ones(HIDDEN, device=dev, dtype=dt)
            return lambda: ops.rms_norm(o, x, w, 1e-5)
        cases.append(("rms_norm", 2 * t * HIDDEN * 2 + HIDDEN * 2, rms))
        def add_rms():
            x = torch.randn(t, HIDDEN, device=dev, dtype=dt); r = torch.randn_like(x) * 0.01
            w = torch.ones(HIDDEN, device=dev, dtype=dt)
            return lambda: ops.fused_add_rms_norm(x, r, w, 1e-5)
        cases.append(("fused_add_rms_norm", 4 * t * HIDDEN * 2 + HIDDEN * 2, add_rms))
        qkv_w = (HEADS + 2 * KV_HEADS) * HEAD
        inv = 500000.0 ** (-torch.arange(0, HEAD, 2, device=dev).float() / HEAD)
        tab = torch.arange(8192, device=dev).float()[:, None] * inv[None]
        cache = torch.cat((tab.cos(), tab.sin()), -1).to(dt)
        pos = torch.randint(0, 8192, (t,), device=dev)
        def rot():
            qkv = torch.randn(t, qkv_w, device=dev, dtype=dt)
            q, k = qkv[:, : HEADS * HEAD], qkv[:, HEADS * HEAD: (HEADS + KV_HEADS) * HEAD]
            return lambda: ops.rotary_embedding(pos, q, k, HEAD, cache, True)
        cases.append(("rotary_embedding", 2 * t * (HEADS + KV_HEADS) * HEAD * 2 + t * HEAD * 2, rot))
        nb = (t + BLOCK - 1) // BLOCK + 1
        slots = torch.randperm(nb * BLOCK, device=dev)[:t]
        def rot_cache():
            qkv = torch.randn(t, qkv_w, device=dev, dtype=dt)
            a, b = HEADS * HEAD, (HEADS + KV_HEADS) * HEAD
            kc = torch.zeros(nb, KV_HEADS, HEAD // 8, BLOCK, 8, device=dev, dtype=dt)
            vc = torch.zeros(nb, KV_HEADS, HEAD, BLOCK, device=dev, dtype=dt)
            return lambda: ops.rotary_embedding_and_cache(pos, qkv[:, :a], qkv[:, a:b], qkv[:, b:], HEAD, cache,
                                                          True, kc, vc, slots, "auto", 1.0)
        cases.append(("rotary_embedding_and_cache",
                      2 * t * (HEADS + KV_HEADS) * HEAD * 2 + t * HEAD * 2 + 3 * t * KV_HEADS * HEAD * 2, rot_cache))
        def silu():
            gu = torch.randn(t, 2 * INTER, device=dev, dtype=dt); o = torch.empty(t, INTER, device=dev, dtype=dt)
            return lambda: ops.silu_and_mul(o, gu)
        cases.append(("silu_and_mul", 3 * t * INTER * 2, silu))
        for name, byts, build in cases:
            n_sets = max(2, min(16, -(-2 * L2_BYTES // max(byts, 1))))
            if byts * n_sets > 24 * 2**30:
                n_sets = max(2, int(24 * 2**30 // byts))
            fns = [build() for _ in range(n_sets)]
            us = time_graph(fns, args.iters)
            print(json.dumps({"op": name, "tokens": t, "us_per_call": round(us, 2), "bytes": byts,
                              "GBps": round(byts / us / 1e3, 1), "hbm_frac": round(byts / us / 1e3 / peak, 3),
                              "buffer_sets": n_sets, "l2_exceeded": byts * n_sets > L2_BYTES, "dtype": "f16"}),
                  flush=True)
            del fns
            torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
