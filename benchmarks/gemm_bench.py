"""Quantized-linear microbench: gptq_marlin_gemm W4A16 at the Llama-3-8B
layer shapes and cutlass_scaled_mm fp8 at Llama-3-70B TP shards, swept over M.

Every line reports the time per call (CUDA events on the current stream, weights rotated through
enough copies to exceed the 126 MB L2), the algorithmic HBM GB/s (weights + scales + A + C) against
the measured copy bandwidth, and the dense TFLOP/s against the measured bf16 peak (x2 for fp8).
The roofline that applies is whichever bound is larger for that M.

  python benchmarks/gemm_bench.py [--op marlin|scaled_mm|all] [--iters 50] [--ms 1,16,64,256,2048]
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
LLAMA3_8B = {"qkv": (4096, 6144), "o": (4096, 4096), "gate_up": (4096, 28672), "down": (14336, 4096)}
LLAMA3_70B_TP8 = {"qkv": (8192, 1280), "o": (1024, 8192), "gate_up": (8192, 7168), "down": (3584, 8192)}
LLAMA3_70B_TP2 = {"qkv": (8192, 5120), "o": (4096, 8192), "gate_up": (8192, 28672), "down": (14336, 8192)}


def peaks():
    hbm, tf = 6545.6, 1628.7  # fallback = this pool's driver-measured numbers
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            d = json.load(f)
        hbm, tf = float(d["hbm_gbs"]), float(d["bf16_tflops"])
    except Exception:
        pass
    return hbm, tf


USE_GRAPH = True
LAYERS = None


def time_calls(fn, n_variants: int, iters: int, warmup: int = 5) -> float:
    """Decode runs from CUDA graphs in the reference (worker/model_runner.py:910-1060), so the
    default is one graph holding `n_variants` calls (one per rotated weight copy), replayed."""
    for i in range(warmup):
        fn(i % n_variants)
    torch.cuda.synchronize()
    if USE_GRAPH:
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for i in range(n_variants):
                fn(i)
        g.replay()
        torch.cuda.synchronize()
        reps = max(1, iters // n_variants)
        st, en = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        st.record()
        for _ in range(reps):
            g.replay()
        en.record()
        torch.cuda.synchronize()
        return st.elapsed_time(en) / (reps * n_variants) * 1e-3
    st, en = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    st.record()
    for i in range(iters):
        fn(i % n_variants)
    en.record()
    torch.cuda.synchronize()
    return st.elapsed_time(en) / iters * 1e-3


def bench_marlin(ops, ms, iters, dtype=torch.float16):
    hbm, tf = peaks()
    out = []
    for name, (k, n) in LLAMA3_8B.items():
        if LAYERS and name not in LAYERS:
            continue
        w_bytes = k * n // 2 + (k // 128) * n * 2
        copies = max(2, -(-2 * L2_BYTES // w_bytes))
        g = torch.Generator(device="cuda").manual_seed(0)
        qs = [torch.randint(-2**31, 2**31 - 1, (k // 16, n * 2), dtype=torch.int32, device="cuda", generator=g)
              for _ in range(copies)]
        ss = [(torch.rand(k // 128, n, device="cuda", generator=g) * 0.02 + 0.001).to(dtype) for _ in range(copies)]
        e = torch.empty(0, dtype=torch.int32, device="cuda")
        ws = torch.zeros(n // 64 * 16, dtype=torch.int32, device="cuda")
        for m in ms:
            a = torch.randn(m, k, dtype=dtype, device="cuda")
            t = time_calls(lambda i: ops.gptq_marlin_gemm(a, qs[i], ss[i], e, e, ws, 4, m, n, k, True), copies, iters)
            byts = w_bytes + m * (k + n) * 2
            fl = 2.0 * m * k * n
            r = {"op": "gptq_marlin_gemm", "layer": name, "m": m, "k": k, "n": n, "us": t * 1e6,
                 "GBps": byts / t / 1e9, "hbm_frac": byts / t / 1e9 / hbm, "TFLOPs": fl / t / 1e12,
                 "tensor_frac": fl / t / 1e12 / tf}
            r["roofline_frac"] = max(r["hbm_frac"], r["tensor_frac"])
            out.append(r)
            print(json.dumps(r), flush=True)
        del qs, ss
    return out


LLAMA2_7B = {"qkv": (4096, 12288), "o": (4096, 4096), "gate_up": (4096, 22016), "down": (11008, 4096)}


def bench_marlin_24(ops, ms, iters):
    """C5: Llama-2-7B 2:4 sparse + W4A16 g128 (random compressed stream: timing only)."""
    hbm, tf = peaks()
    for name, (k, n) in LLAMA2_7B.items():
        if LAYERS and name not in LAYERS:
            continue
        w_bytes = k * n // 4 + k * n // 8 + (k // 128) * n * 2
        copies = max(2, -(-2 * L2_BYTES // w_bytes))
        g = torch.Generator(device="cuda").manual_seed(0)
        qs = [torch.randint(-2**31, 2**31 - 1, (k // 32, n * 2), dtype=torch.int32, device="cuda", generator=g)
              for _ in range(copies)]
        mt = [torch.randint(-2**15, 2**15 - 1, (k // 32, n * 2), dtype=torch.int16, device="cuda", generator=g)
              for _ in range(copies)]
        ss = [(torch.rand(k // 128, n, device="cuda", generator=g) * 0.02 + 0.001).half() for _ in range(copies)]
        ws = torch.zeros(n // 128 * 64, dtype=torch.int32, device="cuda")
        for m in ms:
            a = torch.randn(m, k, dtype=torch.half, device="cuda")
            t = time_calls(lambda i: ops.gptq_marlin_24_gemm(a, qs[i], mt[i], ss[i], ws, 4, m, n, k), copies, iters)
            byts = w_bytes + m * (k + n) * 2
            fl = 2.0 * m * k * n  # dense-equivalent
            r = {"op": "gptq_marlin_24_gemm", "layer": name, "m": m, "k": k, "n": n, "us": t * 1e6,
                 "GBps": byts / t / 1e9, "hbm_frac": byts / t / 1e9 / hbm, "TFLOPs": fl / t / 1e12,
                 "tensor_frac": fl / t / 1e12 / tf}
            r["roofline_frac"] = max(r["hbm_frac"], r["tensor_frac"])
            print(json.dumps(r), flush=True)
        del qs, mt, ss


def bench_awq(ops, ms, iters, with_ref):
    """AWQ W4A16 g128 at the Llama-3-8B layer shapes (random words / zeros: timing only); with
    --awq-ref the reference's own kernel (oracle/_ref/_C_refcuda.so) is timed beside it."""
    hbm, tf = peaks()
    ref = None
    if with_ref:
        from oracle import ref_cuda
        ref = ref_cuda.ops() if ref_cuda.available() else None
    for name, (k, n) in LLAMA3_8B.items():
        if LAYERS and name not in LAYERS:
            continue
        w_bytes = k * n // 2 + (k // 128) * n * 2 + (k // 128) * n // 2
        copies = max(2, -(-2 * L2_BYTES // w_bytes))
        g = torch.Generator(device="cuda").manual_seed(0)
        qs = [torch.randint(-2**31, 2**31 - 1, (k, n // 8), dtype=torch.int32, device="cuda", generator=g)
              for _ in range(copies)]
        zs = [torch.randint(-2**31, 2**31 - 1, (k // 128, n // 8), dtype=torch.int32, device="cuda", generator=g)
              for _ in range(copies)]
        ss = [(torch.rand(k // 128, n, device="cuda", generator=g) * 0.02 + 0.001).half() for _ in range(copies)]
        for m in ms:
            a = torch.randn(m, k, dtype=torch.half, device="cuda")
            t = time_calls(lambda i: ops.awq_gemm(a, qs[i], ss[i], zs[i], 8), copies, iters)
            byts = w_bytes + m * (k + n) * 2
            fl = 2.0 * m * k * n
            r = {"op": "awq_gemm", "layer": name, "m": m, "k": k, "n": n, "us": t * 1e6,
                 "GBps": byts / t / 1e9, "hbm_frac": byts / t / 1e9 / hbm, "TFLOPs": fl / t / 1e12,
                 "tensor_frac": fl / t / 1e12 / tf}
            r["roofline_frac"] = max(r["hbm_frac"], r["tensor_frac"])
            if ref is not None:
                r["reference_cuda_us"] = time_calls(lambda i: ref.awq_gemm(a, qs[i], ss[i], zs[i], 8), copies, iters) * 1e6
            print(json.dumps(r), flush=True)
        del qs, zs, ss


def bench_scaled_mm(ops, ms, iters, shapes, tag, fp8=True):
    hbm, tf = peaks()
    out = []
    for name, (k, n) in shapes.items():
        if LAYERS and name not in LAYERS:
            continue
        copies = max(2, -(-2 * L2_BYTES // (k * n)))
        if fp8:
            bs = [(torch.randn(n, k, device="cuda") * 0.5).to(torch.float8_e4m3fn).t() for _ in range(copies)]
        else:
            bs = [torch.randint(-127, 127, (n, k), dtype=torch.int8, device="cuda").t() for _ in range(copies)]
        sb = torch.rand(1, n, device="cuda") * 0.01
        for m in ms:
            if fp8:
                a = (torch.randn(m, k, device="cuda") * 0.5).to(torch.float8_e4m3fn)
            else:
                a = torch.randint(-127, 127, (m, k), dtype=torch.int8, device="cuda")
            sa = torch.rand(m, 1, device="cuda") * 0.01
            t = time_calls(lambda i: ops.cutlass_scaled_mm(a, bs[i], sa, sb, torch.bfloat16), copies, iters)
            byts = k * n + m * k + m * n * 2 + (m + n) * 4
            fl = 2.0 * m * k * n
            r = {"op": "cutlass_scaled_mm", "kind": "fp8" if fp8 else "int8", "shard": tag, "layer": name,
                 "m": m, "k": k, "n": n, "us": t * 1e6, "GBps": byts / t / 1e9,
                 "hbm_frac": byts / t / 1e9 / hbm, "TFLOPs": fl / t / 1e12,
                 "tensor_frac": fl / t / 1e12 / (2 * tf)}
            r["roofline_frac"] = max(r["hbm_frac"], r["tensor_frac"])
            out.append(r)
            print(json.dumps(r), flush=True)
        del bs
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--op", default="all", choices=["marlin", "marlin24", "scaled_mm", "awq", "all"])
    ap.add_argument("--awq-ref", action="store_true", help="time the reference's AWQ kernel beside ours")
    ap.add_argument("--iters", type=int, default=50)
    ap.add_argument("--ms", default="1,16,64,256,2048")
    ap.add_argument("--no-graph", action="store_true", help="time eager launches (host overhead included)")
    ap.add_argument("--layers", default="", help="comma list of layer names to keep")
    args = ap.parse_args()
    global USE_GRAPH, LAYERS
    USE_GRAPH = not args.no_graph
    LAYERS = set(args.layers.split(",")) if args.layers else None
    ms = [int(x) for x in args.ms.split(",")]
    import neural_magic_vllm_b200 as pkg
    ops = pkg.ops
    torch.cuda.set_device(0)
    if args.op in ("marlin", "all"):
        bench_marlin(ops, ms, args.iters)
    if args.op in ("marlin24", "all"):
        bench_marlin_24(ops, [m for m in ms if m <= 128], args.iters)
    if args.op in ("awq", "all"):
        bench_awq(ops, [m for m in ms if m <= 256], args.iters, args.awq_ref)
    if args.op in ("scaled_mm", "all"):
        bench_scaled_mm(ops, ms, args.iters, LLAMA3_70B_TP8, "70B/TP8")
        bench_scaled_mm(ops, ms, args.iters, LLAMA3_70B_TP2, "70B/TP2")


if __name__ == "__main__":
    main()
