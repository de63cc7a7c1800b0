"""Same-box bar: this repo's kernels beside the reference's OWN CUDA kernels recompiled for sm_100
(oracle/_ref/_C_refcuda.so, built by oracle/build_ref_cuda.py from /root/reference/csrc where it
lies).  C3 layers (gptq_marlin_gemm, Llama-3-8B W4A16 g128) at M in {1, 16, 64, 2048}, C5 layers
(gptq_marlin_24_gemm, Llama-2-7B 2:4) at M in {1, 16, 64}, and the C2 paged_attention_v2 microbench.
Timing: CUDA events around back-to-back eager launches rotating over enough weight copies to
exceed L2 (both arms identically; eager, because the reference GEMM allocates inside the op).

Bench-only use of oracle/ (tier rule 3: the checker beside the thing measured, never on the
product path).

  python benchmarks/vs_ref_cuda.py > profiles/r2_vs_ref_cuda.jsonl
"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
L2 = 126 * 2**20
C3 = {"qkv": (4096, 6144), "o": (4096, 4096), "gate_up": (4096, 28672), "down": (14336, 4096)}
C5 = {"qkv": (4096, 12288), "o": (4096, 4096), "gate_up": (4096, 22016), "down": (11008, 4096)}


def timeit(fn, n_var, iters=40):
    for i in range(4):
        fn(i % n_var)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(iters):
        fn(i % n_var)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3  # us


def main():
    import neural_magic_vllm_b200 as pkg
    from oracle import ref_cuda
    ops = pkg.ops
    assert ref_cuda.available(), ref_cuda.why_unavailable()
    ref = ref_cuda.ops()
    g = torch.Generator(device="cuda").manual_seed(0)
    e = torch.empty(0, dtype=torch.int32, device="cuda")
    for name, (k, n) in C3.items():
        wb = k * n // 2
        copies = max(2, -(-2 * L2 // wb))
        qs = [torch.randint(-2**31, 2**31 - 1, (k // 16, n * 2), dtype=torch.int32, device="cuda", generator=g)
              for _ in range(copies)]
        ss = [(torch.rand(k // 128, n, device="cuda", generator=g) * 0.02 + 0.001).half() for _ in range(copies)]
        ws = torch.zeros(n // 64 * 16, dtype=torch.int32, device="cuda")
        for m in (1, 16, 64, 2048):
            a = torch.randn(m, k, dtype=torch.half, device="cuda")
            ours = timeit(lambda i: ops.gptq_marlin_gemm(a, qs[i], ss[i], e, e, ws, 4, m, n, k, True), copies)
            theirs = timeit(lambda i: ref.gptq_marlin_gemm(a, qs[i], ss[i], e, e, ws, 4, m, n, k, True), copies)
            print(json.dumps({"op": "gptq_marlin_gemm", "config": "C3", "layer": name, "m": m, "k": k, "n": n,
                              "ours_us": ours, "ref_cuda_us": theirs, "speedup": theirs / ours}), flush=True)
        del qs, ss
    for name, (k, n) in C5.items():
        wb = k * n // 4 + k * n // 8
        copies = max(2, -(-2 * L2 // wb))
        qs = [torch.randint(-2**31, 2**31 - 1, (k // 32, n * 2), dtype=torch.int32, device="cuda", generator=g)
              for _ in range(copies)]
        mt = [torch.randint(-2**15, 2**15 - 1, (k // 32, n * 2), dtype=torch.int16, device="cuda", generator=g)
              for _ in range(copies)]
        ss = [(torch.rand(k // 128, n, device="cuda", generator=g) * 0.02 + 0.001).half() for _ in range(copies)]
        ws = torch.zeros(n // 128 * 64, dtype=torch.int32, device="cuda")
        for m in (1, 16, 64):
            a = torch.randn(m, k, dtype=torch.half, device="cuda")
            ours = timeit(lambda i: ops.gptq_marlin_24_gemm(a, qs[i], mt[i], ss[i], ws, 4, m, n, k), copies)
            theirs = timeit(lambda i: ref.gptq_marlin_24_gemm(a, qs[i], mt[i], ss[i], ws, 4, m, n, k), copies)
            print(json.dumps({"op": "gptq_marlin_24_gemm", "config": "C5", "layer": name, "m": m, "k": k, "n": n,
                              "ours_us": ours, "ref_cuda_us": theirs, "speedup": theirs / ours}), flush=True)
        del qs, mt, ss
    # C2: paged_attention_v2, 64 seqs x ctx 4096, 32 q / 8 kv heads, D 128, block 16, fp16 KV
    seqs, heads, kvh, d, blk, ctx = 64, 32, 8, 128, 16, 4096
    s = d ** -0.5
    nb = seqs * ctx // blk
    q = torch.empty(seqs, heads, d, dtype=torch.half, device="cuda").uniform_(-s, s, generator=g)
    kc = torch.empty(nb, kvh, d // 8, blk, 8, dtype=torch.half, device="cuda").uniform_(-s, s, generator=g)
    vc = torch.empty(nb, kvh, d, blk, dtype=torch.half, device="cuda").uniform_(-s, s, generator=g)
    bt = torch.randperm(nb, device="cuda", generator=g).to(torch.int32).view(seqs, ctx // blk).contiguous()
    sl = torch.full((seqs,), ctx, dtype=torch.int32, device="cuda")
    parts = ctx // 512
    out = torch.empty_like(q)
    es = torch.empty(seqs, heads, parts, dtype=torch.float32, device="cuda")
    ml = torch.empty_like(es)
    tmp = torch.empty(seqs, heads, parts, d, dtype=torch.half, device="cuda")
    ours = timeit(lambda i: ops.paged_attention_v2(out, es, ml, tmp, q, kc, vc, kvh, s, bt, sl, blk, ctx, None,
                                                   "auto", 1.0), 1)
    out2 = torch.empty_like(q)
    theirs = timeit(lambda i: ref_cuda.paged_attention_v2(out2, es, ml, tmp, q, kc, vc, kvh, s, bt, sl, blk, ctx), 1)
    nbytes = 2 * seqs * ctx * kvh * d * 2
    print(json.dumps({"op": "paged_attention_v2", "config": "C2 fp16 KV", "ours_us": ours, "ref_cuda_us": theirs,
                      "speedup": theirs / ours, "ours_GBps": nbytes / ours / 1e3, "ref_GBps": nbytes / theirs / 1e3,
                      "max_abs_diff": float((out.float() - out2.float()).abs().max())}), flush=True)


if __name__ == "__main__":
    main()
