"""CPU arm of the decode benchmark: the reference's CPU path for one Llama-3-8B W4A16 decode step.

TEST / BASELINE INFRASTRUCTURE ONLY (see oracle/__init__.py): imported by bench.py's
`cpu_baseline` leg and by `bench.py --impl reference`, never by the product.

What runs (BASELINE.json north_star "csrc/cpu paged attention plus torch
dequant-matmul"):
  * attention + cache write: the reference's OWN compiled CPU kernels (oracle/_ref/_C_ref_cpu.so =
    /root/reference/csrc/cpu/{attention,cache}.cpp): reshape_and_cache + paged_attention_v1/v2,
    bf16 KV (the CPU backend has no fp16: cpu_types_x86.hpp:16-18), version picked by the
    reference's own heuristic (vllm/attention/ops/paged_attn.py:118-119);
  * quantized linears: the reference CPU extension registers no GEMM op, so W4A16 is
    unpack GPTQ int32 [K/8, N] -> (q - 8) * s[g] -> torch.matmul (gptq semantics:
    model_executor/layers/quantization/utils/quant_utils.py:73), in bf16 or fp32, whichever a
    probe finds faster on this host (AMX hosts: bf16);
  * RMSNorm / rotary / SiLU*mul / residual adds: plain torch ops (the reference CPU backend routes
    them to its own kernels; they are a rounding error of the step next to the matmuls).
A full 32-layer step costs several seconds of host time, so the timed sample is `sample_layers`
of the 32 identical layers plus the lm_head, and the step time is extrapolated linearly:
t_step = t_layers * 32 / sample_layers + t_lm_head.  The sample is stated in the result.
"""
from __future__ import annotations

import os
import time

import torch

HIDDEN, HEADS, KV_HEADS, HEAD, INTER, VOCAB, LAYERS, BLOCK = 4096, 32, 8, 128, 14336, 128256, 32, 16
_SHIFTS = torch.arange(0, 32, 4, dtype=torch.int32)


class _QLinearCPU:
    def __init__(self, k, n, gen, dtype):
        self.k, self.n, self.dtype = k, n, dtype
        self.qweight = torch.randint(-2**31, 2**31 - 1, (k // 8, n), dtype=torch.int32, generator=gen)
        self.scales = (torch.rand(k // 128, n, generator=gen) * 0.004 + 0.0005).to(dtype)

    def __call__(self, x):
        # unpack -> (q - 8) * s -> matmul, every call: the weights stay int4 in memory
        # (byte view: nibble 2j of a little-endian int32 is the low half of byte j)
        b = self.qweight.view(torch.uint8).view(self.k // 8, self.n, 4)
        q = torch.stack((b & 15, b >> 4), dim=-1).permute(0, 2, 3, 1)           # [K/8, 4, 2, N] = k-major
        w = (q.reshape(self.k // 128, 128, self.n).to(self.dtype) - 8) * self.scales[:, None, :]
        return x @ w.view(self.k, self.n)


def _rms(x, w, eps=1e-5):
    v = x.float()
    return (v * torch.rsqrt(v.pow(2).mean(-1, keepdim=True) + eps)).to(x.dtype) * w


def _rope(x, cos, sin):
    x1, x2 = x[..., : HEAD // 2], x[..., HEAD // 2:]
    return torch.cat((x1 * cos - x2 * sin, x2 * cos + x1 * sin), dim=-1)


def _pick_dtype():
    a = torch.randn(64, 4096)
    b = torch.randn(4096, 4096)
    best, best_t = torch.float32, None
    for dt in (torch.float32, torch.bfloat16):
        x, y = a.to(dt), b.to(dt)
        x @ y
        t0 = time.perf_counter()
        for _ in range(3):
            x @ y
        t = time.perf_counter() - t0
        if best_t is None or t < best_t:
            best, best_t = dt, t
    return best


def run(batch: int = 64, ctx: int = 1024, sample_layers: int = 2, steps: int = 2, warmup: int = 1):
    """-> dict(tokens_per_s, ms_per_step, cores, sample, ...) or dict(unavailable=why)."""
    from oracle import ref_cpu
    if not ref_cpu.available():
        return {"unavailable": ref_cpu.why_unavailable()}
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    mm_dtype = _pick_dtype()
    gen = torch.Generator().manual_seed(1234)
    blocks_per_seq = (ctx + BLOCK) // BLOCK
    nb = batch * blocks_per_seq
    s = HEAD ** -0.5
    layers = []
    for _ in range(sample_layers):
        L = {
            "qkv": _QLinearCPU(HIDDEN, (HEADS + 2 * KV_HEADS) * HEAD, gen, mm_dtype),
            "o": _QLinearCPU(HEADS * HEAD, HIDDEN, gen, mm_dtype),
            "gate_up": _QLinearCPU(HIDDEN, 2 * INTER, gen, mm_dtype),
            "down": _QLinearCPU(INTER, HIDDEN, gen, mm_dtype),
            "ln1": torch.ones(HIDDEN, dtype=mm_dtype), "ln2": torch.ones(HIDDEN, dtype=mm_dtype),
            "kc": torch.empty(nb, KV_HEADS, HEAD // 8, BLOCK, 8, dtype=torch.bfloat16).uniform_(-s, s, generator=gen),
            "vc": torch.empty(nb, KV_HEADS, HEAD, BLOCK, dtype=torch.bfloat16).uniform_(-s, s, generator=gen),
        }
        layers.append(L)
    lm_head = (torch.randn(VOCAB, HIDDEN, generator=gen) * 0.02).to(mm_dtype)
    embed = (torch.randn(4096, HIDDEN, generator=gen) * 0.02).to(mm_dtype)
    tokens = torch.randint(0, 4096, (batch,), generator=gen)
    bt = torch.randperm(nb, generator=gen).to(torch.int32).view(batch, blocks_per_seq)
    seq_lens = torch.full((batch,), ctx + 1, dtype=torch.int32)
    slots = bt[:, ctx // BLOCK].to(torch.int64) * BLOCK + ctx % BLOCK
    inv = 500000.0 ** (-torch.arange(0, HEAD, 2).float() / HEAD)
    cos, sin = (ctx * inv).cos().to(mm_dtype), (ctx * inv).sin().to(mm_dtype)
    parts = (ctx + 1 + 511) // 512
    use_v1 = parts == 1 or batch * HEADS > 512  # paged_attn.py:118-119
    out = torch.empty(batch, HEADS, HEAD, dtype=torch.bfloat16)
    tmp = torch.empty(batch, HEADS, parts, HEAD, dtype=torch.bfloat16)
    es = torch.empty(batch, HEADS, parts)
    ml = torch.empty_like(es)

    def layer_step(L, res):
        hn = _rms(res, L["ln1"])
        qkv = L["qkv"](hn)
        q, k, v = qkv.split([HEADS * HEAD, KV_HEADS * HEAD, KV_HEADS * HEAD], dim=-1)
        q = _rope(q.view(-1, HEADS, HEAD), cos, sin).to(torch.bfloat16).contiguous()
        k = _rope(k.view(-1, KV_HEADS, HEAD), cos, sin).to(torch.bfloat16).contiguous()
        v = v.reshape(-1, KV_HEADS, HEAD).to(torch.bfloat16).contiguous()
        ref_cpu.reshape_and_cache(k, v, L["kc"], L["vc"], slots)
        if use_v1:
            ref_cpu.paged_attention_v1(out, q, L["kc"], L["vc"], KV_HEADS, s, bt, seq_lens, BLOCK, ctx + 1)
        else:
            ref_cpu.paged_attention_v2(out, es, ml, tmp, q, L["kc"], L["vc"], KV_HEADS, s, bt, seq_lens,
                                       BLOCK, ctx + 1)
        res = res + L["o"](out.view(batch, HEADS * HEAD).to(mm_dtype))
        hn = _rms(res, L["ln2"])
        gu = L["gate_up"](hn)
        act = torch.nn.functional.silu(gu[:, :INTER]) * gu[:, INTER:]
        return res + L["down"](act)

    def head_step(res):
        return (_rms(res, layers[0]["ln1"]) @ lm_head.t()).argmax(-1)

    with torch.no_grad():
        x = embed[tokens]
        for _ in range(max(warmup, 1)):
            r = x
            for L in layers:
                r = layer_step(L, r)
            head_step(r)
        t_layers = t_head = 0.0
        for _ in range(steps):
            r = x
            t0 = time.perf_counter()
            for L in layers:
                r = layer_step(L, r)
            t1 = time.perf_counter()
            head_step(r)
            t2 = time.perf_counter()
            t_layers += t1 - t0
            t_head += t2 - t1
    t_step = t_layers / steps * LAYERS / sample_layers + t_head / steps
    return {
        "tokens_per_s": batch / t_step, "ms_per_step": t_step * 1e3, "cores": cores,
        "ms_per_layer": t_layers / steps / sample_layers * 1e3, "ms_lm_head": t_head / steps * 1e3,
        "matmul_dtype": str(mm_dtype).replace("torch.", ""),
        "attention": "reference csrc/cpu paged_attention_" + ("v1" if use_v1 else "v2") + " bf16 KV",
        "sample": (f"{sample_layers} of {LAYERS} layers + lm_head timed x {steps} step(s) at batch {batch}, ctx {ctx}; "
                   f"step time = layers x {LAYERS}/{sample_layers} + lm_head; linears = torch dequant-matmul "
                   f"({str(mm_dtype).replace('torch.', '')}), attention/cache = compiled reference csrc/cpu ops"),
    }
