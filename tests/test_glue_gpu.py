"""GPU parity tests of the decode-side glue ops against oracle/glue_ref.py
(operator-by-operator restatement of the reference CUDA kernels) and the reference's own compiled
CPU kernels (oracle/_ref).  They mirror tests/kernels/test_layernorm.py, test_pos_encoding.py,
test_activation.py and test_cache.py of the reference.

Tolerances: rotary is bit-exact in fp16/bf16 (same rounding sequence, no transcendental);
RMSNorm / activations differ only through the fp32 summation order / libm ulps, so they must be
within 2 ulp of the element type with >= 99 % of elements identical; against the reference CPU
kernels the north_star bound rtol 1e-2 / atol 1e-3 (bf16: 2e-2, its own ulp) applies."""
import pytest
import torch

from oracle import glue_ref as gr
from oracle import ref_cpu

pytestmark = pytest.mark.gpu
DEV = "cuda"
DTYPES = [torch.half, torch.bfloat16, torch.float]
ULP = {torch.half: 2.0 ** -10, torch.bfloat16: 2.0 ** -7, torch.float: 2.0 ** -20}


def _near(out, ref, dtype, exact_frac=0.99):
    out, ref = out.cpu(), ref.cpu()
    # a 1-ulp (fp32) difference in rstd / libm can flip T(x * rstd) by one T-ulp, and the product
    # with the weight re-rounds: <= 2 ulp of T, on a handful of elements
    torch.testing.assert_close(out.float(), ref.float(), rtol=2 * ULP[dtype], atol=1e-5)
    if dtype is not torch.float and out.numel():
        assert (out == ref).float().mean().item() >= exact_frac


def _ref_cpu_close(out, ref, dtype):
    tol = dict(rtol=2e-2, atol=2e-2) if dtype is torch.bfloat16 else dict(rtol=1e-2, atol=1e-3)
    torch.testing.assert_close(out.cpu().float(), ref.float(), **tol)


# ------------------------------------------------------------------------------------------
# layernorm (test_layernorm.py: tokens [7, 83, 4096], hidden [768, 769, 770, 771, 5120..8199])
# ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("hidden", [8, 768, 769, 770, 4096, 5120, 8192, 8199, 32768])
@pytest.mark.parametrize("tokens", [1, 7, 83, 4096])
@pytest.mark.parametrize("add_residual", [False, True])
def test_rms_norm(ops, dtype, hidden, tokens, add_residual):
    if tokens == 4096 and hidden > 8199:
        pytest.skip("size covered at fewer tokens")
    g = torch.Generator().manual_seed(0)
    scale = 1 / (2 * hidden)
    x = (torch.randn(tokens, hidden, generator=g) * scale).to(dtype)
    res = (torch.randn(tokens, hidden, generator=g) * scale).to(dtype)
    w = (torch.randn(hidden, generator=g) * 0.1 + 1.0).to(dtype)
    if add_residual:
        ref_out, ref_res = gr.fused_add_rms_norm(x, res, w, 1e-6)
        xd, rd = x.to(DEV), res.to(DEV)
        ops.fused_add_rms_norm(xd, rd, w.to(DEV), 1e-6)
        assert torch.equal(rd.cpu(), ref_res)  # the residual stream is bit-exact
        _near(xd, ref_out, dtype)
    else:
        ref_out = gr.rms_norm(x, w, 1e-6)
        out = torch.empty(tokens, hidden, dtype=dtype, device=DEV)
        ops.rms_norm(out, x.to(DEV), w.to(DEV), 1e-6)
        _near(out, ref_out, dtype)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float])
def test_rms_norm_vs_reference_cpu_kernels(ops, dtype):
    if not ref_cpu.available():
        pytest.skip(str(ref_cpu.why_unavailable()))
    g = torch.Generator().manual_seed(1)
    x = torch.randn(83, 4096, generator=g).to(dtype)
    res = torch.randn(83, 4096, generator=g).to(dtype)
    w = (torch.randn(4096, generator=g) * 0.1 + 1.0).to(dtype)
    ref = torch.empty_like(x)
    torch.ops._C_ref.rms_norm(ref, x, w, 1e-5)
    out = torch.empty_like(x, device=DEV)
    ops.rms_norm(out, x.to(DEV), w.to(DEV), 1e-5)
    _ref_cpu_close(out, ref, dtype)
    xi, ri = x.clone(), res.clone()
    torch.ops._C_ref.fused_add_rms_norm(xi, ri, w, 1e-5)
    xd, rd = x.to(DEV), res.to(DEV)
    ops.fused_add_rms_norm(xd, rd, w.to(DEV), 1e-5)
    _ref_cpu_close(rd, ri, dtype)
    _ref_cpu_close(xd, xi, dtype)


@pytest.mark.parametrize("dtype", [torch.half, torch.bfloat16])
@pytest.mark.parametrize("hidden", [8, 768, 4096, 8192, 28672])
@pytest.mark.parametrize("tokens", [1, 7, 83, 2048])
@pytest.mark.parametrize("add_residual", [False, True])
@pytest.mark.parametrize("mode", ["int8_dynamic", "int8_static", "fp8_static"])
def test_rms_norm_quant_is_the_two_ops_back_to_back(ops, dtype, hidden, tokens, add_residual, mode):
    """(fused_add_)rms_norm + activation quant in one launch must be
    bit-identical to the reference sequence layernorm_kernels.cu:21-45/199-249 followed by
    int8_quant_kernels.cu:27-115 / fp8/common.cu:78-105, i.e. to this repo's two separately
    parity-tested ops, and must match the oracle composition (scales exact, values +-1 LSB as
    tests/kernels/test_int8_quant.py:25-71 allows)."""
    from oracle import quant_ref as qr
    if tokens * hidden > 2048 * 8192:
        pytest.skip("size")
    g = torch.Generator().manual_seed(tokens * 131 + hidden)
    x = torch.randn(tokens, hidden, generator=g).to(dtype)
    res = torch.randn(tokens, hidden, generator=g).to(dtype) if add_residual else None
    w = (torch.randn(hidden, generator=g) * 0.1 + 1).to(dtype)
    static_scale = torch.tensor([0.037], dtype=torch.float32)
    # unfused: this repo's ops
    xd = x.to(DEV)
    rd = res.to(DEV) if add_residual else None
    if add_residual:
        ops.fused_add_rms_norm(xd, rd, w.to(DEV), 1e-5)
        y = xd
    else:
        y = torch.empty_like(xd)
        ops.rms_norm(y, xd, w.to(DEV), 1e-5)
    if mode == "int8_dynamic":
        q_ref, s_ref = ops.scaled_int8_quant(y)
    elif mode == "int8_static":
        q_ref, s_ref = ops.scaled_int8_quant(y, static_scale.to(DEV))
    else:
        q_ref, s_ref = ops.scaled_fp8_quant(y, static_scale.to(DEV))
    # fused
    r2 = res.to(DEV) if add_residual else None
    q, s = ops.rms_norm_quant(x.to(DEV), w.to(DEV), 1e-5, residual=r2,
                              scale=None if mode == "int8_dynamic" else static_scale.to(DEV),
                              out_dtype=torch.float8_e4m3fn if mode == "fp8_static" else torch.int8)
    assert torch.equal(q.view(torch.uint8), q_ref.view(torch.uint8))
    assert torch.equal(s, s_ref)
    if add_residual:
        assert torch.equal(r2, rd)
    # oracle composition on the CPU (small cases)
    if tokens <= 83:
        if add_residual:
            y_o, _ = gr.fused_add_rms_norm(x, res, w, 1e-5)
        else:
            y_o = gr.rms_norm(x, w, 1e-5)
        if mode == "fp8_static":
            q_o, _ = qr.scaled_fp8_quant_ref(y_o, static_scale)
            torch.testing.assert_close(q.cpu().float(), q_o.float(), rtol=0.13, atol=1e-3)  # <= 1 fp8 ulp
        else:
            q_o, s_o = qr.scaled_int8_quant_ref(y_o, None if mode == "int8_dynamic" else static_scale)
            assert (q.cpu().int() - q_o.int()).abs().max() <= 1
            if mode == "int8_dynamic":
                torch.testing.assert_close(s.cpu().view(-1), s_o.view(-1).float(), rtol=2e-3, atol=0)


def test_rms_norm_quant_rejects(ops):
    x = torch.randn(4, 64, dtype=torch.half, device=DEV)
    w = torch.ones(64, dtype=torch.half, device=DEV)
    with pytest.raises(RuntimeError):
        ops.rms_norm_quant(x, w, 1e-5, out_dtype=torch.float8_e4m3fn)            # dynamic fp8: not fused
    with pytest.raises(RuntimeError):
        ops.rms_norm_quant(x.float(), w.float(), 1e-5)                            # fp32 rows
    with pytest.raises(RuntimeError):
        ops.rms_norm_quant(x, w, 1e-5, residual=torch.zeros(4, 32, dtype=torch.half, device=DEV))


def test_rms_norm_empty_and_errors(ops):
    x = torch.empty(0, 128, dtype=torch.half, device=DEV)
    ops.rms_norm(torch.empty_like(x), x, torch.ones(128, dtype=torch.half, device=DEV), 1e-6)
    with pytest.raises(RuntimeError):  # dtype mismatch is rejected, not converted
        ops.rms_norm(torch.empty(2, 128, dtype=torch.half, device=DEV),
                     torch.zeros(2, 128, dtype=torch.half, device=DEV),
                     torch.ones(128, dtype=torch.float, device=DEV), 1e-6)


# ------------------------------------------------------------------------------------------
# rotary (test_pos_encoding.py: head sizes [64..256], rotary_dim None/32, neox True/False)
# ------------------------------------------------------------------------------------------
def _cache(rot_dim, max_pos, dtype, base=10000.0):
    inv = 1.0 / (base ** (torch.arange(0, rot_dim, 2).float() / rot_dim))
    ang = torch.arange(max_pos).float()[:, None] * inv[None]
    return torch.cat((ang.cos(), ang.sin()), -1).to(dtype)


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("is_neox", [True, False])
@pytest.mark.parametrize("head_size,rot_dim", [(64, 64), (80, 80), (96, 32), (112, 112), (128, 128),
                                               (128, 64), (192, 192), (256, 256), (128, 20)])
@pytest.mark.parametrize("tokens", [1, 11, 2048])
def test_rotary_embedding(ops, dtype, is_neox, head_size, rot_dim, tokens):
    g = torch.Generator().manual_seed(3)
    heads, kv_heads, max_pos = 8, 2, 8192
    cache = _cache(rot_dim, max_pos, dtype)
    pos = torch.randint(0, max_pos, (tokens,), generator=g)
    # q and k are slices of one qkv row (non-contiguous token stride), as in the model
    qkv = torch.randn(tokens, (heads + 2 * kv_heads) * head_size, generator=g).to(dtype)
    q, k = qkv[:, : heads * head_size], qkv[:, heads * head_size: (heads + kv_heads) * head_size]
    ref_q, ref_k = gr.rotary_embedding(pos, q, k, head_size, cache, is_neox)
    qkv_d = qkv.to(DEV)
    qd, kd = qkv_d[:, : heads * head_size], qkv_d[:, heads * head_size: (heads + kv_heads) * head_size]
    ops.rotary_embedding(pos.to(DEV), qd, kd, head_size, cache.to(DEV), is_neox)
    if dtype is torch.float:  # fp32: the reference build may contract x*cos - y*sin into an fma
        torch.testing.assert_close(qd.cpu(), ref_q, rtol=1e-5, atol=1e-5)
        torch.testing.assert_close(kd.cpu(), ref_k, rtol=1e-5, atol=1e-5)
    else:
        assert torch.equal(qd.cpu(), ref_q) and torch.equal(kd.cpu(), ref_k)
    # v part of the row untouched
    assert torch.equal(qkv_d[:, (heads + kv_heads) * head_size:].cpu(), qkv[:, (heads + kv_heads) * head_size:])


@pytest.mark.parametrize("dtype", [torch.half, torch.bfloat16])
@pytest.mark.parametrize("is_neox", [True, False])
def test_batched_rotary_embedding(ops, dtype, is_neox):
    g = torch.Generator().manual_seed(4)
    tokens, heads, kv_heads, head_size, rot_dim = 37, 8, 8, 128, 128
    cache = torch.cat([_cache(rot_dim, 1024, dtype, b) for b in (10000.0, 500000.0, 1e6)])
    pos = torch.randint(0, 1024, (tokens,), generator=g)
    off = torch.randint(0, 3, (tokens,), generator=g) * 1024
    q = torch.randn(tokens, heads * head_size, generator=g).to(dtype)
    k = torch.randn(tokens, kv_heads * head_size, generator=g).to(dtype)
    ref_q, ref_k = gr.rotary_embedding(pos, q, k, head_size, cache, is_neox, offsets=off)
    qd, kd = q.to(DEV), k.to(DEV)
    ops.batched_rotary_embedding(pos.to(DEV), qd, kd, head_size, cache.to(DEV), is_neox, rot_dim, off.to(DEV))
    assert torch.equal(qd.cpu(), ref_q) and torch.equal(kd.cpu(), ref_k)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float])
@pytest.mark.parametrize("is_neox", [True, False])
def test_rotary_vs_reference_cpu_kernels(ops, dtype, is_neox):
    if not ref_cpu.available():
        pytest.skip(str(ref_cpu.why_unavailable()))
    g = torch.Generator().manual_seed(5)
    tokens, heads, kv_heads, head_size = 64, 32, 8, 128
    cache = _cache(head_size, 4096, dtype, 500000.0)
    pos = torch.randint(0, 4096, (tokens,), generator=g)
    q = torch.randn(tokens, heads * head_size, generator=g).to(dtype)
    k = torch.randn(tokens, kv_heads * head_size, generator=g).to(dtype)
    qi, ki = q.clone(), k.clone()
    torch.ops._C_ref.rotary_embedding(pos, qi, ki, head_size, cache, is_neox)
    qd, kd = q.to(DEV), k.to(DEV)
    ops.rotary_embedding(pos.to(DEV), qd, kd, head_size, cache.to(DEV), is_neox)
    _ref_cpu_close(qd, qi, dtype)
    _ref_cpu_close(kd, ki, dtype)


# fused extension == rotary_embedding then reshape_and_cache, bit for bit (incl. fp8 caches, padding slots)
@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("kv_cache_dtype", ["auto", "fp8", "fp8_e5m2"])
@pytest.mark.parametrize("is_neox", [True, False])
@pytest.mark.parametrize("heads,kv_heads,head_size,rot_dim", [(32, 8, 128, 128), (12, 12, 64, 64), (8, 2, 128, 64)])
def test_rotary_embedding_and_cache(ops, dtype, kv_cache_dtype, is_neox, heads, kv_heads, head_size, rot_dim):
    g = torch.Generator().manual_seed(6)
    tokens, bs, nb = 43, 16, 64
    cache_dt = dtype if kv_cache_dtype == "auto" else torch.uint8
    x = 16 // torch.tensor([], dtype=cache_dt).element_size()
    cache = _cache(rot_dim, 2048, dtype).to(DEV)
    pos = torch.randint(0, 2048, (tokens,), generator=g).to(DEV)
    slots = torch.randperm(nb * bs, generator=g)[:tokens]
    slots[5] = -1  # padding token
    slots = slots.to(DEV)
    qkv = torch.randn(tokens, (heads + 2 * kv_heads) * head_size, generator=g).to(dtype).to(DEV)
    a, b = heads * head_size, (heads + kv_heads) * head_size

    def caches():
        return (torch.zeros(nb, kv_heads, head_size // x, bs, x, dtype=cache_dt, device=DEV),
                torch.zeros(nb, kv_heads, head_size, bs, dtype=cache_dt, device=DEV))

    r = qkv.clone()
    kc0, vc0 = caches()
    ops.rotary_embedding(pos, r[:, :a], r[:, a:b], head_size, cache, is_neox)
    ops.reshape_and_cache(r[:, a:b].view(tokens, kv_heads, head_size), r[:, b:].view(tokens, kv_heads, head_size),
                          kc0, vc0, slots, kv_cache_dtype, 0.5)
    f = qkv.clone()
    kc1, vc1 = caches()
    ops.rotary_embedding_and_cache(pos, f[:, :a], f[:, a:b], f[:, b:], head_size, cache, is_neox, kc1, vc1, slots,
                                   kv_cache_dtype, 0.5)
    assert torch.equal(f, r)
    assert torch.equal(kc1, kc0) and torch.equal(vc1, vc0)
    assert kc1.any() and vc1.any()


# ------------------------------------------------------------------------------------------
# activations (test_activation.py: tokens [7, 83, 2048], d [512, 13824])
# ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("act", ["silu", "gelu", "gelu_tanh"])
@pytest.mark.parametrize("d", [1, 13, 512, 13824, 14336])
@pytest.mark.parametrize("tokens", [1, 7, 83, 2048])
def test_act_and_mul(ops, dtype, act, d, tokens):
    g = torch.Generator().manual_seed(7)
    x = (torch.randn(tokens, 2 * d, generator=g) * 2).to(dtype)
    ref = gr.act_and_mul(x, act)
    out = torch.empty(tokens, d, dtype=dtype, device=DEV)
    getattr(ops, f"{act}_and_mul")(out, x.to(DEV))
    # libm (expf / erff / tanhf) differs between the GPU and the host by an fp32 ulp or two
    _near(out, ref, dtype, exact_frac=0.95)


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("act", ["gelu_new", "gelu_fast", "gelu_quick"])
@pytest.mark.parametrize("d", [13, 512, 13824])
@pytest.mark.parametrize("tokens", [7, 2048])
def test_activation(ops, dtype, act, d, tokens):
    g = torch.Generator().manual_seed(8)
    x = (torch.randn(tokens, d, generator=g) * 2).to(dtype)
    ref = getattr(gr, act)(x)
    out = torch.empty_like(x, device=DEV)
    getattr(ops, act)(out, x.to(DEV))
    # tanh of a T-rounded argument: a 1-ulp libm difference can move the T result by 1 ulp
    out_c = out.cpu()
    torch.testing.assert_close(out_c.float(), ref.float(), rtol=2 * ULP[dtype], atol=2 * ULP[dtype])
    if dtype is not torch.float:
        assert (out_c == ref).float().mean().item() >= 0.95


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float])
def test_activations_vs_reference_cpu_kernels(ops, dtype):
    if not ref_cpu.available():
        pytest.skip(str(ref_cpu.why_unavailable()))
    g = torch.Generator().manual_seed(9)
    d = 14336
    x = (torch.randn(83, 2 * d, generator=g) * 2).to(dtype)
    for name in ("silu", "gelu", "gelu_tanh"):
        ref = torch.empty(83, d, dtype=dtype)
        getattr(torch.ops._C_ref, f"{name}_and_mul")(ref, x)
        out = torch.empty(83, d, dtype=dtype, device=DEV)
        getattr(ops, f"{name}_and_mul")(out, x.to(DEV))
        _ref_cpu_close(out, ref, dtype)


def test_glue_ops_cuda_graph(ops):
    """All glue ops are capturable (no sync / allocation inside)."""
    g = torch.Generator().manual_seed(10)
    x = torch.randn(8, 4096, generator=g).half().to(DEV)
    res = torch.randn(8, 4096, generator=g).half().to(DEV)
    w = torch.ones(4096, dtype=torch.half, device=DEV)
    gu = torch.randn(8, 2 * 14336, generator=g).half().to(DEV)
    out = torch.empty(8, 14336, dtype=torch.half, device=DEV)
    x0, r0 = x.clone(), res.clone()
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        gph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gph, stream=s):
            ops.fused_add_rms_norm(x, res, w, 1e-5)
            ops.silu_and_mul(out, gu)
        x.copy_(x0), res.copy_(r0)
        gph.replay()
        s.synchronize()
    ref_out, ref_res = gr.fused_add_rms_norm(x0.cpu(), r0.cpu(), w.cpu(), 1e-5)
    assert torch.equal(res.cpu(), ref_res)
    _near(x, ref_out, torch.half)
    _near(out, gr.act_and_mul(gu.cpu(), "silu"), torch.half, exact_frac=0.98)


def test_pdl_chain_matches_serial(ops):
    """Programmatic dependent launch (glue kernels and the decode GEMM become resident before their
    predecessor finishes) must not change a single bit: one decoder-layer MLP chain replayed from a
    CUDA graph with PDL edges == the same chain launched eagerly with B200_PDL=0."""
    import os
    g = torch.Generator().manual_seed(11)
    hidden, inter = 4096, 14336

    def qlin(k, n):
        q = torch.randint(-2**31, 2**31 - 1, (k // 16, n * 2), dtype=torch.int32, generator=g).to(DEV)
        s = (torch.rand(k // 128, n, generator=g) * 0.004 + 0.0005).half().to(DEV)
        ws = torch.zeros(max(n // 64, 1) * 16, dtype=torch.int32, device=DEV)
        e = torch.empty(0, dtype=torch.int32, device=DEV)
        return lambda x: ops.gptq_marlin_gemm(x, q, s, e, e, ws, 4, x.shape[0], n, k, True)

    gate_up, down = qlin(hidden, 2 * inter), qlin(inter, hidden)
    w = (torch.randn(hidden, generator=g) * 0.1 + 1).half().to(DEV)
    for m in (1, 8, 33):
        x0 = torch.randn(m, hidden, generator=g).half().to(DEV)
        r0 = torch.randn(m, hidden, generator=g).half().to(DEV)

        def chain(x, res):
            for _ in range(3):  # three "layers": every PDL edge kind (glue->GEMM, GEMM->glue) repeats
                ops.fused_add_rms_norm(x, res, w, 1e-5)
                gu = gate_up(x)
                act = torch.empty(m, inter, dtype=torch.half, device=DEV)
                ops.silu_and_mul(act, gu)
                x = down(act)
            return x, res

        os.environ["B200_PDL"] = "0"
        try:
            ref_x, ref_r = chain(x0.clone(), r0.clone())
            torch.cuda.synchronize()
        finally:
            os.environ.pop("B200_PDL")
        xs, rs = x0.clone(), r0.clone()
        s = torch.cuda.Stream()
        with torch.cuda.stream(s):
            chain(xs.clone(), rs.clone())  # warm-up outside capture
            s.synchronize()
            gph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gph, stream=s):
                out_x, out_r = chain(xs, rs)
            for _ in range(3):
                xs.copy_(x0), rs.copy_(r0)
                gph.replay()
            s.synchronize()
        assert torch.equal(out_x, ref_x) and torch.equal(out_r, ref_r)
        assert torch.isfinite(out_x.float()).all()


# ---- gptq_marlin_gemm + fused_add_rms_norm as one operator (extension) ----

@pytest.mark.parametrize("dtype", [torch.half, torch.bfloat16])
@pytest.mark.parametrize("m", [1, 4, 5, 16, 33, 64, 128])
@pytest.mark.parametrize("k,n", [(4096, 4096), (14336, 4096), (512, 256)])
def test_marlin_gemm_add_rms_norm_matches_the_unfused_pair(m, k, n, dtype):
    """Bit-identical to gptq_marlin_gemm followed by fused_add_rms_norm, whichever path the GEMM takes
    (in-kernel reduce at M <= 4, split-K partials summed inside the norm kernel, or no split)."""
    import neural_magic_vllm_b200 as pkg
    from oracle import quant_ref as qr
    ops = pkg.ops
    torch.manual_seed(m + n)
    w = torch.randn(k, n, dtype=dtype) * 0.05
    _, mq, ms, _, _, _ = qr.marlin_quantize(w, 4, 128, False)
    mq, ms = mq.to(DEV), ms.to(DEV)
    e = torch.empty(0, dtype=torch.int32, device=DEV)
    ws = torch.zeros(n // 64 * 16, dtype=torch.int32, device=DEV)
    a = torch.randn(m, k, dtype=dtype, device=DEV)
    res0 = torch.randn(m, n, dtype=dtype, device=DEV)
    nw = (torch.rand(n, device=DEV) + 0.5).to(dtype)
    # unfused
    c = ops.gptq_marlin_gemm(a, mq, ms, e, e, ws, 4, m, n, k, True)
    res_a = res0.clone()
    ops.fused_add_rms_norm(c, res_a, nw, 1e-5)
    # fused, twice (the workspace / lock buffer must come back clean)
    for _ in range(2):
        res_b = res0.clone()
        out = ops.gptq_marlin_gemm_add_rms_norm(a, mq, ms, e, e, ws, 4, m, n, k, True, res_b, nw, 1e-5)
        assert torch.equal(res_a, res_b)
        assert torch.equal(c, out)
    assert int(ws.abs().max()) == 0
