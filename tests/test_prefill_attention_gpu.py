"""Prefill attention: the tcgen05 varlen causal kernel against the oracle
restatement of the reference's CPU prefill path (vllm/attention/backends/torch_sdpa.py:203-218,
per-sequence scaled_dot_product_attention) on identical inputs, and against torch's own fp32 SDPA
on the GPU at the BASELINE C3 prefill size (2k tokens, 32 q / 8 kv heads, D 128).
Tolerance: north_star's rtol 1e-2 / atol 1e-3 for fp16 (bf16: its ulp, 2e-2)."""
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda"


@pytest.fixture(scope="module")
def ops():
    import neural_magic_vllm_b200 as pkg
    return pkg.ops


def _run(ops, lens_q, lens_k, heads, kv_heads, dtype, causal, seed=0, scale_in=1.0, qkv_packed=False, d=128):
    from oracle import attention_ref as ar
    g = torch.Generator().manual_seed(seed)
    tq, tk = sum(lens_q), sum(lens_k)
    if qkv_packed:  # q / k / v are column slices of one fused qkv row, as the model produces them
        assert lens_q == lens_k
        qkv = (torch.randn(tq, (heads + 2 * kv_heads) * d, generator=g) * scale_in).to(dtype)
        q = qkv[:, :heads * d].view(tq, heads, d)
        k = qkv[:, heads * d:(heads + kv_heads) * d].view(tk, kv_heads, d)
        v = qkv[:, (heads + kv_heads) * d:].view(tk, kv_heads, d)
        dev = qkv.to(DEV)
        qd = dev[:, :heads * d].view(tq, heads, d)
        kd = dev[:, heads * d:(heads + kv_heads) * d].view(tk, kv_heads, d)
        vd = dev[:, (heads + kv_heads) * d:].view(tk, kv_heads, d)
    else:
        q = (torch.randn(tq, heads, d, generator=g) * scale_in).to(dtype)
        k = (torch.randn(tk, kv_heads, d, generator=g) * scale_in).to(dtype)
        v = torch.randn(tk, kv_heads, d, generator=g).to(dtype)
        qd, kd, vd = q.to(DEV), k.to(DEV), v.to(DEV)
    cq = torch.tensor([0] + list(torch.tensor(lens_q).cumsum(0)), dtype=torch.int32)
    ck = torch.tensor([0] + list(torch.tensor(lens_k).cumsum(0)), dtype=torch.int32)
    scale = d ** -0.5
    out = ops.flash_attn_varlen_func(qd, kd, vd, cq.to(DEV), ck.to(DEV), max(lens_q), max(lens_k), scale, causal)
    ref = ar.varlen_attention_ref(q, k, v, cq, ck, scale, causal)
    tol = dict(rtol=1e-2, atol=1e-3) if dtype == torch.half else dict(rtol=2e-2, atol=2e-2)
    torch.testing.assert_close(out.cpu().float(), ref, **tol)


@pytest.mark.parametrize("dtype", [torch.half, torch.bfloat16])
@pytest.mark.parametrize("causal", [True, False])
@pytest.mark.parametrize("lens", [[1], [7], [128], [129], [1, 100, 513, 31], [257, 256, 255], [640]])
def test_varlen_attention_ragged(ops, dtype, causal, lens):
    _run(ops, lens, lens, 8, 2, dtype, causal)


@pytest.mark.parametrize("heads,kv_heads", [(4, 4), (8, 1), (32, 8)])
def test_varlen_attention_gqa(ops, heads, kv_heads):
    _run(ops, [200, 77], [200, 77], heads, kv_heads, torch.half, True, seed=3)


@pytest.mark.parametrize("dtype", [torch.half, torch.bfloat16])
@pytest.mark.parametrize("causal", [True, False])
def test_varlen_attention_head_size_64(ops, dtype, causal):
    _run(ops, [1, 100, 513, 31], [1, 100, 513, 31], 8, 2, dtype, causal, seed=9, d=64)
    _run(ops, [5, 130], [300, 130], 4, 4, dtype, causal, seed=10, d=64)


def test_varlen_attention_q_shorter_than_k(ops):
    # chunked prefill / prefix: the causal diagonal is aligned to the END of the keys
    _run(ops, [5, 130, 1], [300, 130, 64], 8, 2, torch.half, True, seed=4)
    _run(ops, [5, 130, 1], [300, 130, 64], 8, 2, torch.half, False, seed=5)


def test_varlen_attention_fused_qkv_rows_and_large_logits(ops):
    _run(ops, [300, 45], [300, 45], 8, 2, torch.half, True, seed=6, qkv_packed=True)
    _run(ops, [260], [260], 4, 2, torch.half, True, seed=7, scale_in=4.0)  # peaked softmax


def test_varlen_attention_empty_sequence_in_batch(ops):
    _run(ops, [64, 0, 3], [64, 0, 3], 4, 2, torch.half, True, seed=8)


def test_varlen_attention_baseline_prefill_shape(ops):
    # BASELINE configs[2]: 2k-token prefill at the Llama-3-8B shape; GPU fp32 SDPA as the checker
    torch.manual_seed(0)
    t, h, kvh, d = 2048, 32, 8, 128
    q = torch.randn(t, h, d, dtype=torch.half, device=DEV)
    k = torch.randn(t, kvh, d, dtype=torch.half, device=DEV)
    v = torch.randn(t, kvh, d, dtype=torch.half, device=DEV)
    cu = torch.tensor([0, t], dtype=torch.int32, device=DEV)
    out = ops.flash_attn_varlen_func(q, k, v, cu, cu, t, t, d ** -0.5, True)
    ref = torch.nn.functional.scaled_dot_product_attention(
        q.float().transpose(0, 1), k.float().transpose(0, 1).repeat_interleave(h // kvh, 0),
        v.float().transpose(0, 1).repeat_interleave(h // kvh, 0), is_causal=True).transpose(0, 1)
    torch.testing.assert_close(out.float(), ref, rtol=1e-2, atol=1e-3)


def test_varlen_attention_rejects(ops):
    q = torch.zeros(4, 2, 96, dtype=torch.half, device=DEV)
    cu = torch.tensor([0, 4], dtype=torch.int32, device=DEV)
    with pytest.raises(NotImplementedError):
        ops.flash_attn_varlen_func(q, q, q, cu, cu, 4, 4, 0.1, True)             # head_size 96 is not built
    q = torch.zeros(4, 3, 128, dtype=torch.half, device=DEV)
    k = torch.zeros(4, 2, 128, dtype=torch.half, device=DEV)
    with pytest.raises(RuntimeError):
        ops.flash_attn_varlen_func(q, k, k, cu, cu, 4, 4, 0.1, True)             # 3 heads over 2 kv heads


@pytest.mark.parametrize("d", [64, 128])
def test_varlen_attention_writes_only_its_rows(ops, d):
    """NaN canaries around the output and past the last token of q / k / v: nothing outside
    out[0 : total_q] is written, and rows the TMA boxes read past the tensors' ends (zero-filled) never
    reach a result (lengths that leave partly empty 128-row tiles)."""
    torch.manual_seed(3)
    lens = [130, 1, 257]
    t, h, kvh = sum(lens), 4, 2
    pad = 300
    buf = torch.full((pad + t + pad, h, d), float("nan"), dtype=torch.half, device=DEV)
    out = buf[pad:pad + t]
    q = torch.randn(t, h, d, dtype=torch.half, device=DEV)
    k = torch.randn(t, kvh, d, dtype=torch.half, device=DEV)
    v = torch.randn(t, kvh, d, dtype=torch.half, device=DEV)
    cu = torch.tensor([0, 130, 131, 388], dtype=torch.int32, device=DEV)
    ops.flash_attn_varlen_func(q, k, v, cu, cu, max(lens), max(lens), d ** -0.5, True, out)
    torch.cuda.synchronize()
    assert torch.isnan(buf[:pad]).all() and torch.isnan(buf[pad + t:]).all()
    assert not torch.isnan(out).any()
    from oracle import attention_ref as ar
    ref = ar.varlen_attention_ref(q.cpu(), k.cpu(), v.cpu(), cu.cpu(), cu.cpu(), d ** -0.5, True)
    torch.testing.assert_close(out.cpu().float(), ref, rtol=1e-2, atol=1e-3)
