"""Pins oracle/glue_ref.py (restatement of the reference's layernorm / rotary / activation CUDA
kernels) against the reference's OWN CPU kernels (csrc/cpu/{layernorm,pos_encoding,activation}.cpp,
compiled in place into oracle/_ref) on fp32 and bf16 -- mirrors tests/kernels/test_layernorm.py,
test_pos_encoding.py, test_activation.py of the reference."""
import pytest
import torch

from oracle import glue_ref as gr
from oracle import ref_cpu

needs_ref = pytest.mark.skipif(not ref_cpu.available(), reason=str(ref_cpu.why_unavailable()))
TOL = {torch.float: dict(atol=1e-5, rtol=1e-5), torch.bfloat16: dict(atol=2e-2, rtol=2e-2)}


def _close(a, b, dtype):
    torch.testing.assert_close(a.float(), b.float(), **TOL[dtype])


@needs_ref
@pytest.mark.parametrize("dtype", [torch.float, torch.bfloat16])
@pytest.mark.parametrize("hidden", [768, 4096, 5120])
@pytest.mark.parametrize("tokens", [1, 7, 83])
def test_rms_norm_vs_reference_cpu(dtype, hidden, tokens):
    g = torch.Generator().manual_seed(0)
    x = (torch.randn(tokens, hidden, generator=g) * 0.5).to(dtype)
    w = (torch.randn(hidden, generator=g) * 0.1 + 1).to(dtype)
    out = torch.empty_like(x)
    torch.ops._C_ref.rms_norm(out, x, w, 1e-6)
    _close(gr.rms_norm(x, w, 1e-6), out, dtype)
    res = (torch.randn(tokens, hidden, generator=g) * 0.5).to(dtype)
    xi, ri = x.clone(), res.clone()
    torch.ops._C_ref.fused_add_rms_norm(xi, ri, w, 1e-6)
    o, z = gr.fused_add_rms_norm(x, res, w, 1e-6)
    _close(z, ri, dtype)
    _close(o, xi, dtype)


@needs_ref
@pytest.mark.parametrize("dtype", [torch.float, torch.bfloat16])
@pytest.mark.parametrize("is_neox", [True, False])
@pytest.mark.parametrize("head_size,rot_dim", [(64, 64), (128, 128), (128, 64), (80, 32)])
def test_rotary_vs_reference_cpu(dtype, is_neox, head_size, rot_dim):
    g = torch.Generator().manual_seed(1)
    tokens, heads, kv_heads, max_pos = 11, 8, 2, 512
    inv = 1.0 / (10000 ** (torch.arange(0, rot_dim, 2).float() / rot_dim))
    ang = torch.arange(max_pos).float()[:, None] * inv[None]
    cache = torch.cat((ang.cos(), ang.sin()), -1).to(dtype)
    pos = torch.randint(0, max_pos, (tokens,), generator=g)
    q = torch.randn(tokens, heads * head_size, generator=g).to(dtype)
    k = torch.randn(tokens, kv_heads * head_size, generator=g).to(dtype)
    qr_, kr_ = gr.rotary_embedding(pos, q, k, head_size, cache, is_neox)
    qi, ki = q.clone(), k.clone()
    torch.ops._C_ref.rotary_embedding(pos, qi, ki, head_size, cache, is_neox)
    _close(qr_, qi, dtype)
    _close(kr_, ki, dtype)


@needs_ref
@pytest.mark.parametrize("dtype", [torch.float, torch.bfloat16])
@pytest.mark.parametrize("d", [512, 13824])
def test_activations_vs_reference_cpu(dtype, d):
    g = torch.Generator().manual_seed(2)
    x = (torch.randn(7, 2 * d, generator=g) * 2).to(dtype)
    for name in ("silu", "gelu", "gelu_tanh"):
        out = torch.empty(7, d, dtype=dtype)
        getattr(torch.ops._C_ref, f"{name}_and_mul")(out, x)
        _close(gr.act_and_mul(x, name), out, dtype)
    y = x[:, :d].contiguous()
    for name in ("gelu_new", "gelu_fast", "gelu_quick"):
        out = torch.empty_like(y)
        getattr(torch.ops._C_ref, name)(out, y)
        _close(getattr(gr, name)(y), out, dtype)
