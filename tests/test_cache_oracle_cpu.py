"""Pins oracle/cache_ref.py (CPU, no GPU needed).

* reshape_and_cache / copy_blocks against the reference's OWN compiled CPU kernels
  (oracle/_ref = csrc/cpu/cache.cpp) on fp32 and bf16, incl. negative slots and a strided key.
* fp8 conversions against torch's float8 casts on every byte code and against the reference
  test's round-trip bound (tests/kernels/test_cache.py:369-394: U(-224, 224), atol 1e-3 rtol 0.1).
"""
import random

import pytest
import torch

from oracle import cache_ref as cr
from oracle import ref_cpu

needs_ref = pytest.mark.skipif(not ref_cpu.available(), reason=str(ref_cpu.why_unavailable()))


@needs_ref
@pytest.mark.parametrize("dtype", [torch.float, torch.bfloat16])
@pytest.mark.parametrize("head_size", [64, 80, 128])
def test_reshape_and_cache_vs_reference_cpu(dtype, head_size):
    torch.manual_seed(0)
    T, H, NB, bs = 42, 8, 64, 16
    x = 16 // torch.tensor([], dtype=dtype).element_size()
    qkv = torch.randn(T, 3, H, head_size).to(dtype)
    _, key, value = qkv.unbind(dim=1)  # strided rows, as the reference test builds them
    kc = torch.randn(NB, H, head_size // x, bs, x).to(dtype)
    vc = torch.randn(NB, H, head_size, bs).to(dtype)
    slots = torch.tensor(random.Random(0).sample(range(NB * bs), T), dtype=torch.int64)
    slots[5] = -1  # padding token: must be skipped (csrc/cpu/cache.cpp:54)
    kc2, vc2 = kc.clone(), vc.clone()
    ref_cpu.reshape_and_cache(key, value, kc, vc, slots)
    cr.reshape_and_cache(key, value, kc2, vc2, slots)
    assert torch.equal(kc.view(torch.uint8), kc2.view(torch.uint8))
    assert torch.equal(vc.view(torch.uint8), vc2.view(torch.uint8))


@needs_ref
@pytest.mark.parametrize("dtype", [torch.float, torch.bfloat16])
def test_copy_blocks_vs_reference_cpu(dtype):
    torch.manual_seed(0)
    rnd = random.Random(0)
    NB, H, D, bs, L, P = 128, 8, 64, 16, 3, 24
    x = 16 // torch.tensor([], dtype=dtype).element_size()
    kcs = [torch.randn(NB, H, D // x, bs, x).to(dtype) for _ in range(L)]
    vcs = [torch.randn(NB, H, D, bs).to(dtype) for _ in range(L)]
    src = rnd.sample(range(NB), P)
    rest = sorted(set(range(NB)) - set(src))
    dst = rnd.sample(rest, 2 * P)
    mapping = torch.tensor([(s, dst[2 * i + j]) for i, s in enumerate(src) for j in range(2)],
                           dtype=torch.int64)
    k2, v2 = [k.clone() for k in kcs], [v.clone() for v in vcs]
    ref_cpu.copy_blocks(kcs, vcs, mapping)
    cr.copy_blocks(k2, v2, mapping)
    for a, b in zip(kcs + vcs, k2 + v2):
        assert torch.equal(a.view(torch.uint8), b.view(torch.uint8))


@pytest.mark.parametrize("kv", ["fp8", "fp8_e5m2"])
def test_fp8_dequant_every_code(kv):
    codes = torch.arange(256, dtype=torch.uint8)
    got = cr.fp8_dequantize(codes, torch.float, 1.0, kv)
    want = codes.view(cr.FP8_DTYPES[kv]).float()
    ok = (got == want) | (torch.isnan(got) & torch.isnan(want))
    assert bool(ok.all())
    # re-quantising a representable value is the identity (NaN codes aside)
    back = cr.fp8_quantize(want, 1.0, kv)
    finite = torch.isfinite(want)
    assert torch.equal(back[finite], codes[finite])


@pytest.mark.parametrize("dtype", [torch.half, torch.bfloat16, torch.float])
def test_fp8_e4m3_round_trip_bound(dtype):
    g = torch.Generator().manual_seed(0)
    cache = ((torch.rand(64, 8, 64, 16, generator=g) * 2 - 1) * 224.0).to(dtype)
    u8 = torch.empty_like(cache, dtype=torch.uint8)
    cr.convert_fp8(u8, cache)
    back = torch.empty_like(cache)
    cr.convert_fp8(back, u8)
    assert torch.allclose(cache, back, atol=0.001, rtol=0.1)
    # saturating: out-of-range values clamp to +-448, never NaN/inf
    big = torch.tensor([1e6, -1e6, 449.0, -500.0]).to(dtype)
    q = cr.fp8_quantize(big, 1.0, "fp8")
    assert torch.equal(cr.fp8_dequantize(q, torch.float), torch.tensor([448.0, -448.0, 448.0, -448.0]))


def test_fp8_scale_is_divided_then_multiplied():
    x = torch.tensor([1.0, -3.0, 100.0, 0.3])
    q = cr.fp8_quantize(x, 0.5, "fp8")  # cvt(x / 0.5)
    assert torch.equal(q, (x / 0.5).to(torch.float8_e4m3fn).view(torch.uint8))
    d = cr.fp8_dequantize(q, torch.half, 0.5, "fp8")
    assert torch.equal(d, (q.view(torch.float8_e4m3fn).to(torch.half).float() * 0.5).half())


def test_flash_layout_and_swap_semantics():
    torch.manual_seed(0)
    T, H, D, NB, bs = 11, 4, 32, 6, 8
    key, value = torch.randn(T, H, D), torch.randn(T, H, D)
    kc, vc = torch.zeros(NB, bs, H, D), torch.zeros(NB, bs, H, D)
    slots = torch.randperm(NB * bs)[:T].to(torch.int64)
    slots[0] = -7
    cr.reshape_and_cache_flash(key, value, kc, vc, slots)
    for i, s in enumerate(slots.tolist()):
        if s < 0:
            continue
        assert torch.equal(kc[s // bs, s % bs], key[i]) and torch.equal(vc[s // bs, s % bs], value[i])
    assert int((kc != 0).any(-1).any(-1).sum()) == T - 1
    src, dst = torch.randn(NB, 100), torch.zeros(NB, 100)
    cr.swap_blocks(src, dst, torch.tensor([[0, 3], [2, 1]]))
    assert torch.equal(dst[3], src[0]) and torch.equal(dst[1], src[2]) and not dst[0].any()
