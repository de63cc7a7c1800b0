"""GPU parity tests of the paged KV-cache maintenance ops
against oracle/cache_ref.py (pinned in tests/test_cache_oracle_cpu.py) and, where the compiled
reference CPU backend covers the case (fp32 / bf16, "auto"), against oracle/_ref itself.

Mirrors the reference's tests/kernels/test_cache.py:
  test_copy_blocks               :47-113   exact
  test_reshape_and_cache         :124-204  exact ("auto"); fp8: the reference's own bound
                                           atol 1e-3 / rtol 0.1 after convert_fp8 AND bit-exact bytes
                                           against the oracle's cvt.satfinite.rn(x / scale)
  test_reshape_and_cache_flash   :215-289  exact
  test_swap_blocks               :299-358  exact, cuda->cpu / cuda->cuda / cpu->cuda
  test_fp8_e4m3_conversion       :369-394  U(-224, 224) round trip, atol 1e-3 / rtol 0.1
plus what the reference kernels handle but its tests do not sweep: negative (padding) slots
(cache_kernels.cu:166-169, :217-220), strided key/value rows (a slice of a qkv tensor, :268-269),
kv_scale != 1, fp8_e5m2 caches, unaligned rows (scalar path), and > 2^31-element caches.
"""
import random

import pytest
import torch

from oracle import cache_ref as cr
from oracle import ref_cpu

pytestmark = pytest.mark.gpu
DEV = "cuda"
DTYPES = [torch.half, torch.bfloat16, torch.float]
HEAD_SIZES = [64, 80, 96, 112, 128, 192, 256]
BLOCK_SIZES = [8, 16, 32]
NUM_HEADS = 8
NUM_TOKENS = 42


def _bytes_equal(a: torch.Tensor, b: torch.Tensor) -> bool:
    return torch.equal(a.cpu().contiguous().view(torch.uint8), b.cpu().contiguous().view(torch.uint8))


def _caches(num_blocks, block_size, layers, head_size, kv, dtype, seed=0):
    return cr.make_kv_caches(num_blocks, block_size, layers, NUM_HEADS, head_size, kv, dtype, seed)


# ------------------------------------------------------------------------------------------
# copy_blocks (reference test :47-113: every source block copied to two destinations)
# ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("kv", ["auto", "fp8"])
@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("block_size", BLOCK_SIZES)
@pytest.mark.parametrize("head_size", HEAD_SIZES)
def test_copy_blocks(ops, head_size, block_size, dtype, kv):
    rnd = random.Random(0)
    num_blocks, num_mappings, layers = 160, 48, 2
    src = rnd.sample(range(num_blocks), num_mappings)
    rest = sorted(set(range(num_blocks)) - set(src))
    dst = rnd.sample(rest, 2 * num_mappings)
    mapping = [(s, dst[2 * i + j]) for i, s in enumerate(src) for j in range(2)]
    kcs, vcs = _caches(num_blocks, block_size, layers, head_size, kv, dtype)
    gk, gv = [k.to(DEV) for k in kcs], [v.to(DEV) for v in vcs]
    bm = torch.tensor(mapping, dtype=torch.int64, device=DEV).view(-1, 2)
    ops.copy_blocks(gk, gv, bm)
    cr.copy_blocks(kcs, vcs, torch.tensor(mapping, dtype=torch.int64))
    for a, b in zip(gk + gv, kcs + vcs):
        assert _bytes_equal(a, b)


def test_copy_blocks_many_layers_large_and_empty(ops):
    # Llama-3-8B-like: 32 layers, 8 kv heads x 128 x 16 fp16 blocks (32 KiB of K per block)
    rnd = random.Random(1)
    num_blocks, layers = 512, 32
    g = torch.Generator(device=DEV).manual_seed(0)
    gk = [torch.randn(num_blocks, 8, 16, 16, 8, device=DEV, dtype=torch.half, generator=g) for _ in range(layers)]
    gv = [torch.randn(num_blocks, 8, 128, 16, device=DEV, dtype=torch.half, generator=g) for _ in range(layers)]
    k0, v0 = [k.clone() for k in gk], [v.clone() for v in gv]
    src = rnd.sample(range(num_blocks), 100)
    rest = sorted(set(range(num_blocks)) - set(src))
    dst = rnd.sample(rest, 200)
    mapping = torch.tensor([(s, dst[2 * i + j]) for i, s in enumerate(src) for j in range(2)], dtype=torch.int64)
    ops.copy_blocks(gk, gv, mapping.to(DEV))
    s_idx, d_idx = mapping[:, 0].to(DEV), mapping[:, 1].to(DEV)
    for a, a0 in zip(gk + gv, k0 + v0):
        want = a0.clone()
        want[d_idx] = a0[s_idx]
        assert torch.equal(a, want)
    # empty mapping / empty layer list are no-ops (cache_kernels.cu:108-111)
    ops.copy_blocks(gk, gv, torch.empty(0, 2, dtype=torch.int64, device=DEV))
    ops.copy_blocks([], [], mapping.to(DEV))
    torch.cuda.synchronize()


@pytest.mark.parametrize("dtype", [torch.float, torch.bfloat16])
def test_copy_blocks_vs_compiled_reference_cpu(ops, dtype):
    if not ref_cpu.available():
        pytest.skip(str(ref_cpu.why_unavailable()))
    rnd = random.Random(2)
    kcs, vcs = _caches(96, 16, 3, 128, "auto", dtype)
    gk, gv = [k.to(DEV) for k in kcs], [v.to(DEV) for v in vcs]
    src = rnd.sample(range(96), 20)
    rest = sorted(set(range(96)) - set(src))
    dst = rnd.sample(rest, 40)
    mapping = torch.tensor([(s, dst[2 * i + j]) for i, s in enumerate(src) for j in range(2)], dtype=torch.int64)
    ref_cpu.copy_blocks(kcs, vcs, mapping)
    ops.copy_blocks(gk, gv, mapping.to(DEV))
    for a, b in zip(gk + gv, kcs + vcs):
        assert _bytes_equal(a, b)


# ------------------------------------------------------------------------------------------
# reshape_and_cache (reference test :124-204)
# ------------------------------------------------------------------------------------------
def _slots(num_blocks, block_size, n, seed=0, negatives=()):
    s = random.Random(seed).sample(range(num_blocks * block_size), n)
    for i in negatives:
        s[i] = -1 - i
    return torch.tensor(s, dtype=torch.int64)


@pytest.mark.parametrize("kv", ["auto", "fp8", "fp8_e5m2"])
@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("block_size", BLOCK_SIZES)
@pytest.mark.parametrize("head_size", HEAD_SIZES)
def test_reshape_and_cache(ops, head_size, block_size, dtype, kv):
    torch.manual_seed(0)
    num_blocks = 128
    slots = _slots(num_blocks, block_size, NUM_TOKENS, negatives=(3, 17))
    qkv = torch.randn(NUM_TOKENS, 3, NUM_HEADS, head_size).to(dtype)
    _, key, value = qkv.unbind(dim=1)  # rows strided by 3 * H * D: a slice of the qkv projection
    kcs, vcs = _caches(num_blocks, block_size, 1, head_size, kv, dtype)
    kc, vc = kcs[0], vcs[0]
    kv_scale = 1.0 if kv == "auto" else 0.75
    gq = qkv.to(DEV)
    _, gkey, gval = gq.unbind(dim=1)
    assert not gkey.is_contiguous()
    gkc, gvc = kc.to(DEV), vc.to(DEV)
    ops.reshape_and_cache(gkey, gval, gkc, gvc, slots.to(DEV), kv, kv_scale)
    cr.reshape_and_cache(key, value, kc, vc, slots, kv, kv_scale)
    # bit-exact: plain copies for "auto", cvt.satfinite.rn(float(x) / scale) bytes for fp8
    assert _bytes_equal(gkc, kc)
    assert _bytes_equal(gvc, vc)
    if kv == "fp8":
        # and the reference test's own criterion: dequantise through the product's convert_fp8,
        # compare with the unquantised scatter at atol 1e-3 / rtol 0.1 (test_cache.py:193-200)
        rk = torch.empty_like(gkc, dtype=torch.float16)
        rv = torch.empty_like(gvc, dtype=torch.float16)
        ops.convert_fp8(rk, gkc, kv_scale, "fp8")
        ops.convert_fp8(rv, gvc, kv_scale, "fp8")
        wk = cr.fp8_dequantize(kcs[0], torch.float16, kv_scale)  # untouched slots: old cache
        wv = cr.fp8_dequantize(vcs[0], torch.float16, kv_scale)
        for i, s in enumerate(slots.tolist()):
            if s < 0:
                continue
            b, o = divmod(s, block_size)
            wk[b, :, :, o, :] = key[i].reshape(NUM_HEADS, -1, wk.shape[-1]).to(torch.float16)
            wv[b, :, :, o] = value[i].to(torch.float16)
        assert torch.allclose(rk.cpu(), wk, atol=0.001, rtol=0.1)
        assert torch.allclose(rv.cpu(), wv, atol=0.001, rtol=0.1)


@pytest.mark.parametrize("dtype", [torch.float, torch.bfloat16])
@pytest.mark.parametrize("head_size", [64, 128, 256])
def test_reshape_and_cache_vs_compiled_reference_cpu(ops, head_size, dtype):
    if not ref_cpu.available():
        pytest.skip(str(ref_cpu.why_unavailable()))
    torch.manual_seed(1)
    num_blocks, bs = 64, 16
    slots = _slots(num_blocks, bs, NUM_TOKENS, seed=1, negatives=(0,))
    qkv = torch.randn(NUM_TOKENS, 3, NUM_HEADS, head_size).to(dtype)
    _, key, value = qkv.unbind(dim=1)
    kcs, vcs = _caches(num_blocks, bs, 1, head_size, "auto", dtype)
    gkc, gvc = kcs[0].to(DEV), vcs[0].to(DEV)
    _, gkey, gval = qkv.to(DEV).unbind(dim=1)
    ops.reshape_and_cache(gkey, gval, gkc, gvc, slots.to(DEV), "auto", 1.0)
    ref_cpu.reshape_and_cache(key, value, kcs[0], vcs[0], slots)
    assert _bytes_equal(gkc, kcs[0]) and _bytes_equal(gvc, vcs[0])


@pytest.mark.parametrize("kv", ["auto", "fp8"])
@pytest.mark.parametrize("dtype", [torch.half, torch.float])
def test_reshape_and_cache_unaligned_rows(ops, dtype, kv):
    """key/value rows that start 2 (fp16) / 4 (fp32) bytes off a 16-byte boundary take the scalar
    kernel; results must not change."""
    torch.manual_seed(2)
    head_size, bs, num_blocks, T = 128, 16, 32, 9
    flat = torch.randn(2, T * NUM_HEADS * head_size + 1).to(dtype)
    key = flat[0, 1:].view(T, NUM_HEADS, head_size)
    value = flat[1, 1:].view(T, NUM_HEADS, head_size)
    gflat = flat.to(DEV)
    gkey = gflat[0, 1:].view(T, NUM_HEADS, head_size)
    gval = gflat[1, 1:].view(T, NUM_HEADS, head_size)
    assert gkey.data_ptr() % 16 != 0
    kcs, vcs = _caches(num_blocks, bs, 1, head_size, kv, dtype)
    gkc, gvc = kcs[0].to(DEV), vcs[0].to(DEV)
    slots = _slots(num_blocks, bs, T, seed=2)
    ops.reshape_and_cache(gkey, gval, gkc, gvc, slots.to(DEV), kv, 1.0)
    cr.reshape_and_cache(key, value, kcs[0], vcs[0], slots, kv, 1.0)
    assert _bytes_equal(gkc, kcs[0]) and _bytes_equal(gvc, vcs[0])


def test_reshape_and_cache_decode_and_prefill_sizes(ops):
    """C2/C3 shapes: H=8, D=128, block 16; T = 1 (decode batch 1), 256, 2048 (prefill)."""
    g = torch.Generator(device=DEV).manual_seed(0)
    num_blocks, bs, H, D = 4096, 16, 8, 128
    for T in (1, 256, 2048):
        key = torch.randn(T, H, D, device=DEV, dtype=torch.half, generator=g)
        value = torch.randn(T, H, D, device=DEV, dtype=torch.half, generator=g)
        kc = torch.zeros(num_blocks, H, D // 8, bs, 8, device=DEV, dtype=torch.half)
        vc = torch.zeros(num_blocks, H, D, bs, device=DEV, dtype=torch.half)
        slots = torch.randperm(num_blocks * bs, device=DEV, generator=g)[:T].to(torch.int64)
        ops.reshape_and_cache(key, value, kc, vc, slots, "auto", 1.0)
        b, o = slots // bs, slots % bs
        assert torch.equal(kc[b, :, :, o, :], key.view(T, H, D // 8, 8))
        assert torch.equal(vc[b, :, :, o], value)
        assert int((kc != 0).sum()) == int((key != 0).sum())  # nothing else was written


def test_reshape_and_cache_beyond_int32_elements(ops):
    """A cache with more than 2^31 elements: block offsets must be computed in 64 bits."""
    H, D, bs = 8, 128, 16
    num_blocks = (1 << 31) // (H * D * bs) + 64  # 131,136 blocks = 4.0 GiB of fp16 K
    kc = torch.zeros(num_blocks, H, D // 8, bs, 8, device=DEV, dtype=torch.half)
    vc = torch.zeros(num_blocks, H, D, bs, device=DEV, dtype=torch.half)
    T = 8
    key = torch.randn(T, H, D, device=DEV, dtype=torch.half)
    value = torch.randn(T, H, D, device=DEV, dtype=torch.half)
    slots = torch.tensor([(num_blocks - 1 - i) * bs + i for i in range(T)], dtype=torch.int64, device=DEV)
    ops.reshape_and_cache(key, value, kc, vc, slots, "auto", 1.0)
    b, o = slots // bs, slots % bs
    assert torch.equal(kc[b, :, :, o, :], key.view(T, H, D // 8, 8))
    assert torch.equal(vc[b, :, :, o], value)
    assert not kc[: num_blocks - T].any() and not vc[: num_blocks - T].any()
    del kc, vc
    torch.cuda.empty_cache()


# ------------------------------------------------------------------------------------------
# reshape_and_cache_flash (reference test :215-289; layout [NB, bs, H, D])
# ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("block_size", BLOCK_SIZES)
@pytest.mark.parametrize("head_size", HEAD_SIZES)
def test_reshape_and_cache_flash(ops, head_size, block_size, dtype):
    torch.manual_seed(0)
    num_blocks = 128
    slots = _slots(num_blocks, block_size, NUM_TOKENS, negatives=(5,))
    qkv = torch.randn(NUM_TOKENS, 3, NUM_HEADS, head_size).to(dtype)
    _, key, value = qkv.unbind(dim=1)
    s = head_size ** -0.5
    kc = ((torch.rand(num_blocks, block_size, NUM_HEADS, head_size) * 2 - 1) * s).to(dtype)
    vc = ((torch.rand(num_blocks, block_size, NUM_HEADS, head_size) * 2 - 1) * s).to(dtype)
    gkc, gvc = kc.to(DEV), vc.to(DEV)
    _, gkey, gval = qkv.to(DEV).unbind(dim=1)
    ops.reshape_and_cache_flash(gkey, gval, gkc, gvc, slots.to(DEV), "auto")
    cr.reshape_and_cache_flash(key, value, kc, vc, slots)
    assert _bytes_equal(gkc, kc) and _bytes_equal(gvc, vc)


def test_reshape_and_cache_flash_rejects_fp8(ops):
    # cache_kernels.cu:288-290: "Unsupported data type of kv cache"
    k = torch.zeros(2, 8, 64, device=DEV, dtype=torch.half)
    c = torch.zeros(4, 16, 8, 64, device=DEV, dtype=torch.half)
    with pytest.raises(RuntimeError):
        ops.reshape_and_cache_flash(k, k, c, c.clone(), torch.zeros(2, dtype=torch.int64, device=DEV), "fp8")


# ------------------------------------------------------------------------------------------
# swap_blocks (reference test :299-358)
# ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("kv", ["auto", "fp8"])
@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("head_size", [64, 80, 128, 256])
@pytest.mark.parametrize("block_size", BLOCK_SIZES)
@pytest.mark.parametrize("direction", [("cuda", "cpu"), ("cuda", "cuda"), ("cpu", "cuda")])
def test_swap_blocks(ops, direction, block_size, head_size, dtype, kv):
    if kv == "fp8" and "cpu" in direction:
        pytest.skip("as the reference test (:324-325)")
    rnd = random.Random(0)
    num_blocks, num_mappings = 128, 48
    src_blocks = rnd.sample(range(num_blocks), num_mappings)
    if direction[0] == direction[1]:
        rest = sorted(set(range(num_blocks)) - set(src_blocks))
        dst_blocks = rnd.sample(rest, num_mappings)
    else:
        dst_blocks = rnd.sample(range(num_blocks), num_mappings)
    mapping = torch.tensor(list(zip(src_blocks, dst_blocks)), dtype=torch.int64).view(-1, 2)
    sk, sv = _caches(num_blocks, block_size, 1, head_size, kv, dtype, seed=0)
    dk, dv = _caches(num_blocks, block_size, 1, head_size, kv, dtype, seed=1)
    pin = lambda t: t.pin_memory()
    src_k = sk[0].to(DEV) if direction[0] == "cuda" else pin(sk[0])
    src_v = sv[0].to(DEV) if direction[0] == "cuda" else pin(sv[0])
    dst_k = dk[0].to(DEV) if direction[1] == "cuda" else pin(dk[0])
    dst_v = dv[0].to(DEV) if direction[1] == "cuda" else pin(dv[0])
    ops.swap_blocks(src_k, dst_k, mapping)
    ops.swap_blocks(src_v, dst_v, mapping)
    torch.cuda.synchronize()  # the op is asynchronous (cache_kernels.cu:56 NOTE)
    cr.swap_blocks(sk[0], dk[0], mapping)
    cr.swap_blocks(sv[0], dv[0], mapping)
    assert _bytes_equal(dst_k, dk[0]) and _bytes_equal(dst_v, dv[0])  # incl. untouched blocks
    assert _bytes_equal(src_k, sk[0])


def test_swap_blocks_rejects_device_mapping_and_cpu_cpu(ops):
    a = torch.zeros(4, 64, device=DEV)
    with pytest.raises(RuntimeError):  # cache_kernels.cu:44
        ops.swap_blocks(a, a.clone(), torch.zeros(1, 2, dtype=torch.int64, device=DEV))
    with pytest.raises((RuntimeError, NotImplementedError)):  # :30-42 "Invalid device combination"
        ops.swap_blocks(torch.zeros(4, 64), torch.zeros(4, 64), torch.zeros(1, 2, dtype=torch.int64))


# ------------------------------------------------------------------------------------------
# convert_fp8 (reference test :369-394)
# ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("block_size", BLOCK_SIZES)
@pytest.mark.parametrize("head_size", HEAD_SIZES)
def test_fp8_e4m3_conversion(ops, head_size, block_size, dtype):
    g = torch.Generator().manual_seed(0)
    shape = (256, NUM_HEADS, head_size, block_size)
    cache = ((torch.rand(shape, generator=g) * 2 - 1) * 224.0).to(dtype)
    gc = cache.to(DEV)
    g8 = torch.empty_like(gc, dtype=torch.uint8)
    ops.convert_fp8(g8, gc)
    back = torch.empty_like(gc)
    ops.convert_fp8(back, g8)
    assert torch.allclose(gc, back, atol=0.001, rtol=0.1)  # the reference's criterion
    # bit-exact in both directions against the oracle
    assert torch.equal(g8.cpu(), cr.fp8_quantize(cache, 1.0, "fp8"))
    assert _bytes_equal(back, cr.fp8_dequantize(g8.cpu(), dtype, 1.0, "fp8"))


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("scale", [1.0, 0.5, 3.7])
def test_convert_fp8_scaled_specials_and_every_code(ops, dtype, scale):
    codes = torch.arange(256, dtype=torch.uint8).repeat(64)
    out = torch.empty(codes.numel(), dtype=dtype, device=DEV)
    ops.convert_fp8(out, codes.to(DEV), scale, "fp8_e4m3")
    want = cr.fp8_dequantize(codes, dtype, scale, "fp8")
    got = out.cpu()
    same = (got == want) | (torch.isnan(got) & torch.isnan(want))
    assert bool(same.all())
    # towards fp8: saturation (no inf / NaN from finite inputs), zero, denormals, huge values
    x = torch.tensor([0.0, -0.0, 1e-9, 2.0 ** -9, 2.0 ** -10, 447.0, 448.0, 449.0, 464.0, 465.0, 1e4,
                      -1e4, 6e4, -6e4, 0.017, -0.33, 239.9, 240.1], dtype=torch.float).to(dtype)
    q = torch.empty(x.numel(), dtype=torch.uint8, device=DEV)
    ops.convert_fp8(q, x.to(DEV), scale, "fp8")
    assert torch.equal(q.cpu(), cr.fp8_quantize(x, scale, "fp8"))


def test_convert_fp8_rejects_what_the_reference_rejects(ops):
    a = torch.zeros(64, device=DEV, dtype=torch.half)
    b = torch.zeros(64, device=DEV, dtype=torch.uint8)
    for kv in ("fp8_e5m2", "auto", "int8"):  # cache_kernels.cu:386-388; quant_utils.cuh:513-523
        with pytest.raises(RuntimeError):
            ops.convert_fp8(b, a, 1.0, kv)
    with pytest.raises(RuntimeError):  # cache_kernels.cu:343-347
        ops.convert_fp8(b, a.cpu(), 1.0, "fp8")
