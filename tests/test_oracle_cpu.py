"""CPU suite: pins oracle/ against the reference-generated golden vectors (tests/golden, made by
tests/golden/make_golden.py from the reference's own Python) and against the compiled reference
CPU kernels (oracle/_ref), and checks that the C-ABI library exports what include/*.h declares.
No GPU compute here."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch

from oracle import attention_ref as ar
from oracle import quant_ref as qr
from oracle import ref_cpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")


def _cases(npz, prefix):
    keys = sorted({re.match(rf"({prefix}_[a-z0-9_\-]+?)_(w|w_ref|marlin_q|marlin_s|g_idx|sort_idx|"
                            r"rand_perm|gptq|expect|meta)$", k).group(1) for k in npz.files
                   if k.startswith(prefix)})
    return keys


DENSE = np.load(os.path.join(GOLD, "marlin_dense.npz"))
S24 = np.load(os.path.join(GOLD, "marlin_24.npz"))


def _parse(key):
    m = re.search(r"k(\d+)_n(\d+)_b(\d+)(?:_g(-?\d+))?(?:_a(\d))?", key)
    k, n, b = int(m.group(1)), int(m.group(2)), int(m.group(3))
    g = int(m.group(4)) if m.group(4) else None
    a = bool(int(m.group(5))) if m.group(5) else False
    return k, n, b, g, a


@pytest.mark.parametrize("key", _cases(DENSE, "dense"))
def test_marlin_quantize_matches_reference(key):
    k, n, bits, group, act = _parse(key)
    w = torch.from_numpy(DENSE[key + "_w"])
    rand_perm = torch.from_numpy(DENSE[key + "_rand_perm"])
    if act:
        # same permutation as the reference drew: replay it instead of torch.randperm
        w_ref, q, s, _, _ = qr.quantize_weights(w, bits, group, False)
        g_idx = (torch.arange(k) // group).to(torch.int32)[rand_perm]
        q, w_ref = q[rand_perm], w_ref[rand_perm]
        q, g_idx, sort_idx = qr.sort_weights(q, g_idx)
        perm, sp, sps = qr.get_perms(bits)
        mq = qr.marlin_weights(q, k, n, bits, perm)
        ms = qr.marlin_permute_scales(s, k, n, group, sp, sps)
        assert torch.equal(g_idx, torch.from_numpy(DENSE[key + "_g_idx"]))
        assert torch.equal(sort_idx, torch.from_numpy(DENSE[key + "_sort_idx"]))
    else:
        w_ref, mq, ms, _, _, _ = qr.marlin_quantize(w, bits, group, False)
    assert torch.equal(mq, torch.from_numpy(DENSE[key + "_marlin_q"]))  # bit-exact layout
    assert torch.equal(ms, torch.from_numpy(DENSE[key + "_marlin_s"]))
    assert torch.equal(w_ref, torch.from_numpy(DENSE[key + "_w_ref"]))
    # the inverse used by the GEMM oracle: unpack(marlin) == integer weights
    if not act:
        q_back = qr.marlin_unpack(mq, k, n, bits)
        gs = k if group == -1 else group
        s_back = qr.marlin_unpermute_scales(ms, k, n, gs, bits)
        zp = 1 << (bits - 1)
        w_back = (q_back - zp).half() * s_back.repeat_interleave(gs, 0)
        assert torch.equal(w_back, w_ref)


@pytest.mark.parametrize("key", _cases(DENSE, "repack"))
def test_repack_oracle_matches_reference(key):
    k, n, bits, _, act = _parse(key)
    gptq = torch.from_numpy(DENSE[key + "_gptq"])
    sort_idx = torch.from_numpy(DENSE[key + "_sort_idx"])
    q = qr.gptq_unpack(gptq, bits)
    assert torch.equal(qr.gptq_pack(q, bits, k, n), gptq)
    if act:
        q = q[sort_idx.long()]
    perm, _, _ = qr.get_perms(bits)
    assert torch.equal(qr.marlin_weights(q, k, n, bits, perm), torch.from_numpy(DENSE[key + "_expect"]))


@pytest.mark.parametrize("key", _cases(S24, "s24"))
def test_marlin_24_quantize_matches_reference(key):
    k, n, bits, group, _ = _parse(key)
    w = torch.from_numpy(S24[key + "_w"])
    w_ref, mq, meta, ms = qr.marlin_24_quantize(w, bits, group)
    assert torch.equal(w_ref, torch.from_numpy(S24[key + "_w_ref"]))
    assert torch.equal(mq, torch.from_numpy(S24[key + "_marlin_q"]))
    assert torch.equal(meta, torch.from_numpy(S24[key + "_meta"]))
    assert torch.equal(ms, torch.from_numpy(S24[key + "_marlin_s"]))


def test_compress_24_roundtrip():
    torch.manual_seed(0)
    w = torch.randn(64, 128)
    w24, mask = qr.inject_24(w.t().contiguous().t())
    assert int(mask.t().reshape(-1, 4).sum(1).max()) == 2
    q = torch.randint(-7, 8, (64, 128), dtype=torch.int32)
    q = q * qr.mask_24(q.float()).int()
    comp, meta = qr.compress_24(q)
    assert torch.equal(qr.decompress_24(comp, meta), q)


def test_baseline_scaled_mm_and_quant_refs():
    torch.manual_seed(0)
    x = torch.randn(17, 64, dtype=torch.half) * 3
    q, s = qr.scaled_int8_quant_ref(x)
    assert q.dtype == torch.int8 and s.shape == (17, 1)
    assert torch.equal(s, x.float().abs().max(-1, keepdim=True)[0] / 127.0)
    q8, s8 = qr.scaled_fp8_quant_ref(x)
    assert torch.allclose(q8.float() * s8, x.float(), rtol=0.07, atol=1e-2)
    a = torch.randint(-5, 5, (4, 32), dtype=torch.int8)
    b = torch.randint(-5, 5, (32, 8), dtype=torch.int8)
    out = qr.baseline_scaled_mm(a, b, torch.ones(4, 1), torch.ones(1, 8), torch.float32)
    assert torch.equal(out, (a.int() @ b.int()).float())


# ---------------------------------------------------------------------------------------------
# attention / cache oracle vs the reference's own compiled CPU kernels (csrc/cpu/*.cpp)
# ---------------------------------------------------------------------------------------------
needs_ref = pytest.mark.skipif(not ref_cpu.available(), reason=str(ref_cpu.why_unavailable()))


@needs_ref
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("heads", [(8, 8), (16, 4)])
@pytest.mark.parametrize("version", ["v1", "v2"])
def test_attention_oracle_vs_reference_cpu(dtype, heads, version):
    nh, nkv = heads
    D = 64
    seq_lens = [1, 17, 300, 700]
    q, kc, vc, bt, sl = ar.make_paged_inputs(len(seq_lens), nh, nkv, D, 16, seq_lens, dtype, seed=3)
    scale = D ** -0.5
    ours = ar.paged_attention_ref(q, kc, vc, bt, sl, scale, None, "auto", 1.0)
    out = torch.empty_like(q)
    mx = max(seq_lens)
    if version == "v1":
        ref_cpu.paged_attention_v1(out, q, kc, vc, nkv, scale, bt, sl, 16, mx)
    else:
        parts = (mx + 511) // 512
        es = torch.empty(len(seq_lens), nh, parts)
        ml = torch.empty_like(es)
        tmp = torch.empty(len(seq_lens), nh, parts, D, dtype=dtype)
        ref_cpu.paged_attention_v2(out, es, ml, tmp, q, kc, vc, nkv, scale, bt, sl, 16, mx)
    tol = dict(atol=1e-5, rtol=1e-5) if dtype == torch.float32 else dict(atol=2e-3, rtol=2e-2)
    torch.testing.assert_close(ours.float(), out.float(), **tol)


@needs_ref
def test_reshape_and_cache_oracle_vs_reference_cpu():
    torch.manual_seed(0)
    T, H, D, NB, bs, x = 9, 4, 64, 8, 16, 4  # fp32: x = 16 / 4
    key = torch.randn(T, H, D)
    value = torch.randn(T, H, D)
    kc = torch.zeros(NB, H, D // x, bs, x)
    vc = torch.zeros(NB, H, D, bs)
    slots = torch.randperm(NB * bs)[:T].to(torch.int64)
    ref_cpu.reshape_and_cache(key, value, kc, vc, slots)
    kc2, vc2 = torch.zeros_like(kc), torch.zeros_like(vc)
    for t, s in enumerate(slots.tolist()):
        b, o = divmod(s, bs)
        kc2[b, :, :, o, :] = key[t].reshape(H, D // x, x)
        vc2[b, :, :, o] = value[t]
    assert torch.equal(kc, kc2) and torch.equal(vc, vc2)


# ---------------------------------------------------------------------------------------------
# C ABI: every function include/b200_vllm.h declares is exported by libb200vllm.so
# ---------------------------------------------------------------------------------------------
def test_c_abi_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "b200_vllm.h")).read()
    names = re.findall(r"B200_API\s+[\w\s\*]+?\b(b200_\w+)\s*\(", hdr)
    assert len(names) >= 15
    lib_path = os.path.join(ROOT, "neural_magic_vllm_b200", "libb200vllm.so")
    assert os.path.exists(lib_path), "libb200vllm.so not built: run python __graft_entry__.py"
    lib = ctypes.CDLL(lib_path)
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, f"declared but not exported: {missing}"
    lib.b200_abi_version.restype = ctypes.c_int
    assert lib.b200_abi_version() >= 2
    lib.b200_parse_kv_cache_dtype.argtypes = [ctypes.c_char_p]
    assert lib.b200_parse_kv_cache_dtype(b"fp8") == 1
    assert lib.b200_parse_kv_cache_dtype(b"auto") == 0
    assert lib.b200_parse_kv_cache_dtype(b"int3") < 0
    assert lib.b200_scaled_mm_supports_fp8(100) == 1 and lib.b200_scaled_mm_supports_fp8(80) == 0


def test_torch_ops_registered_with_reference_schemas():
    import neural_magic_vllm_b200  # noqa: F401  loads _C.so (fails loudly when missing)
    want = ["_C::paged_attention_v1", "_C::paged_attention_v2", "_C::gptq_marlin_gemm",
            "_C::gptq_marlin_repack", "_C::gptq_marlin_24_gemm", "_C::fp8_marlin_gemm", "_C::cutlass_scaled_mm",
            "_C::cutlass_scaled_mm_supports_fp8", "_C::static_scaled_fp8_quant",
            "_C::dynamic_scaled_fp8_quant", "_C::static_scaled_int8_quant",
            "_C::dynamic_scaled_int8_quant", "_C_cache_ops::reshape_and_cache",
            "_C_cache_ops::reshape_and_cache_flash", "_C_cache_ops::copy_blocks",
            "_C_cache_ops::swap_blocks", "_C_cache_ops::convert_fp8",
            "_C_cuda_utils::get_device_attribute"]
    for w in want:
        ns, name = w.split("::")
        assert hasattr(getattr(torch.ops, ns), name), w
    s = str(torch.ops._C.cutlass_scaled_mm.default._schema)
    assert "Tensor(a0!) out" in s or "Tensor! out" in s or "out" in s


def test_awq_pack_roundtrip_and_dequantize_ref():
    """AWQ restatement (oracle/quant_ref.py): nibble j of a word holds channel 8c + {0,2,4,6,1,3,5,7}[j]
    (csrc/quantization/awq/dequantize.cuh:17-101), w = (q - z) * s rounded to fp16 once."""
    import torch
    from oracle import quant_ref as qr
    g = torch.Generator().manual_seed(0)
    q = torch.randint(0, 16, (64, 32), generator=g)
    packed = qr.awq_pack(q)
    assert packed.dtype == torch.int32 and tuple(packed.shape) == (64, 4)
    assert torch.equal(qr.awq_unpack(packed), q)
    # the reference's lop3 pairs nibbles (0,4) (1,5) (2,6) (3,7) into consecutive channels
    word = int(packed[0, 0]) & 0xFFFFFFFF
    pairs = [((word >> (4 * j)) & 0xF, (word >> (4 * (j + 4))) & 0xF) for j in range(4)]
    assert [v for p in pairs for v in p] == q[0, :8].tolist()
    w = torch.randn(128, 64, generator=g) * 0.1
    w_ref, qw, s, qz = qr.awq_quantize(w, 32)
    assert w_ref.dtype == torch.half and tuple(qz.shape) == (4, 8) and tuple(s.shape) == (4, 64)
    assert float((w_ref.float() - w).abs().max()) < float(s.float().max()) * 0.51 + 1e-3
    manual = ((qr.awq_unpack(qw) - qr.awq_unpack(qz).repeat_interleave(32, 0)).float()
              * s.float().repeat_interleave(32, 0)).half()
    assert torch.equal(manual, w_ref)
