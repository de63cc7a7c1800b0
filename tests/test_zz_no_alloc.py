"""b200_paged_attention_v2_static: the no-allocation variant of paged_attention_v2 (scratch in a
device array built into the library).  Called through the C-ABI and checked against the existing
entry point b200_paged_attention_v2 on the same inputs -- same arithmetic, so bit-identical -- and
against the oracle at the tolerance of the other attention tests (rtol 1e-2 / atol 1e-3), on two
calls in a row that share the array.  This file sorts last on
purpose."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda"


class Params(ctypes.Structure):  # b200_paged_attention_params (include/b200_vllm.h)
    _fields_ = [
        ("out", ctypes.c_void_p), ("query", ctypes.c_void_p), ("key_cache", ctypes.c_void_p),
        ("value_cache", ctypes.c_void_p), ("block_tables", ctypes.c_void_p), ("seq_lens", ctypes.c_void_p),
        ("alibi_slopes", ctypes.c_void_p), ("exp_sums", ctypes.c_void_p), ("max_logits", ctypes.c_void_p),
        ("tmp_out", ctypes.c_void_p),
        ("num_seqs", ctypes.c_int32), ("num_heads", ctypes.c_int32), ("num_kv_heads", ctypes.c_int32),
        ("head_size", ctypes.c_int32), ("block_size", ctypes.c_int32),
        ("max_num_blocks_per_seq", ctypes.c_int32), ("max_seq_len", ctypes.c_int32),
        ("q_stride", ctypes.c_int64), ("kv_block_stride", ctypes.c_int64), ("kv_head_stride", ctypes.c_int64),
        ("scale", ctypes.c_float), ("kv_scale", ctypes.c_float),
        ("dtype", ctypes.c_int32), ("kv_dtype", ctypes.c_int32),
        ("tp_rank", ctypes.c_int32), ("blocksparse_local_blocks", ctypes.c_int32),
        ("blocksparse_vert_stride", ctypes.c_int32), ("blocksparse_block_size", ctypes.c_int32),
        ("blocksparse_head_sliding_step", ctypes.c_int32),
    ]


@pytest.fixture(scope="module")
def lib():
    import neural_magic_vllm_b200 as pkg
    lib = pkg.c_abi()
    for name in ("b200_paged_attention_v2", "b200_paged_attention_v2_static"):
        fn = getattr(lib, name)
        fn.argtypes = [ctypes.POINTER(Params), ctypes.c_void_p]
        fn.restype = ctypes.c_int
    lib.b200_paged_attention_v2_static_scratch_max_bytes.restype = ctypes.c_int64
    lib.b200_paged_attention_v2_static_scratch_bytes.restype = ctypes.c_int64
    lib.b200_paged_attention_v2_static_scratch_bytes.argtypes = [ctypes.c_int32] * 5
    lib.b200_last_error.restype = ctypes.c_char_p
    return lib


def _case(lib, num_seqs, heads, kv_heads, seq_lens, seed):
    from oracle import attention_ref as ar
    d, bs = 128, 16
    q, kc, vc, bt, sl = ar.make_paged_inputs(num_seqs, heads, kv_heads, d, bs, seq_lens, torch.float16, seed=seed)
    ref = ar.paged_attention_ref(q, kc, vc, bt, sl, d ** -0.5, None, "auto", 1.0)
    qd, kcd, vcd, btd, sld = (t.to(DEV) for t in (q, kc, vc, bt, sl))
    max_len = int(sl.max())
    parts = (max_len + 511) // 512
    # all device memory is set up here, before either entry point runs
    out_a, out_b = torch.full_like(qd, float("nan")), torch.full_like(qd, float("nan"))
    tmp = torch.empty(num_seqs, heads, parts, d, dtype=torch.float16, device=DEV)
    es = torch.empty(num_seqs, heads, parts, dtype=torch.float32, device=DEV)
    ml = torch.empty_like(es)
    p = Params()
    p.query, p.key_cache, p.value_cache = qd.data_ptr(), kcd.data_ptr(), vcd.data_ptr()
    p.block_tables, p.seq_lens, p.alibi_slopes = btd.data_ptr(), sld.data_ptr(), None
    p.num_seqs, p.num_heads, p.num_kv_heads, p.head_size, p.block_size = num_seqs, heads, kv_heads, d, bs
    p.max_num_blocks_per_seq, p.max_seq_len = bt.shape[1], max_len
    p.q_stride, p.kv_block_stride, p.kv_head_stride = qd.stride(0), kcd.stride(0), kcd.stride(1)
    p.scale, p.kv_scale, p.dtype, p.kv_dtype = d ** -0.5, 1.0, 0, 0
    p.tp_rank, p.blocksparse_vert_stride = 0, 0
    stream = torch.cuda.current_stream().cuda_stream
    # existing entry point, caller-owned scratch
    p.out, p.exp_sums, p.max_logits, p.tmp_out = out_a.data_ptr(), es.data_ptr(), ml.data_ptr(), tmp.data_ptr()
    assert lib.b200_paged_attention_v2(ctypes.byref(p), stream) == 0, lib.b200_last_error()
    # new entry point: no scratch passed
    p.out, p.exp_sums, p.max_logits, p.tmp_out = out_b.data_ptr(), None, None, None
    before = torch.cuda.memory_allocated()
    assert lib.b200_paged_attention_v2_static(ctypes.byref(p), stream) == 0, lib.b200_last_error()
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before
    assert torch.equal(out_a, out_b), "static-scratch variant differs from b200_paged_attention_v2"
    torch.testing.assert_close(out_b.cpu().float(), ref.float(), rtol=1e-2, atol=1e-3)


def test_static_scratch_matches_v2_on_two_calls_in_a_row(lib):
    # the second call is smaller and has other partition counts: the array still holds the first
    # call's contents, which must not leak into the result
    _case(lib, 6, 32, 8, [1, 100, 513, 1024, 2000, 777], seed=1)
    _case(lib, 3, 16, 16, [1500, 5, 640], seed=2)


def test_static_scratch_rejects_oversize(lib):
    cap = lib.b200_paged_attention_v2_static_scratch_max_bytes()
    assert cap == 16 << 20
    assert lib.b200_paged_attention_v2_static_scratch_bytes(6, 32, 128, 2000, 0) <= cap
    assert lib.b200_paged_attention_v2_static_scratch_bytes(256, 64, 128, 32768, 0) > cap
    p = Params()
    p.num_seqs, p.num_heads, p.num_kv_heads, p.head_size, p.block_size = 256, 64, 8, 128, 16
    p.max_seq_len, p.max_num_blocks_per_seq, p.dtype = 32768, 2048, 0
    rc = lib.b200_paged_attention_v2_static(ctypes.byref(p), None)  # rejected before any pointer is used
    assert rc != 0
    assert b"built-in array holds" in lib.b200_last_error()
