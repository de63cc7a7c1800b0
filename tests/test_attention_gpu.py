"""GPU parity of paged_attention_v1/v2 against the oracle (oracle/attention_ref.py, itself
pinned to the compiled reference CPU kernels in tests/test_oracle.py).

Mirrors the sweep of the reference's tests/kernels/test_attention.py:118-285 (heads
(40,40)/(64,8), head sizes 64..256, block sizes, dtypes, ALiBi, fp8 KV) and adds the
B200-specific edge cases: partition boundaries (511/512/513), single-token and empty
sequences, block_size 8, non-contiguous query rows, GQA group 4 (Llama-3).
Tolerances: BASELINE.json north_star -- rtol 1e-2 / atol 1e-3 for fp16/bf16; fp8 KV
atol 1e-2 (the reference's own bound, test_attention.py:279-283).
"""
import pytest
import torch

from oracle import attention_ref as ar

pytestmark = pytest.mark.gpu

DTYPES = {"half": torch.float16, "bf16": torch.bfloat16, "float": torch.float32}


def run_case(ops, version, dtype, num_heads, num_kv_heads, head_size, block_size, seq_lens,
             kv_cache_dtype="auto", use_alibi=False, kv_scale=1.0, seed=0, q_pad=0,
             max_seq_len=None):
    q, kc, vc, bt, sl = ar.make_paged_inputs(len(seq_lens), num_heads, num_kv_heads, head_size,
                                             block_size, seq_lens, dtype, seed=seed,
                                             kv_cache_dtype=kv_cache_dtype)
    scale = head_size ** -0.5
    alibi = None
    if use_alibi:
        alibi = torch.randn(num_heads, generator=torch.Generator().manual_seed(seed + 1))
    ref = ar.paged_attention_ref(q, kc, vc, bt, sl, scale, alibi, kv_cache_dtype, kv_scale)

    dev = "cuda"
    if q_pad:  # q as a slice of a wider qkv row (attention_kernels.cu:178 honours q_stride)
        wide = torch.zeros(q.shape[0], num_heads * head_size + q_pad, dtype=dtype, device=dev)
        wide[:, :num_heads * head_size] = q.to(dev).view(q.shape[0], -1)
        qd = wide[:, :num_heads * head_size].view(q.shape[0], num_heads, head_size)
        assert not qd.is_contiguous() or q.shape[0] == 1
    else:
        qd = q.to(dev)
    kcd, vcd, btd, sld = kc.to(dev), vc.to(dev), bt.to(dev), sl.to(dev)
    alibid = alibi.to(dev) if alibi is not None else None
    out = torch.full_like(q, float("nan"), device=dev)
    max_len = max_seq_len or max(int(sl.max()), 1)
    if version == "v1":
        ops.paged_attention_v1(out, qd, kcd, vcd, num_kv_heads, scale, btd, sld, block_size,
                               max_len, alibid, kv_cache_dtype, kv_scale)
    else:
        parts = (max_len + 511) // 512
        tmp = torch.empty(len(seq_lens), num_heads, parts, head_size, dtype=dtype, device=dev)
        es = torch.empty(len(seq_lens), num_heads, parts, dtype=torch.float32, device=dev)
        ml = torch.empty_like(es)
        ops.paged_attention_v2(out, es, ml, tmp, qd, kcd, vcd, num_kv_heads, scale, btd, sld,
                               block_size, max_len, alibid, kv_cache_dtype, kv_scale)
    torch.cuda.synchronize()
    out = out.cpu()
    nz = [i for i, n in enumerate(seq_lens) if n > 0]
    atol = 1e-2 if kv_cache_dtype != "auto" else 1e-3
    torch.testing.assert_close(out[nz].float(), ref[nz].float(), atol=atol, rtol=1e-2)
    return out, ref


SEQ_LENS = [1, 15, 16, 17, 511, 512, 513, 1000]


@pytest.mark.parametrize("version", ["v1", "v2"])
@pytest.mark.parametrize("dtype", ["half", "bf16", "float"])
@pytest.mark.parametrize("heads", [(40, 40), (64, 8), (32, 8)])
@pytest.mark.parametrize("use_alibi", [False, True])
def test_paged_attention_heads_dtypes(ops, version, dtype, heads, use_alibi):
    run_case(ops, version, DTYPES[dtype], heads[0], heads[1], 128, 16, SEQ_LENS,
             use_alibi=use_alibi)


@pytest.mark.parametrize("version", ["v1", "v2"])
@pytest.mark.parametrize("head_size", [64, 80, 96, 112, 128, 192, 256])
@pytest.mark.parametrize("block_size", [8, 16, 32])
@pytest.mark.parametrize("dtype", ["half", "bf16"])
def test_paged_attention_head_and_block_sizes(ops, version, head_size, block_size, dtype):
    run_case(ops, version, DTYPES[dtype], 16, 4, head_size, block_size,
             [1, 7, 9, 33, 500, 529, 777], seed=head_size + block_size)


@pytest.mark.parametrize("version", ["v1", "v2"])
@pytest.mark.parametrize("kv", ["fp8", "fp8_e5m2"])
@pytest.mark.parametrize("dtype", ["half", "bf16"])
@pytest.mark.parametrize("head_size,block_size,heads",
                         [(128, 16, (32, 8)), (64, 16, (40, 40)), (80, 32, (64, 8)),
                          (256, 8, (16, 4)), (128, 8, (8, 8))])
def test_paged_attention_fp8_kv(ops, version, kv, dtype, head_size, block_size, heads):
    run_case(ops, version, DTYPES[dtype], heads[0], heads[1], head_size, block_size,
             [3, 16, 100, 513, 900], kv_cache_dtype=kv, seed=7)


def test_fp8_kv_scale_folding(ops):
    # kv_scale != 1 is folded into the logits / output instead of a per-element multiply
    run_case(ops, "v2", torch.float16, 32, 8, 128, 16, [64, 700], kv_cache_dtype="fp8",
             kv_scale=0.5, seed=3)
    run_case(ops, "v1", torch.bfloat16, 32, 8, 128, 16, [64, 700], kv_cache_dtype="fp8",
             kv_scale=2.0, seed=4)


def test_float_with_fp8_cache(ops):
    run_case(ops, "v2", torch.float32, 8, 2, 64, 16, [5, 600], kv_cache_dtype="fp8", seed=5)


@pytest.mark.parametrize("version", ["v1", "v2"])
def test_empty_and_ragged_batch(ops, version):
    out, _ = run_case(ops, version, torch.float16, 32, 8, 128, 16, [0, 1, 0, 4096, 2], seed=11)
    assert torch.isfinite(out[[1, 3, 4]].float()).all()


@pytest.mark.parametrize("version", ["v1", "v2"])
def test_query_row_stride(ops, version):
    run_case(ops, version, torch.float16, 32, 8, 128, 16, [33, 1025], q_pad=2 * 8 * 128, seed=12)


def test_v2_scratch_sized_for_longer_max_seq_len(ops):
    # callers size scratch by max_seq_len of the batch bucket, not of these sequences
    run_case(ops, "v2", torch.float16, 32, 8, 128, 16, [10, 300], max_seq_len=4096, seed=13)


def test_large_group_mqa(ops):
    # 64 query heads on 1 kv head: several 16-head tiles per kv head
    run_case(ops, "v2", torch.float16, 64, 1, 128, 16, [100, 1030], seed=14)
    run_case(ops, "v1", torch.bfloat16, 24, 2, 64, 16, [100, 1030], seed=15)


def test_tail_garbage_does_not_leak(ops):
    """Slots past seq_len inside the last block may hold NaN/Inf (uninitialised cache):
    they must not reach the output (attention_kernels.cu:420-430)."""
    dtype = torch.float16
    q, kc, vc, bt, sl = ar.make_paged_inputs(2, 32, 8, 128, 16, [5, 530], dtype, seed=21)
    ref = ar.paged_attention_ref(q, kc, vc, bt, sl, 128 ** -0.5)
    for i, n in enumerate([5, 530]):
        blk = int(bt[i, n // 16])
        kc[blk, :, :, n % 16:, :] = float("nan")
        vc[blk, :, :, n % 16:] = float("inf")
    out = torch.empty_like(q, device="cuda")
    for version in ("v1", "v2"):
        if version == "v1":
            ops.paged_attention_v1(out, q.cuda(), kc.cuda(), vc.cuda(), 8, 128 ** -0.5, bt.cuda(),
                                   sl.cuda(), 16, 530, None, "auto", 1.0)
        else:
            tmp = torch.empty(2, 32, 2, 128, dtype=dtype, device="cuda")
            es = torch.empty(2, 32, 2, device="cuda")
            ml = torch.empty_like(es)
            ops.paged_attention_v2(out, es, ml, tmp, q.cuda(), kc.cuda(), vc.cuda(), 8,
                                   128 ** -0.5, bt.cuda(), sl.cuda(), 16, 530, None, "auto", 1.0)
        torch.testing.assert_close(out.cpu().float(), ref.float(), atol=1e-3, rtol=1e-2)


def test_blocksparse_matches_dense_when_everything_is_local(ops):
    # local window covering the whole sequence == dense attention (attention_kernels.cu:232-238)
    dtype = torch.float16
    q, kc, vc, bt, sl = ar.make_paged_inputs(3, 16, 4, 128, 16, [40, 200, 513], dtype, seed=31)
    ref = ar.paged_attention_ref(q, kc, vc, bt, sl, 128 ** -0.5)
    out = torch.empty_like(q, device="cuda")
    ops.paged_attention_v1(out, q.cuda(), kc.cuda(), vc.cuda(), 4, 128 ** -0.5, bt.cuda(),
                           sl.cuda(), 16, 513, None, "auto", 1.0, 0, 1000, 8, 64, 0)
    torch.testing.assert_close(out.cpu().float(), ref.float(), atol=1e-3, rtol=1e-2)


def test_rejects_what_the_reference_rejects(ops):
    q, kc, vc, bt, sl = ar.make_paged_inputs(1, 8, 8, 128, 16, [16], torch.float16)
    out = torch.empty_like(q, device="cuda")
    with pytest.raises(RuntimeError, match="Unsupported data type of kv cache"):
        ops.paged_attention_v1(out, q.cuda(), kc.cuda(), vc.cuda(), 8, 0.1, bt.cuda(), sl.cuda(),
                               16, 16, None, "int4", 1.0)
    with pytest.raises((RuntimeError, NotImplementedError), match="Unsupported block size"):
        kc4 = kc.view(-1, 8, 16, 4, 8)[: kc.shape[0]].contiguous()
        ops.paged_attention_v1(out, q.cuda(), kc4.cuda(), vc.cuda(), 8, 0.1, bt.cuda(), sl.cuda(),
                               4, 16, None, "auto", 1.0)


def test_cuda_graph_capture(ops):
    """Decode is replayed from CUDA graphs in the reference (worker/model_runner.py:910-1060)."""
    dtype = torch.float16
    q, kc, vc, bt, sl = ar.make_paged_inputs(4, 32, 8, 128, 16, [100, 600, 1024, 7], dtype, seed=41)
    ref = ar.paged_attention_ref(q, kc, vc, bt, sl, 128 ** -0.5)
    qd, kcd, vcd, btd, sld = (t.cuda() for t in (q, kc, vc, bt, sl))
    out = torch.zeros_like(qd)
    tmp = torch.empty(4, 32, 2, 128, dtype=dtype, device="cuda")
    es = torch.empty(4, 32, 2, device="cuda")
    ml = torch.empty_like(es)
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        ops.paged_attention_v2(out, es, ml, tmp, qd, kcd, vcd, 8, 128 ** -0.5, btd, sld, 16, 1024,
                               None, "auto", 1.0)
        s.synchronize()
        g = torch.cuda.CUDAGraph()
        out.zero_()
        with torch.cuda.graph(g, stream=s):
            ops.paged_attention_v2(out, es, ml, tmp, qd, kcd, vcd, 8, 128 ** -0.5, btd, sld, 16,
                                   1024, None, "auto", 1.0)
        out.zero_()
        g.replay()
        s.synchronize()
    torch.testing.assert_close(out.cpu().float(), ref.float(), atol=1e-3, rtol=1e-2)


# ------------------------------------------------------------------------------------------
# BASELINE.json configs[0] (C1): opt-125m on the reference CPU backend -- 12 heads, D 64, fp32,
# batch 8, block 16, context growing 128 -> 256.  Op-level replay of decode steps, GPU ops against
# the reference's OWN compiled CPU kernels (oracle/_ref), not just the restatement.
# ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("dtype", ["float", "bf16"])
def test_c1_opt125m_decode_replay_vs_reference_cpu(ops, dtype):
    from oracle import ref_cpu
    if not ref_cpu.available():
        pytest.skip(f"compiled reference CPU kernels unavailable: {ref_cpu.why_unavailable()}")
    dt = DTYPES[dtype]
    heads, d, batch, bs = 12, 64, 8, 16
    x = 16 // torch.tensor([], dtype=dt).element_size()
    nb = batch * (256 // bs)
    g = torch.Generator().manual_seed(0)
    kc = torch.zeros(nb, heads, d // x, bs, x, dtype=dt)
    vc = torch.zeros(nb, heads, d, bs, dtype=dt)
    bt = torch.randperm(nb, generator=g).to(torch.int32).view(batch, 256 // bs)
    # prefill 128 tokens of context through the CPU reference's cache writer
    ctx = 128
    k0 = (torch.rand(batch * ctx, heads, d, generator=g) - 0.5).to(dt)
    v0 = (torch.rand(batch * ctx, heads, d, generator=g) - 0.5).to(dt)
    pos = torch.arange(ctx).repeat(batch)
    seq = torch.arange(batch).repeat_interleave(ctx)
    slots0 = bt[seq, pos // bs].to(torch.int64) * bs + pos % bs
    ref_cpu.reshape_and_cache(k0, v0, kc, vc, slots0)
    kcd, vcd = kc.cuda(), vc.cuda()
    ops.reshape_and_cache(k0.cuda(), v0.cuda(), torch.zeros_like(kcd), torch.zeros_like(vcd), slots0.cuda(),
                          "auto", 1.0)  # exercised below on the live caches
    kcd.zero_(), vcd.zero_()
    ops.reshape_and_cache(k0.cuda(), v0.cuda(), kcd, vcd, slots0.cuda(), "auto", 1.0)
    assert torch.equal(kcd.cpu(), kc) and torch.equal(vcd.cpu(), vc)  # cache bytes identical
    for step in (0, 1, 63, 127):  # decode steps: ctx 128+step, write then attend
        n = ctx + step
        q = (torch.rand(batch, heads, d, generator=g) - 0.5).to(dt)
        k1 = (torch.rand(batch, heads, d, generator=g) - 0.5).to(dt)
        v1 = (torch.rand(batch, heads, d, generator=g) - 0.5).to(dt)
        slots = bt[:, n // bs].to(torch.int64) * bs + n % bs
        ref_cpu.reshape_and_cache(k1, v1, kc, vc, slots)
        ops.reshape_and_cache(k1.cuda(), v1.cuda(), kcd, vcd, slots.cuda(), "auto", 1.0)
        sl = torch.full((batch,), n + 1, dtype=torch.int32)
        ref = torch.empty_like(q)
        ref_cpu.paged_attention_v1(ref, q, kc, vc, heads, d ** -0.5, bt, sl, bs, n + 1)
        out = torch.empty_like(q, device="cuda")
        ops.paged_attention_v1(out, q.cuda(), kcd, vcd, heads, d ** -0.5, bt.cuda(), sl.cuda(), bs, n + 1,
                               None, "auto", 1.0)
        torch.testing.assert_close(out.cpu().float(), ref.float(), atol=1e-3, rtol=1e-2)
    assert torch.equal(kcd.cpu(), kc) and torch.equal(vcd.cpu(), vc)


@pytest.mark.parametrize("version", ["v1", "v2"])
@pytest.mark.parametrize("kv", ["auto", "fp8"])
@pytest.mark.parametrize("dtype", ["half", "bf16"])
def test_one_wave_grid_takes_three_warp_ctas(ops, version, kv, dtype):
    """64 sequences x 8 kv heads = 512 CTAs: between three and four CTAs per SM, where the tensor-core
    kernel runs 3-warp CTAs (four per SM) so that the grid is one wave (paged_attention.cu).  Ragged
    lengths, including ones that leave a warp without any 16-token item."""
    dt = DTYPES[dtype]
    g = torch.Generator().manual_seed(7)
    lens = torch.randint(1, 400, (64,), generator=g).tolist()
    lens[0], lens[1], lens[2], lens[3] = 1, 16, 17, 48
    run_case(ops, version, dt, 32, 8, 128, 16, lens, kv_cache_dtype=kv, seed=11)
