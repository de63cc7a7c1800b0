"""CPU restatement of the reference's paged-attention semantics (TEST INFRASTRUCTURE ONLY).

Follows ``tests/kernels/test_attention.py:46-115`` of the reference
(``ref_masked_attention`` / ``ref_single_query_cached_kv_attention``): per sequence, gather
K/V through the block table, logits = scale * q.k (+ ALiBi bias ``slope * (pos - len + 1)``),
softmax in fp32, probabilities cast to the value dtype, out = P.V.  fp8 caches are first
dequantised exactly as ``ops.convert_fp8`` does (``tests/kernels/test_attention.py:243-261``,
``csrc/quantization/fp8/nvidia/quant_utils.cuh:295-301``): fp8 -> half -> float * kv_scale ->
model dtype.

Pinned against the compiled reference CPU kernels (``oracle/_ref``) in
``tests/test_oracle.py`` and against the committed golden vectors in ``tests/golden``.

Layouts (reference ``vllm/attention/ops/paged_attn.py:36-60``):
  key_cache   [num_blocks, num_kv_heads, head_size/x, block_size, x]
  value_cache [num_blocks, num_kv_heads, head_size, block_size]
"""
from __future__ import annotations

from typing import Optional

import torch

FP8_DTYPES = {"fp8": torch.float8_e4m3fn, "fp8_e4m3": torch.float8_e4m3fn,
              "fp8_e5m2": torch.float8_e5m2}


def fp8_quantize(x: torch.Tensor, scale: float = 1.0, kv_dtype: str = "fp8") -> torch.Tensor:
    """``convert_fp8`` towards fp8 (quant_utils.cuh:440-475): cvt_satfinite_rn(float(x)/scale)."""
    ft = FP8_DTYPES[kv_dtype]
    v = x.float() / scale
    lim = torch.finfo(ft).max
    v = torch.where(torch.isnan(v), v, v.clamp(-lim, lim))
    return v.to(ft).view(torch.uint8)


def fp8_dequantize(x_u8: torch.Tensor, dtype: torch.dtype, scale: float = 1.0,
                   kv_dtype: str = "fp8") -> torch.Tensor:
    """``convert_fp8`` from fp8 (quant_utils.cuh:295-301): dtype(float(half(fp8)) * scale)."""
    return (x_u8.view(FP8_DTYPES[kv_dtype]).to(torch.float16).float() * scale).to(dtype)


def gather_kv(key_cache: torch.Tensor, value_cache: torch.Tensor, block_table: torch.Tensor,
              seq_len: int):
    """-> keys [seq_len, kv_heads, D], values [seq_len, kv_heads, D] for one sequence."""
    nb, kvh, dx, bs, x = key_cache.shape
    n_blocks = (seq_len + bs - 1) // bs
    blocks = block_table[:n_blocks].long()
    k = key_cache[blocks]                       # [n, kvh, D/x, bs, x]
    k = k.permute(0, 3, 1, 2, 4).reshape(n_blocks * bs, kvh, dx * x)[:seq_len]
    v = value_cache[blocks]                     # [n, kvh, D, bs]
    v = v.permute(0, 3, 1, 2).reshape(n_blocks * bs, kvh, -1)[:seq_len]
    return k, v


def paged_attention_ref(query: torch.Tensor, key_cache: torch.Tensor, value_cache: torch.Tensor,
                        block_tables: torch.Tensor, seq_lens: torch.Tensor, scale: float,
                        alibi_slopes: Optional[torch.Tensor] = None, kv_cache_dtype: str = "auto",
                        kv_scale: float = 1.0) -> torch.Tensor:
    """Returns out [num_seqs, num_heads, head_size] in ``query.dtype`` (CPU tensors)."""
    dtype = query.dtype
    if kv_cache_dtype != "auto":
        key_cache = fp8_dequantize(key_cache, dtype, kv_scale, kv_cache_dtype)
        value_cache = fp8_dequantize(value_cache, dtype, kv_scale, kv_cache_dtype)
    num_seqs, num_heads, head_size = query.shape
    kvh = value_cache.shape[1]
    group = num_heads // kvh
    out = torch.zeros_like(query)
    for i in range(num_seqs):
        n = int(seq_lens[i])
        if n == 0:
            continue
        k, v = gather_kv(key_cache, value_cache, block_tables[i], n)
        if group > 1:
            k = k.repeat_interleave(group, dim=1)
            v = v.repeat_interleave(group, dim=1)
        # test_attention.py:46-60 computes the einsum in the model dtype and softmax in fp32;
        # the kernels accumulate q.k in fp32 (attention_utils.cuh:33-49) -- use fp32 here.
        logits = scale * torch.einsum("hd,khd->hk", query[i].float(), k.float())
        if alibi_slopes is not None:
            pos = torch.arange(n, dtype=torch.float32)
            logits = logits + alibi_slopes.float().view(-1, 1) * (pos - n + 1).view(1, -1)
        p = torch.softmax(logits, dim=-1).to(v.dtype)
        out[i] = torch.einsum("hk,khd->hd", p.float(), v.float()).to(dtype)
    return out


def make_paged_inputs(num_seqs: int, num_heads: int, num_kv_heads: int, head_size: int,
                      block_size: int, seq_lens, dtype: torch.dtype, seed: int = 0,
                      kv_cache_dtype: str = "auto", num_blocks: Optional[int] = None,
                      shuffle_blocks: bool = True):
    """Seeded synthetic inputs shaped like ``vllm/utils.py:515-563`` produces them:
    q, K, V ~ U(-s, s) with s = head_size**-0.5; block tables = distinct random blocks."""
    g = torch.Generator().manual_seed(seed)
    s = head_size ** -0.5
    seq_lens = torch.as_tensor(seq_lens, dtype=torch.int32)
    max_len = int(seq_lens.max()) if len(seq_lens) else 0
    max_blocks = max((max_len + block_size - 1) // block_size, 1)
    need = num_seqs * max_blocks
    num_blocks = num_blocks or need
    assert num_blocks >= need
    x = 16 // (1 if kv_cache_dtype != "auto" else torch.empty((), dtype=dtype).element_size())
    q = (torch.rand(num_seqs, num_heads, head_size, generator=g) * 2 - 1).mul_(s).to(dtype)
    kc = (torch.rand(num_blocks, num_kv_heads, head_size // x, block_size, x, generator=g) * 2 - 1
          ).mul_(s).to(dtype)
    vc = (torch.rand(num_blocks, num_kv_heads, head_size, block_size, generator=g) * 2 - 1
          ).mul_(s).to(dtype)
    if kv_cache_dtype != "auto":
        kc = fp8_quantize(kc, 1.0, kv_cache_dtype)
        vc = fp8_quantize(vc, 1.0, kv_cache_dtype)
    ids = torch.randperm(num_blocks, generator=g) if shuffle_blocks else torch.arange(num_blocks)
    bt = ids[:need].to(torch.int32).view(num_seqs, max_blocks).contiguous()
    return q, kc, vc, bt, seq_lens


def gather_last_tokens(key_cache, value_cache, block_tables, seq_lens):
    """-> key, value [num_seqs, kv_heads, D]: the last cached token of every sequence."""
    ks, vs = [], []
    for i, n in enumerate(seq_lens.tolist()):
        k, v = gather_kv(key_cache, value_cache, block_tables[i], n)
        ks.append(k[-1])
        vs.append(v[-1])
    return torch.stack(ks).contiguous(), torch.stack(vs).contiguous()


def varlen_attention_ref(q, k, v, cu_seqlens_q, cu_seqlens_k, softmax_scale, causal):
    """Per-sequence softmax(scale * Q K^T + mask) V in fp32, the arithmetic of the reference's CPU
    prefill path (vllm/attention/backends/torch_sdpa.py:203-218: scaled_dot_product_attention per
    sequence, K/V heads repeated for GQA) with the bottom-right aligned causal mask of
    flash_attn_varlen_func (backends/flash_attn.py:320-346).  q [Tq, H, D], k / v [Tk, KVH, D]."""
    out = torch.zeros(q.shape, dtype=torch.float32)
    h, kvh = q.shape[1], k.shape[1]
    cq, ck = cu_seqlens_q.tolist(), cu_seqlens_k.tolist()
    for s in range(len(cq) - 1):
        qs = q[cq[s]:cq[s + 1]].float().transpose(0, 1)                                   # [H, Lq, D]
        ks = k[ck[s]:ck[s + 1]].float().transpose(0, 1).repeat_interleave(h // kvh, 0)    # [H, Lk, D]
        vs = v[ck[s]:ck[s + 1]].float().transpose(0, 1).repeat_interleave(h // kvh, 0)
        lq, lk = qs.shape[1], ks.shape[1]
        if lq == 0:
            continue
        logits = torch.einsum("hqd,hkd->hqk", qs, ks) * softmax_scale
        if causal:
            mask = torch.ones(lq, lk, dtype=torch.bool).tril(diagonal=lk - lq)
            logits = logits.masked_fill(~mask, float("-inf"))
        p = torch.softmax(logits, dim=-1)
        p = torch.nan_to_num(p, nan=0.0)  # rows with no visible key
        out[cq[s]:cq[s + 1]] = torch.einsum("hqk,hkd->hqd", p, vs).transpose(0, 1)
    return out
