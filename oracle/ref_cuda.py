"""Loader for ``oracle/_ref/_C_refcuda.so``: the reference's OWN CUDA kernels
(``/root/reference/csrc/**/*.cu``) recompiled for sm_100 by ``oracle/build_ref_cuda.py``.

TEST / BASELINE INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).

The ops are registered under ``torch.ops._C_refcuda`` / ``torch.ops._C_refcuda_cache_ops`` with the
reference's own schemas (``oracle/ref_cuda_bindings.cpp``), so they coexist with the product's
``_C`` registration: a GPU-side oracle on identical inputs at full size, and the same-box baseline
("the recompiled mma.sync / cp.async kernels") that benchmarks/*_bench.py time beside ours.
"""
from __future__ import annotations

import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(_HERE, "_ref", "_C_refcuda.so")
_state = {"loaded": False, "why": None}


def available() -> bool:
    if _state["loaded"]:
        return True
    if _state["why"] is not None:
        return False
    if not os.path.exists(SO):
        _state["why"] = "oracle/_ref/_C_refcuda.so not built (run oracle/build_ref_cuda.py where /root/reference exists)"
        return False
    if not torch.cuda.is_available():
        _state["why"] = "no GPU"
        return False
    try:
        torch.ops.load_library(SO)
    except OSError as e:  # pragma: no cover - depends on the box
        _state["why"] = f"load failed: {e}"
        return False
    _state["loaded"] = True
    return True


def why_unavailable() -> str | None:
    available()
    return _state["why"]


def ops():
    assert available(), why_unavailable()
    return torch.ops._C_refcuda


def cache_ops():
    assert available(), why_unavailable()
    return torch.ops._C_refcuda_cache_ops


def paged_attention_v1(out, query, key_cache, value_cache, num_kv_heads, scale, block_tables, seq_lens,
                       block_size, max_seq_len, alibi_slopes=None, kv_cache_dtype="auto", kv_scale=1.0):
    ops().paged_attention_v1(out, query, key_cache, value_cache, num_kv_heads, scale, block_tables,
                             seq_lens, block_size, max_seq_len, alibi_slopes, kv_cache_dtype, kv_scale,
                             0, 0, 0, 64, 0)


def paged_attention_v2(out, exp_sums, max_logits, tmp_out, query, key_cache, value_cache, num_kv_heads,
                       scale, block_tables, seq_lens, block_size, max_seq_len, alibi_slopes=None,
                       kv_cache_dtype="auto", kv_scale=1.0):
    ops().paged_attention_v2(out, exp_sums, max_logits, tmp_out, query, key_cache, value_cache,
                             num_kv_heads, scale, block_tables, seq_lens, block_size, max_seq_len,
                             alibi_slopes, kv_cache_dtype, kv_scale, 0, 0, 0, 64, 0)
