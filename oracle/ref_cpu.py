"""Loader for ``oracle/_ref/_C_ref_cpu.so``: the reference's OWN CPU backend
(``/root/reference/csrc/cpu/*.cpp``) compiled in place by ``oracle/build_ref.py``.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).

The ops are registered under ``torch.ops._C_ref`` / ``torch.ops._C_ref_cache_ops`` so they
coexist with the product's ``_C`` registration.  Coverage of the compiled reference
(``csrc/cpu/attention.cpp:412-432``, ``cache.cpp:110``, ``cpu_types_x86.hpp:16-18``):
fp32 and bf16, block_size 16, kv_scale 1.0, kv_cache_dtype "auto", no block-sparse.
"""
from __future__ import annotations

import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(_HERE, "_ref", "_C_ref_cpu.so")
_state = {"loaded": False, "why": None}


def _host_has(flags: list[str]) -> bool:
    try:
        cpuinfo = open("/proc/cpuinfo").read()
    except OSError:
        return False
    names = {"avx512f": "avx512f", "avx512vl": "avx512vl", "avx512bw": "avx512bw",
             "avx512dq": "avx512dq", "avx512bf16": "avx512_bf16", "avx2": "avx2"}
    return all(names.get(f, f) in cpuinfo for f in flags)


def available() -> bool:
    """True when the compiled reference exists and this host can execute it."""
    if _state["loaded"]:
        return True
    if _state["why"] is not None:
        return False
    if not os.path.exists(SO):
        _state["why"] = "oracle/_ref/_C_ref_cpu.so not built (run oracle/build_ref.py)"
        return False
    isa_file = os.path.join(os.path.dirname(SO), "isa.txt")
    flags = open(isa_file).read().split() if os.path.exists(isa_file) else []
    if not _host_has(flags):
        _state["why"] = f"host CPU lacks {flags}"
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


def paged_attention_v1(out, query, key_cache, value_cache, num_kv_heads, scale, block_tables,
                       seq_lens, block_size, max_seq_len, alibi_slopes=None):
    assert available(), why_unavailable()
    torch.ops._C_ref.paged_attention_v1(out, query, key_cache, value_cache, num_kv_heads, scale,
                                        block_tables, seq_lens, block_size, max_seq_len,
                                        alibi_slopes, "auto", 1.0, 0, 0, 0, 64, 0)


def paged_attention_v2(out, exp_sums, max_logits, tmp_out, query, key_cache, value_cache,
                       num_kv_heads, scale, block_tables, seq_lens, block_size, max_seq_len,
                       alibi_slopes=None):
    assert available(), why_unavailable()
    torch.ops._C_ref.paged_attention_v2(out, exp_sums, max_logits, tmp_out, query, key_cache,
                                        value_cache, num_kv_heads, scale, block_tables, seq_lens,
                                        block_size, max_seq_len, alibi_slopes, "auto", 1.0, 0, 0, 0,
                                        64, 0)


def reshape_and_cache(key, value, key_cache, value_cache, slot_mapping):
    assert available(), why_unavailable()
    torch.ops._C_ref_cache_ops.reshape_and_cache(key, value, key_cache, value_cache,
                                                 slot_mapping, "auto", 1.0)


def copy_blocks(key_caches, value_caches, block_mapping):
    assert available(), why_unavailable()
    torch.ops._C_ref_cache_ops.copy_blocks(key_caches, value_caches, block_mapping)
