"""Host-side mirror of the reference's op surface (``vllm/_custom_ops.py``) for the hot path.

Same function names, argument order, defaults and return conventions as the reference
wrappers, so code written against ``from vllm import _custom_ops as ops`` runs against
``from neural_magic_vllm_b200 import ops`` unchanged.  Every function is a thin call into the
``torch.ops._C*`` ops registered by ``_C.so``; nothing here computes, and nothing falls back.

Reference line numbers refer to ``vllm/_custom_ops.py``.
"""
from __future__ import annotations

from typing import List, Optional, Tuple

import torch

_C = torch.ops._C
_cache = torch.ops._C_cache_ops
_utils = torch.ops._C_cuda_utils


def is_custom_op_supported(op_name: str) -> bool:  # :24-26
    ns, _, name = op_name.partition("::")
    return hasattr(getattr(torch.ops, ns), name)


# ------------------------------------------------------------------ attention (:74-129)
def paged_attention_v1(out, query, key_cache, value_cache, num_kv_heads: int, scale: float,
                       block_tables, seq_lens, block_size: int, max_seq_len: int,
                       alibi_slopes: Optional[torch.Tensor], kv_cache_dtype: str,
                       kv_scale: float, tp_rank: int = 0, blocksparse_local_blocks: int = 0,
                       blocksparse_vert_stride: int = 0, blocksparse_block_size: int = 64,
                       blocksparse_head_sliding_step: int = 0) -> None:
    _C.paged_attention_v1(out, query, key_cache, value_cache, num_kv_heads, scale, block_tables,
                          seq_lens, block_size, max_seq_len, alibi_slopes, kv_cache_dtype,
                          kv_scale, tp_rank, blocksparse_local_blocks, blocksparse_vert_stride,
                          blocksparse_block_size, blocksparse_head_sliding_step)


def paged_attention_v2(out, exp_sum, max_logits, tmp_out, query, key_cache, value_cache,
                       num_kv_heads: int, scale: float, block_tables, seq_lens, block_size: int,
                       max_seq_len: int, alibi_slopes: Optional[torch.Tensor],
                       kv_cache_dtype: str, kv_scale: float, tp_rank: int = 0,
                       blocksparse_local_blocks: int = 0, blocksparse_vert_stride: int = 0,
                       blocksparse_block_size: int = 64,
                       blocksparse_head_sliding_step: int = 0) -> None:
    _C.paged_attention_v2(out, exp_sum, max_logits, tmp_out, query, key_cache, value_cache,
                          num_kv_heads, scale, block_tables, seq_lens, block_size, max_seq_len,
                          alibi_slopes, kv_cache_dtype, kv_scale, tp_rank,
                          blocksparse_local_blocks, blocksparse_vert_stride,
                          blocksparse_block_size, blocksparse_head_sliding_step)


# ------------------------------------------------------------------ cache ops (:370-420)
def reshape_and_cache(key, value, key_cache, value_cache, slot_mapping, kv_cache_dtype: str,
                      kv_scale: float) -> None:
    _cache.reshape_and_cache(key, value, key_cache, value_cache, slot_mapping, kv_cache_dtype,
                             kv_scale)


def reshape_and_cache_flash(key, value, key_cache, value_cache, slot_mapping,
                            kv_cache_dtype: str) -> None:
    _cache.reshape_and_cache_flash(key, value, key_cache, value_cache, slot_mapping,
                                   kv_cache_dtype)


def copy_blocks(key_caches: List[torch.Tensor], value_caches: List[torch.Tensor],
                block_mapping: torch.Tensor) -> None:
    _cache.copy_blocks(key_caches, value_caches, block_mapping)


def swap_blocks(src: torch.Tensor, dst: torch.Tensor, block_mapping: torch.Tensor) -> None:
    _cache.swap_blocks(src, dst, block_mapping)


def convert_fp8(output: torch.Tensor, input: torch.Tensor, scale: float = 1.0,
                kv_dtype: str = "fp8") -> None:
    _cache.convert_fp8(output, input, scale, kv_dtype)


def get_device_attribute(attribute: int, device: int) -> int:
    return _utils.get_device_attribute(attribute, device)


def get_max_shared_memory_per_block_device_attribute(device: int) -> int:
    return _utils.get_max_shared_memory_per_block_device_attribute(device)
