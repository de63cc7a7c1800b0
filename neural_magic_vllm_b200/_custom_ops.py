"""Host-side mirror of the reference's op surface (``vllm/_custom_ops.py``) for the hot path.

Same function names, argument order, defaults and return conventions as the reference
wrappers, so code written against ``from vllm import _custom_ops as ops`` runs against
``from neural_magic_vllm_b200 import ops`` unchanged.  Every function is a thin call into the
``torch.ops._C*`` ops registered by ``_C.so``; nothing here computes, and nothing falls back.

Reference line numbers refer to ``vllm/_custom_ops.py``.
"""
from __future__ import annotations

from typing import List, Optional, Tuple, Type

import torch

_C = torch.ops._C
_cache = torch.ops._C_cache_ops
_utils = torch.ops._C_cuda_utils


def is_custom_op_supported(op_name: str) -> bool:  # :24-26
    ns, _, name = op_name.partition("::")
    return hasattr(getattr(torch.ops, ns), name)


# ------------------------------------------------------------------ activation ops (:49-70)
def silu_and_mul(out: torch.Tensor, x: torch.Tensor) -> None:
    _C.silu_and_mul(out, x)


def gelu_and_mul(out: torch.Tensor, x: torch.Tensor) -> None:
    _C.gelu_and_mul(out, x)


def gelu_tanh_and_mul(out: torch.Tensor, x: torch.Tensor) -> None:
    _C.gelu_tanh_and_mul(out, x)


def gelu_fast(out: torch.Tensor, x: torch.Tensor) -> None:
    _C.gelu_fast(out, x)


def gelu_new(out: torch.Tensor, x: torch.Tensor) -> None:
    _C.gelu_new(out, x)


def gelu_quick(out: torch.Tensor, x: torch.Tensor) -> None:
    _C.gelu_quick(out, x)


# ------------------------------------------------------------------ pos encoding ops (:133-152)
def rotary_embedding(positions: torch.Tensor, query: torch.Tensor, key: torch.Tensor,
                     head_size: int, cos_sin_cache: torch.Tensor, is_neox: bool) -> None:
    _C.rotary_embedding(positions, query, key, head_size, cos_sin_cache, is_neox)


def batched_rotary_embedding(positions: torch.Tensor, query: torch.Tensor, key: torch.Tensor,
                             head_size: int, cos_sin_cache: torch.Tensor, is_neox: bool,
                             rot_dim: int, cos_sin_cache_offsets: torch.Tensor) -> None:
    _C.batched_rotary_embedding(positions, query, key, head_size, cos_sin_cache, is_neox, rot_dim,
                                cos_sin_cache_offsets)


def rotary_embedding_and_cache(positions: torch.Tensor, query: torch.Tensor, key: torch.Tensor,
                               value: torch.Tensor, head_size: int, cos_sin_cache: torch.Tensor,
                               is_neox: bool, key_cache: torch.Tensor, value_cache: torch.Tensor,
                               slot_mapping: torch.Tensor, kv_cache_dtype: str,
                               kv_scale: float) -> None:
    """B200 extension (no reference wrapper): ``rotary_embedding`` followed by
    ``reshape_and_cache`` in one launch; results identical to the two reference ops in sequence."""
    _C.rotary_embedding_and_cache(positions, query, key, value, head_size, cos_sin_cache, is_neox,
                                  key_cache, value_cache, slot_mapping, kv_cache_dtype, kv_scale)


# ------------------------------------------------------------------ layer norm ops (:155-163)
def rms_norm(out: torch.Tensor, input: torch.Tensor, weight: torch.Tensor,
             epsilon: float) -> None:
    _C.rms_norm(out, input, weight, epsilon)


def rms_norm_quant(input: torch.Tensor, weight: torch.Tensor, epsilon: float,
                   residual: Optional[torch.Tensor] = None, scale: Optional[torch.Tensor] = None,
                   out_dtype: torch.dtype = torch.int8) -> Tuple[torch.Tensor, torch.Tensor]:
    """Extension: (fused_add_)rms_norm followed by scaled_int8_quant /
    scaled_fp8_quant in ONE launch, bit-identical to the two reference ops.  `residual` is updated
    in place when given.  int8: scale=None -> per-token dynamic scales [tokens, 1]; a scalar tensor
    -> static.  fp8: static only (scale required)."""
    out = torch.empty(input.shape, dtype=out_dtype, device=input.device)
    dynamic = scale is None
    if dynamic:
        scale = torch.empty((input.numel() // input.shape[-1], 1), dtype=torch.float32, device=input.device)
    _C.rms_norm_quant(out, scale, input, residual, weight, epsilon, dynamic)
    return out, scale


def fused_add_rms_norm(input: torch.Tensor, residual: torch.Tensor, weight: torch.Tensor,
                       epsilon: float) -> None:
    _C.fused_add_rms_norm(input, residual, weight, epsilon)


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


def flash_attn_varlen_func(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, cu_seqlens_q: torch.Tensor,
                           cu_seqlens_k: torch.Tensor, max_seqlen_q: int, max_seqlen_k: int,
                           softmax_scale: Optional[float] = None, causal: bool = False,
                           out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Prefill attention over packed variable-length sequences, with the call shape of the
    flash_attn_varlen_func the reference's FlashAttention backend uses
    (vllm/attention/backends/flash_attn.py:320-346): q [total_q, H, D], k / v [total_k, KVH, D],
    int32 cu_seqlens.  Extension op: the reference calls a library here."""
    if softmax_scale is None:
        softmax_scale = q.shape[-1] ** -0.5
    if out is None:
        out = torch.empty_like(q)
    _C.varlen_attention_fwd(out, q, k, v, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k,
                            softmax_scale, causal)
    return out


# ------------------------------------------------------------------ quantized linears (:208-352)
def gptq_marlin_24_gemm(a, b_q_weight, b_meta, b_scales, workspace, num_bits: int, size_m: int,
                        size_n: int, size_k: int) -> torch.Tensor:  # :208-215
    return _C.gptq_marlin_24_gemm(a, b_q_weight, b_meta, b_scales, workspace, num_bits, size_m,
                                  size_n, size_k)


def cutlass_scaled_mm_supports_fp8(cuda_device_capability: int) -> bool:  # :219-220
    return _C.cutlass_scaled_mm_supports_fp8(cuda_device_capability)


def cutlass_scaled_mm(a, b, scale_a, scale_b, out_dtype: Type[torch.dtype],
                      bias: Optional[torch.Tensor] = None) -> torch.Tensor:  # :223-238
    assert (b.shape[0] % 16 == 0 and b.shape[1] % 16 == 0)
    assert (out_dtype is torch.bfloat16 or out_dtype is torch.float16)
    out = torch.empty((a.shape[0], b.shape[1]), dtype=out_dtype, device=a.device)
    _C.cutlass_scaled_mm(out, a, b, scale_a, scale_b, bias)
    return out


def gptq_marlin_repack(b_q_weight, perm, size_k: int, size_n: int, num_bits: int) -> torch.Tensor:
    return _C.gptq_marlin_repack(b_q_weight, perm, size_k, size_n, num_bits)  # :257-261


def gptq_marlin_gemm(a, b_q_weight, b_scales, g_idx, perm, workspace, num_bits: int, size_m: int,
                     size_n: int, size_k: int, is_k_full: bool) -> torch.Tensor:  # :264-271
    return _C.gptq_marlin_gemm(a, b_q_weight, b_scales, g_idx, perm, workspace, num_bits, size_m,
                               size_n, size_k, is_k_full)


def gptq_marlin_gemm_add_rms_norm(a, b_q_weight, b_scales, g_idx, perm, workspace, num_bits: int, size_m: int,
                                  size_n: int, size_k: int, is_k_full: bool, residual: torch.Tensor,
                                  norm_weight: torch.Tensor, epsilon: float) -> torch.Tensor:
    """Extension: ``gptq_marlin_gemm`` followed by ``fused_add_rms_norm(out, residual, norm_weight, epsilon)``
    (residual updated in place, returns the normalised rows) -- one kernel less where the GEMM splits K."""
    return _C.gptq_marlin_gemm_add_rms_norm(a, b_q_weight, b_scales, g_idx, perm, workspace, num_bits, size_m,
                                            size_n, size_k, is_k_full, residual, norm_weight, epsilon)


# awq (vllm/_custom_ops.py:167-177; the parameter NAMES of awq_gemm are swapped there, the positional
# order -- input, qweight, scales, qzeros -- is what awq.py:172 passes and what the C++ op takes)
def awq_dequantize(qweight: torch.Tensor, scales: torch.Tensor, zeros: torch.Tensor, split_k_iters: int,
                   thx: int, thy: int) -> torch.Tensor:
    return _C.awq_dequantize(qweight, scales, zeros, split_k_iters, thx, thy)


def awq_gemm(input: torch.Tensor, qweight: torch.Tensor, qzeros: torch.Tensor, scales: torch.Tensor,
             split_k_iters: int) -> torch.Tensor:
    return _C.awq_gemm(input, qweight, qzeros, scales, split_k_iters)


def fp8_marlin_gemm(a, b_q_weight, b_scales, workspace, num_bits: int, size_m: int, size_n: int,
                    size_k: int) -> torch.Tensor:  # :275-280
    return _C.fp8_marlin_gemm(a, b_q_weight, b_scales, workspace, num_bits, size_m, size_n, size_k)


def scaled_fp8_quant(input: torch.Tensor, scale: Optional[torch.Tensor] = None,
                     batch_dim_padding: Optional[int] = None
                     ) -> Tuple[torch.Tensor, torch.Tensor]:  # :284-320
    if batch_dim_padding:
        shape = (max(batch_dim_padding, input.shape[0]), *input.shape[1:])
        output = torch.empty(shape, device=input.device, dtype=torch.float8_e4m3fn)
    else:
        output = torch.empty_like(input, dtype=torch.float8_e4m3fn)
    if scale is None:
        scale = torch.zeros(1, device=input.device, dtype=torch.float32)
        _C.dynamic_scaled_fp8_quant(output, input, scale)
    else:
        _C.static_scaled_fp8_quant(output, input, scale)
    return output, scale


def scaled_int8_quant(input: torch.Tensor, scale: Optional[torch.Tensor] = None
                      ) -> Tuple[torch.Tensor, torch.Tensor]:  # :324-350
    output = torch.empty_like(input, dtype=torch.int8)
    if scale is not None:
        _C.static_scaled_int8_quant(output, input, scale)
        return output, scale
    input_scales = torch.empty((input.numel() // input.shape[-1], 1), device=input.device,
                               dtype=torch.float32)
    _C.dynamic_scaled_int8_quant(output, input, input_scales)
    return output, input_scales


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
