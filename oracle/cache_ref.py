"""CPU restatement of the reference's KV-cache maintenance ops (TEST INFRASTRUCTURE ONLY).

Each function follows the reference kernel it names, element by element, on CPU tensors:

* ``reshape_and_cache``        csrc/cache_kernels.cu:152-204 (+ host :253-278)
* ``reshape_and_cache_flash``  csrc/cache_kernels.cu:206-237 (+ host :280-314)
* ``copy_blocks``              csrc/cache_kernels.cu:67-148
* ``swap_blocks``              csrc/cache_kernels.cu:24-63
* ``convert_fp8``              csrc/cache_kernels.cu:316-389 with
                               csrc/quantization/fp8/nvidia/quant_utils.cuh:288-478

fp8 arithmetic (quant_utils.cuh): towards fp8 ``cvt.satfinite.rn(float(x) / scale)``
(:440-475); from fp8 ``T(float(half(fp8)) * scale)`` (:295-301, :352-361).

Pinned against the reference's OWN compiled CPU kernels (``oracle/_ref``: ``reshape_and_cache``,
``copy_blocks`` for fp32 / bf16, csrc/cpu/cache.cpp) in ``tests/test_cache_oracle_cpu.py``; the fp8
conversions are pinned there against torch's ``float8_e4m3fn`` / ``float8_e5m2`` casts on every
one of the 256 byte codes and on the reference test's own U(-224, 224) round trip
(tests/kernels/test_cache.py:369-394).
"""
from __future__ import annotations

from typing import List, Sequence

import torch

from .attention_ref import FP8_DTYPES, fp8_dequantize, fp8_quantize

__all__ = ["reshape_and_cache", "reshape_and_cache_flash", "copy_blocks", "swap_blocks",
           "convert_fp8", "fp8_quantize", "fp8_dequantize"]


def reshape_and_cache(key: torch.Tensor, value: torch.Tensor, key_cache: torch.Tensor,
                      value_cache: torch.Tensor, slot_mapping: torch.Tensor,
                      kv_cache_dtype: str = "auto", kv_scale: float = 1.0) -> None:
    """In place on ``key_cache [NB, H, D/x, bs, x]`` / ``value_cache [NB, H, D, bs]``.

    ``key`` / ``value`` are ``[T, H, D]`` (any stride(0)); ``slot_mapping[T]`` int64, negative
    = padding token, skipped (cache_kernels.cu:166-169).  fp8 caches are uint8 and receive
    ``fp8_quantize(x, kv_scale)`` (:190-199)."""
    nb, h, dx, bs, x = key_cache.shape
    d = dx * x
    t = key.shape[0]
    assert value_cache.shape == (nb, h, d, bs)
    for i in range(t):
        slot = int(slot_mapping[i])
        if slot < 0:
            continue
        b, off = slot // bs, slot % bs
        k = key[i].reshape(h, dx, x)
        v = value[i].reshape(h, d)
        if kv_cache_dtype == "auto":
            key_cache[b, :, :, off, :] = k
            value_cache[b, :, :, off] = v
        else:
            key_cache[b, :, :, off, :] = fp8_quantize(k, kv_scale, kv_cache_dtype)
            value_cache[b, :, :, off] = fp8_quantize(v, kv_scale, kv_cache_dtype)


def reshape_and_cache_flash(key: torch.Tensor, value: torch.Tensor, k_cache: torch.Tensor,
                            v_cache: torch.Tensor, slot_mapping: torch.Tensor) -> None:
    """Flash layout ``[NB, bs, H, D]`` (cache_kernels.cu:206-237); negative slots skipped
    (:217-220)."""
    nb, bs, h, d = k_cache.shape
    for i in range(key.shape[0]):
        slot = int(slot_mapping[i])
        if slot < 0:
            continue
        b, off = slot // bs, slot % bs
        k_cache[b, off] = key[i].reshape(h, d)
        v_cache[b, off] = value[i].reshape(h, d)


def copy_blocks(key_caches: Sequence[torch.Tensor], value_caches: Sequence[torch.Tensor],
                block_mapping: torch.Tensor) -> None:
    """``block_mapping [P, 2]`` int64 (src, dst); pairs applied in order to every layer
    (cache_kernels.cu:67-99). The reference test guarantees src and dst sets are disjoint."""
    for src, dst in block_mapping.tolist():
        for kc in key_caches:
            kc[dst].copy_(kc[src])
        for vc in value_caches:
            vc[dst].copy_(vc[src])


def swap_blocks(src: torch.Tensor, dst: torch.Tensor, block_mapping: torch.Tensor) -> None:
    """Whole-block memcpy ``dst[d] = src[s]`` per mapping row (cache_kernels.cu:46-62)."""
    for s, d in block_mapping.tolist():
        dst[d].copy_(src[s])


def convert_fp8(dst: torch.Tensor, src: torch.Tensor, scale: float = 1.0,
                kv_dtype: str = "fp8") -> None:
    """Direction chosen by which side is uint8 (cache_kernels.cu:372-385). The reference accepts
    only "fp8" / "fp8_e4m3" here ("auto" reaches a device ``assert(false)``; anything else is a
    host error, :386-388); the e5m2 branch below exists for the reshape_and_cache fp8_e5m2 tests."""
    if kv_dtype not in FP8_DTYPES:
        raise RuntimeError(f"Unsupported data type: {kv_dtype}")
    if dst.dtype == torch.uint8:
        dst.copy_(fp8_quantize(src, scale, kv_dtype))
    else:
        dst.copy_(fp8_dequantize(src, dst.dtype, scale, kv_dtype))


def make_kv_caches(num_blocks: int, block_size: int, num_layers: int, num_heads: int,
                   head_size: int, cache_dtype: str, model_dtype: torch.dtype, seed: int
                   ) -> tuple[List[torch.Tensor], List[torch.Tensor]]:
    """CPU twin of ``vllm/utils.py:create_kv_caches_with_random`` (:407-455): U(-s, s) with
    s = head_size**-0.5; fp8 caches are the uniform fp16 values pushed through convert_fp8."""
    g = torch.Generator().manual_seed(seed)
    s = head_size ** -0.5
    wide = model_dtype if cache_dtype == "auto" else torch.float16
    x = 16 // (torch.tensor([], dtype=wide if cache_dtype == "auto" else torch.uint8).element_size())
    kcs, vcs = [], []
    for _ in range(num_layers):
        k = (torch.rand(num_blocks, num_heads, head_size // x, block_size, x, generator=g) * 2 - 1) * s
        v = (torch.rand(num_blocks, num_heads, head_size, block_size, generator=g) * 2 - 1) * s
        k, v = k.to(wide), v.to(wide)
        if cache_dtype != "auto":
            k, v = fp8_quantize(k, 1.0, cache_dtype), fp8_quantize(v, 1.0, cache_dtype)
        kcs.append(k)
        vcs.append(v)
    return kcs, vcs
