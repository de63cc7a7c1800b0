"""CPU restatement (numpy/torch) of the reference's quantized-linear layouts and math.

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline leg as the checker -- never by the product path.

Pinned: tests/golden/make_golden.py imports the reference's own Python utilities by file path
(vllm/model_executor/layers/quantization/utils/{quant_utils,marlin_perms,marlin_utils,
marlin_24_perms,format_24}.py) in the build container, runs them on seeded inputs and commits the
outputs under tests/golden/; tests/test_oracle_cpu.py checks every function here against them.

All paths cited are relative to the reference checkout.
"""
from __future__ import annotations

import numpy as np
import torch

MARLIN_TILE = 16


def get_pack_factor(num_bits: int) -> int:  # quant_utils.py:9-11
    assert num_bits in (4, 8)
    return 32 // num_bits


def _wrap_i32(x: torch.Tensor) -> torch.Tensor:
    """int64 holding a uint32 bit pattern -> the int32 with the same 32 bits (what the reference's
    numpy ``astype(np.int32)`` does).  Every function below is plain torch and runs on whatever
    device its input lives on, so full-size shapes can be quantised on the GPU by the tests."""
    x = x & 0xFFFFFFFF
    return torch.where(x >= 2 ** 31, x - 2 ** 32, x).to(torch.int32)


# ------------------------------------------------------------------------------------------
# quantize_weights (quant_utils.py:41-110): symmetric per-group quantisation, zero point 2^(b-1)
# ------------------------------------------------------------------------------------------
def quantize_weights(w: torch.Tensor, num_bits: int, group_size: int, act_order: bool = False,
                     generator: torch.Generator | None = None):
    size_k, size_n = w.shape
    if group_size == -1:
        group_size = size_k
    max_q = 2 ** num_bits - 1
    half_q = (max_q + 1) // 2
    wg = w
    if group_size < size_k:
        wg = w.reshape(-1, group_size, size_n).permute(1, 0, 2).reshape(group_size, -1)
    s = wg.abs().max(0, keepdim=True)[0]
    s = s * (2 / max_q)
    q = torch.round(wg / s).int() + half_q
    q = torch.clamp(q, 0, max_q)
    w_ref = (q - half_q).half() * s
    if group_size < size_k:
        def back(t):
            return t.reshape(group_size, -1, size_n).permute(1, 0, 2).reshape(size_k, size_n).contiguous()
        q, w_ref = back(q), back(w_ref)
    s = s.reshape(-1, size_n).contiguous()
    g_idx = torch.empty(0, dtype=torch.int32, device=w.device)
    rand_perm = torch.empty(0, dtype=torch.int64, device=w.device)
    if act_order:  # permute_rows (quant_utils.py:14-38)
        assert group_size < size_k
        g_idx = (torch.arange(size_k, device=w.device) // group_size).to(torch.int32)
        rand_perm = torch.randperm(size_k, generator=generator).to(w.device)
        g_idx = g_idx[rand_perm].contiguous()
        q = q[rand_perm].contiguous()
        w_ref = w_ref[rand_perm].contiguous()
    return w_ref, q, s, g_idx, rand_perm


def sort_weights(q_w: torch.Tensor, g_idx: torch.Tensor):  # quant_utils.py:113-126
    sort_indices = torch.argsort(g_idx).to(torch.int32)
    return q_w[sort_indices.long()].contiguous(), g_idx[sort_indices.long()].contiguous(), sort_indices


def gptq_pack(q_w: torch.Tensor, num_bits: int, size_k: int, size_n: int) -> torch.Tensor:
    """[K, N] ints -> [K/pack, N] int32, packed along K (quant_utils.py:129-152)."""
    pack = get_pack_factor(num_bits)
    q = q_w.to(torch.int64)
    res = torch.zeros((size_k // pack, size_n), dtype=torch.int64, device=q_w.device)
    for i in range(pack):
        res |= q[i::pack, :] << (num_bits * i)
    return _wrap_i32(res)


def gptq_unpack(packed: torch.Tensor, num_bits: int) -> torch.Tensor:
    """inverse of gptq_pack: [K/pack, N] int32 -> [K, N] int32."""
    pack = get_pack_factor(num_bits)
    p = packed.to(torch.int64) & 0xFFFFFFFF
    k = p.shape[0] * pack
    out = torch.zeros((k, p.shape[1]), dtype=torch.int32, device=packed.device)
    mask = (1 << num_bits) - 1
    for i in range(pack):
        out[i::pack, :] = ((p >> (num_bits * i)) & mask).to(torch.int32)
    return out


# ------------------------------------------------------------------------------------------
# Marlin permutations (marlin_perms.py:16-47)
# ------------------------------------------------------------------------------------------
def get_perms(num_bits: int):
    perm_list = []
    for i in range(32):
        perm1 = []
        col = i // 4
        for block in (0, 1):
            for row in (2 * (i % 4), 2 * (i % 4) + 1, 2 * (i % 4 + 4), 2 * (i % 4 + 4) + 1):
                perm1.append(16 * row + col + 8 * block)
        for j in range(4):
            perm_list.extend(p + 256 * j for p in perm1)
    perm = np.array(perm_list)
    interleave = np.array([0, 2, 4, 6, 1, 3, 5, 7]) if num_bits == 4 else np.array([0, 2, 1, 3])
    perm = perm.reshape(-1, len(interleave))[:, interleave].ravel()
    scale_perm = [i + 8 * j for i in range(8) for j in range(8)]
    scale_perm_single = [2 * i + j for i in range(4) for j in (0, 1, 8, 9, 16, 17, 24, 25)]
    return torch.from_numpy(perm), scale_perm, scale_perm_single


def marlin_permute_weights(q_w, size_k, size_n, perm, tile=MARLIN_TILE):  # marlin_utils.py:25-37
    q_w = q_w.reshape(size_k // tile, tile, size_n // tile, tile).permute(0, 2, 1, 3)
    q_w = q_w.reshape(size_k // tile, size_n * tile)
    return q_w.reshape(-1, perm.numel())[:, perm.to(q_w.device)].reshape(q_w.shape)


def marlin_weights(q_w, size_k, size_n, num_bits, perm):  # marlin_utils.py:40-58
    q_w = marlin_permute_weights(q_w, size_k, size_n, perm)
    pack = get_pack_factor(num_bits)
    q = q_w.to(torch.int64)
    packed = torch.zeros((q.shape[0], q.shape[1] // pack), dtype=torch.int64, device=q_w.device)
    for i in range(pack):
        packed |= q[:, i::pack] << (num_bits * i)
    return _wrap_i32(packed)


def marlin_permute_scales(s, size_k, size_n, group_size, scale_perm, scale_perm_single):
    # marlin_utils.py:61-69
    if group_size < size_k and group_size != -1:
        s = s.reshape(-1, len(scale_perm))[:, scale_perm]
    else:
        s = s.reshape(-1, len(scale_perm_single))[:, scale_perm_single]
    return s.reshape(-1, size_n).contiguous()


def marlin_quantize(w, num_bits, group_size, act_order, generator=None):  # marlin_utils.py:72-107
    size_k, size_n = w.shape
    if group_size == -1:
        group_size = size_k
    w_ref, q_w, s, g_idx, rand_perm = quantize_weights(w, num_bits, group_size, act_order, generator)
    sort_indices = torch.empty(0, dtype=torch.int32, device=w.device)
    if act_order:
        q_w, g_idx, sort_indices = sort_weights(q_w, g_idx)
    perm, sp, sps = get_perms(num_bits)
    marlin_q_w = marlin_weights(q_w, size_k, size_n, num_bits, perm)
    marlin_s = marlin_permute_scales(s, size_k, size_n, group_size, sp, sps)
    return w_ref, marlin_q_w, marlin_s, g_idx, sort_indices, rand_perm


def fp8_marlin_quantize(w: torch.Tensor):
    """fp8 weight-only Marlin operands as tests/kernels/test_marlin_gemm.py:256-283 of the reference
    builds them: per-tensor dynamic fp8 quantisation (fp8/common.cu:129-165: scale = absmax / 448),
    the fp8 bytes packed four k rows per int32 (marlin_utils.py:227-247 pack_fp8_to_int32) and
    repacked to the 8-bit Marlin layout (gptq_marlin_repack == marlin_weights on the same bytes,
    pinned by the repack golden test), the scale expanded to channel-wise and permuted
    (marlin_permute_scales, group_size -1).  -> (w_ref [K, N] = fl(fp8 value x scale) in w.dtype,
    marlin_q [K/16, 4N] int32, marlin_s [1, N])."""
    size_k, size_n = w.shape
    scale = (w.float().abs().max() / 448.0).clamp(min=1e-12)
    w8 = (w.float() / scale).clamp(-448.0, 448.0).to(torch.float8_e4m3fn)
    q = w8.view(torch.uint8).to(torch.int32)
    perm, sp, sps = get_perms(8)
    marlin_q = marlin_weights(q, size_k, size_n, 8, perm)
    s = scale.to(w.dtype).reshape(1, 1).repeat(1, size_n)
    marlin_s = marlin_permute_scales(s, size_k, size_n, -1, sp, sps)
    w_ref = (w8.float() * s.float()).to(w.dtype)  # exact product, one rounding
    return w_ref, marlin_q, marlin_s


def marlin_unpack(marlin_q_w: torch.Tensor, size_k: int, size_n: int, num_bits: int) -> torch.Tensor:
    """inverse of marlin_weights: Marlin tiles [K/16, N*16/pack] int32 -> [K, N] ints."""
    perm, _, _ = get_perms(num_bits)
    pack = get_pack_factor(num_bits)
    p = marlin_q_w.to(torch.int64) & 0xFFFFFFFF
    q = torch.zeros((p.shape[0], p.shape[1] * pack), dtype=torch.int32, device=marlin_q_w.device)
    mask = (1 << num_bits) - 1
    for i in range(pack):
        q[:, i::pack] = ((p >> (num_bits * i)) & mask).to(torch.int32)
    q = q.reshape(-1, perm.numel())
    inv = torch.empty_like(perm)
    inv[perm] = torch.arange(perm.numel())
    q = q[:, inv.to(q.device)].reshape(size_k // 16, size_n * 16)
    q = q.reshape(size_k // 16, size_n // 16, 16, 16).permute(0, 2, 1, 3).reshape(size_k, size_n)
    return q.contiguous()


def marlin_unpermute_scales(marlin_s, size_k, size_n, group_size, num_bits=4):
    _, sp, sps = get_perms(num_bits)
    use = sp if (group_size < size_k and group_size != -1) else sps
    inv = np.argsort(np.array(use))
    return marlin_s.reshape(-1, len(use))[:, inv].reshape(-1, size_n).contiguous()


# ------------------------------------------------------------------------------------------
# 2:4 sparse Marlin (marlin_24_perms.py:16-48, format_24.py, marlin_utils.py:110-204)
# ------------------------------------------------------------------------------------------
def get_perms_24(num_bits: int):
    perm_list = []
    for i in range(32):
        perm1 = []
        col = i // 4
        col_o = col // 2
        for block in (0, 1):
            for row in (2 * (i % 4), 2 * (i % 4) + 1, 2 * (i % 4 + 4), 2 * (i % 4 + 4) + 1):
                perm1.append(16 * row + col_o * 256 + 8 * (col % 2) + 4 * block)
        for j in range(4):
            perm_list.extend(p + j for p in perm1)
    perm = np.array(perm_list)
    interleave = np.array([0, 2, 4, 6, 1, 3, 5, 7]) if num_bits == 4 else np.array([0, 2, 1, 3])
    perm = perm.reshape(-1, len(interleave))[:, interleave].ravel()
    scale_perm = [i * 8 + j for i in range(8) for j in (0, 4, 1, 5, 2, 6, 3, 7)]
    scale_perm_single = list(range(64))
    return torch.from_numpy(perm), scale_perm, scale_perm_single


def mask_24(w_t: torch.Tensor) -> torch.Tensor:
    """keep the 2 largest |w| of every 4 consecutive elements of the flattened tensor
    (mask_creator, format_24.py:278-308)."""
    g = w_t.detach().abs().reshape(-1, 4)
    drop = torch.argsort(g, dim=1)[:, :2]
    m = torch.ones_like(g)
    m.scatter_(1, drop, 0)
    return m.reshape(w_t.shape)


def inject_24(w: torch.Tensor):  # marlin_utils.py:110-116 (sparsity runs along K)
    mask = mask_24(w.t().contiguous()).t().bool()
    return (mask * w).contiguous(), mask.contiguous()


def _meta_scatter_offsets(m: int, meta_ncols: int, device=None) -> torch.Tensor:
    """int16-metadata reordering of the CUTLASS sparse GEMM (format_24.py:22-49)."""
    r = torch.arange(m, device=device)[:, None].repeat(1, meta_ncols)
    c = torch.arange(meta_ncols, device=device).repeat(m, 1)
    gx, gy = 64, 32
    r = r // gx * gx + (r % 2) * 2 + (r % 8) // 4 + ((r % gy) % 4) // 2 * 32 + ((r % gx) // 8) * 4
    tr = ((r % 2 == 0) & (c % 2 == 1)).to(torch.int8)
    bl = ((r % 2 == 1) & (c % 2 == 0)).to(torch.int8)
    r = r + (tr - bl)
    c = c - (tr - bl)
    return ((c // 2) * m * 2 + r * 2 + (c % 2)).reshape(-1)


def compress_24(dense: torch.Tensor):
    """dense [m, k] int32 (2:4 along k) -> (values [m, k/2], meta int16 [m, k/16] reordered)
    (sparse_semi_structured_from_dense_cutlass, format_24.py:55-173, int16 branch)."""
    m, k = dense.shape
    assert m % 32 == 0 and k % 16 == 0
    d4 = dense.reshape(-1, k // 4, 4)
    m0, m1, m2, m3 = (d4 != 0).unbind(-1)
    e0, e1, e2 = m0 & m1, ~m0 & m1, ~m0 & ~m1
    idx0 = e1.to(torch.int64) | (e2.to(torch.int64) << 1)
    idx1 = (e0 | e2 | m3).to(torch.int64) | ((e1 | ~m1).to(torch.int64) << 1)
    s0 = d4.gather(-1, idx0.unsqueeze(-1))
    s1 = d4.gather(-1, idx1.unsqueeze(-1))
    sparse = torch.stack((s0, s1), dim=-1).reshape(m, k // 2)
    meta4 = (idx0 | (idx1 << 2)).reshape(-1, k // 16, 4).to(torch.int16)
    meta = meta4[:, :, 0] | (meta4[:, :, 1] << 4) | (meta4[:, :, 2] << 8) | (meta4[:, :, 3] << 12)
    out = meta.new_empty(m * (k // 16))
    out.scatter_(0, _meta_scatter_offsets(m, k // 16, dense.device), meta.reshape(-1))
    return sparse, out.reshape(m, k // 16)


def decompress_24(sparse: torch.Tensor, meta_reordered: torch.Tensor) -> torch.Tensor:
    """inverse of compress_24 (sparse_semi_structured_to_dense_cutlass, format_24.py:180-275)."""
    m, kh = sparse.shape
    ncols = meta_reordered.shape[1]
    meta = torch.gather(meta_reordered.reshape(-1), 0, _meta_scatter_offsets(m, ncols, sparse.device)).reshape(m, ncols)
    meta = meta.to(torch.int32) & 0xFFFF
    idx = torch.stack([(meta >> (2 * i)) & 3 for i in range(8)], dim=-1).reshape(-1)  # [m*kh]
    base = (torch.arange(m * kh // 2, device=sparse.device) * 4).reshape(-1, 1).repeat(1, 2).reshape(-1)
    dense = torch.zeros(m * 2 * kh, dtype=sparse.dtype, device=sparse.device)
    dense.scatter_(0, (idx.long() + base), sparse.reshape(-1))
    return dense.reshape(m, 2 * kh)


def compress_quantized_24_weight(q_24: torch.Tensor, size_k: int, size_n: int, num_bits: int):
    # marlin_utils.py:144-164
    zp = (1 << num_bits) // 2
    comp, meta = compress_24((q_24 - zp).t().contiguous())
    comp = comp.t().contiguous() + zp
    # resize_ without moving data: [n, k/16] -> [k/32, 2n]
    meta = meta.reshape(-1).reshape(meta.shape[1] // 2, meta.shape[0] * 2)
    return comp, meta


def marlin_24_quantize(w: torch.Tensor, num_bits: int, group_size: int):  # marlin_utils.py:167-204
    size_k, size_n = w.shape
    if group_size == -1:
        group_size = size_k
    w_24, _ = inject_24(w)
    w_24_ref, q_w_24, s, _, _ = quantize_weights(w_24, num_bits, group_size, False)
    q_comp, meta = compress_quantized_24_weight(q_w_24, size_k, size_n, num_bits)
    perm, sp, sps = get_perms_24(num_bits)
    marlin_q = marlin_weights(q_comp, size_k // 2, size_n, num_bits, perm)
    marlin_s = marlin_permute_scales(s, size_k, size_n, group_size, sp, sps)
    return w_24_ref, marlin_q, meta, marlin_s


def compute_max_diff(output, output_ref):  # marlin_utils.py:207-209
    return torch.mean(torch.abs(output - output_ref)) / torch.mean(torch.abs(output_ref))


# ------------------------------------------------------------------------------------------
# W8A8 (tests/kernels/test_cutlass.py:36-47) and activation quantisation
# (tests/kernels/test_int8_quant.py:25-71; csrc/quantization/fp8/common.cu:18-24,78-105)
# ------------------------------------------------------------------------------------------
def baseline_scaled_mm(a, b, scale_a, scale_b, out_dtype, bias=None):
    out = (scale_a * (scale_b * torch.mm(a.to(torch.float32), b.to(torch.float32)))).to(out_dtype)
    if bias is not None:
        out = out + bias
    return out


def scaled_int8_quant_ref(x: torch.Tensor, scale: torch.Tensor | None = None):
    """dynamic (scale None): per-token absmax/127; static: given per-tensor scale."""
    xf = x.float()
    if scale is None:
        scales = xf.abs().max(dim=-1, keepdim=True)[0] / 127.0
        q = torch.round(xf / scales).clamp(-128, 127).to(torch.int8)
        return q, scales
    q = torch.round(xf / scale.float()).clamp(-128, 127).to(torch.int8)
    return q, scale


def scaled_fp8_quant_ref(x: torch.Tensor, scale: torch.Tensor | None = None):
    """per-tensor e4m3; dynamic scale = absmax/448 (fp8/common.cu:55-60), value = clamp(x/scale)."""
    xf = x.float()
    if scale is None:
        scale = (xf.abs().max() / 448.0).reshape(1)
    q = (xf * (1.0 / scale.float())).clamp(-448.0, 448.0).to(torch.float8_e4m3fn)
    return q, scale


# ---- AWQ ------------------------------------------------------------------
# Format restated from the reference's device code, the only place it is written down:
# csrc/quantization/awq/dequantize.cuh:17-101 (a 32-bit word holds eight 4-bit values of ONE k row;
# the lop3 pairs nibbles (0,4) (1,5) (2,6) (3,7) and the results are eight consecutive output
# channels, i.e. nibble j holds channel AWQ_ORDER[j]) and gemm_kernels.cu:369-435
# (dequantize_weights: w = (q - z) * s in fp16, sub.f16x2 then fma.rn.f16x2 with 0: (q - z) is an
# exact integer, so the product is rounded once).  No test or golden vector for AWQ exists in the
# reference tree; the GPU tests pin this restatement against the reference's own kernels compiled
# for sm_100 (oracle/_ref/_C_refcuda.so).
AWQ_ORDER = [0, 2, 4, 6, 1, 3, 5, 7]


def awq_pack(q: torch.Tensor) -> torch.Tensor:
    """q [R, C] integers in [0, 16) -> int32 [R, C / 8]: nibble j of a word = column 8 c + AWQ_ORDER[j]."""
    r, c = q.shape
    assert c % 8 == 0
    q = q.to(torch.int64).view(r, c // 8, 8)
    word = torch.zeros(r, c // 8, dtype=torch.int64)
    for j, col in enumerate(AWQ_ORDER):
        word |= q[:, :, col] << (4 * j)
    return _wrap_i32(word)


def awq_unpack(packed: torch.Tensor) -> torch.Tensor:
    """int32 [R, C / 8] -> int64 [R, C] in [0, 16)."""
    r, c8 = packed.shape
    w = packed.to(torch.int64) & 0xFFFFFFFF
    out = torch.zeros(r, c8, 8, dtype=torch.int64)
    for j, col in enumerate(AWQ_ORDER):
        out[:, :, col] = (w >> (4 * j)) & 0xF
    return out.view(r, c8 * 8)


def awq_quantize(w: torch.Tensor, group_size: int):
    """Asymmetric 4-bit quantisation of w [K, N] per (group of K, column), as AutoAWQ checkpoints
    hold it -> (w_ref fp16 [K, N], qweight int32 [K, N/8], scales fp16 [K/G, N], qzeros int32 [K/G, N/8])."""
    k, n = w.shape
    g = k // group_size
    wg = w.float().view(g, group_size, n)
    lo, hi = wg.amin(1), wg.amax(1)
    scales = ((hi - lo).clamp(min=1e-5) / 15).half()
    zeros = (-lo / scales.float()).round().clamp(0, 15).to(torch.int64)
    q = (wg / scales.float().unsqueeze(1) + zeros.unsqueeze(1)).round().clamp(0, 15).to(torch.int64).view(k, n)
    qweight, qzeros = awq_pack(q), awq_pack(zeros)
    return awq_dequantize_ref(qweight, scales, qzeros), qweight, scales, qzeros


def awq_dequantize_ref(qweight: torch.Tensor, scales: torch.Tensor, qzeros: torch.Tensor) -> torch.Tensor:
    """-> fp16 [K, N]: (q - z) exactly, times the group scale, rounded to fp16 once."""
    k = qweight.shape[0]
    group_size = k // scales.shape[0]
    q = awq_unpack(qweight)
    z = awq_unpack(qzeros).repeat_interleave(group_size, 0)
    s = scales.float().repeat_interleave(group_size, 0)
    return ((q - z).float() * s).half()  # fp32 product of an integer |.| <= 15 and an fp16 value is exact
