"""CPU restatement of the reference's layernorm / rotary / activation CUDA kernels.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``): imported by ``tests/`` and
``__graft_entry__.smoke()`` as the checker, never by the product path.

Each function follows the reference CUDA kernel operator by operator.  The reference computes in
``scalar_t`` (c10::Half / c10::BFloat16), whose every operator is "fp32 operation, round to
scalar_t" -- which is exactly what torch's CPU half/bfloat16 tensor arithmetic does, so plain tensor
expressions in the input dtype reproduce the rounding sequence.

Pinned (``tests/test_oracle_cpu.py``) against the reference's own CPU kernels
(``csrc/cpu/{layernorm,pos_encoding,activation}.cpp``, compiled in place into ``oracle/_ref``) for
fp32 and bf16 -- the dtypes that backend supports.
"""
from __future__ import annotations

import math

import torch


def rms_norm(x: torch.Tensor, weight: torch.Tensor, eps: float) -> torch.Tensor:
    """csrc/layernorm_kernels.cu:21-45: out = ((scalar_t)(x * rstd)) * weight."""
    xf = x.float()
    var = (xf * xf).sum(-1, keepdim=True) / x.shape[-1]
    rstd = torch.rsqrt(var + eps)
    return (xf * rstd).to(x.dtype) * weight


def fused_add_rms_norm(x: torch.Tensor, residual: torch.Tensor, weight: torch.Tensor, eps: float):
    """csrc/layernorm_kernels.cu:199-287: z = x + residual (in scalar_t); returns (out, z)."""
    z = x + residual
    return rms_norm(z, weight, eps), z


def rotary_embedding(positions, query, key, head_size: int, cos_sin_cache, is_neox: bool,
                     offsets=None):
    """csrc/pos_encoding_kernels.cu:10-125.  query/key [tokens, heads*head_size]; returns rotated
    copies.  ``x*cos - y*sin`` / ``y*cos + x*sin`` with every operator rounded to the dtype."""
    rot_dim = cos_sin_cache.shape[1]
    embed = rot_dim // 2
    pos = positions.reshape(-1)
    if offsets is not None:
        pos = pos + offsets.reshape(-1)
    cs = cos_sin_cache[pos]  # [tokens, rot_dim]
    cos, sin = cs[:, None, :embed], cs[:, None, embed:]

    def rot(t):
        tokens = t.shape[0]
        t = t.reshape(tokens, -1, head_size).clone()
        if is_neox:
            x, y = t[..., :embed], t[..., embed:rot_dim]
        else:
            x, y = t[..., 0:rot_dim:2], t[..., 1:rot_dim:2]
        xo = x * cos - y * sin
        yo = y * cos + x * sin
        if is_neox:
            t[..., :embed], t[..., embed:rot_dim] = xo, yo
        else:
            t[..., 0:rot_dim:2], t[..., 1:rot_dim:2] = xo, yo
        return t.reshape(tokens, -1)

    return rot(query), rot(key)


def _silu(x):  # activation_kernels.cu:26-30
    f = x.float()
    return (f / (1.0 + torch.exp(-f))).to(x.dtype)


def _gelu(x):  # :32-40
    f = x.float()
    return (f * 0.5 * (1.0 + torch.erf(f * math.sqrt(0.5)))).to(x.dtype)


def _gelu_tanh(x):  # :42-53
    f = x.float()
    beta = torch.tensor(math.sqrt(2.0), dtype=torch.float32) * torch.tensor(2.0 / math.sqrt(math.pi), dtype=torch.float32) * 0.5
    inner = beta * (f + 0.044715 * (f * f * f))
    return (0.5 * f * (1.0 + torch.tanh(inner))).to(x.dtype)


def act_and_mul(x: torch.Tensor, act: str) -> torch.Tensor:
    """csrc/activation_kernels.cu:12-24: out = ACT(x[..., :d]) * x[..., d:] (product in scalar_t)."""
    d = x.shape[-1] // 2
    fn = {"silu": _silu, "gelu": _gelu, "gelu_tanh": _gelu_tanh}[act]
    return fn(x[..., :d]) * x[..., d:]


def gelu_new(x):  # :118-123
    dt = x.dtype
    x3 = (x * x * x).float()
    a = x + (0.044715 * x3).to(dt)
    t = torch.tanh((0.79788456 * a.float()).to(dt).float()).to(dt)
    return (torch.tensor(0.5, dtype=dt) * x) * (torch.tensor(1.0, dtype=dt) + t)


def gelu_fast(x):  # :125-131
    dt = x.dtype
    f = x.float()
    a = (f * 0.79788456).to(dt)
    b = (0.044715 * f).to(dt) * x
    t = torch.tanh((a * (torch.tensor(1.0, dtype=dt) + b)).float()).to(dt)
    return (torch.tensor(0.5, dtype=dt) * x) * (torch.tensor(1.0, dtype=dt) + t)


def gelu_quick(x):  # :133-137
    f = x.float()
    return (f / (1.0 + torch.exp(-1.702 * f))).to(x.dtype)
