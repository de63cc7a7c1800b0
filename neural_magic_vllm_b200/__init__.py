"""B200-native kernels behind nm-vllm's custom-op surface.

Importing this package loads ``_C.so`` (built in-tree by ``build.py``), whose static
initialisers register the reference's op schemas under ``torch.ops._C``,
``torch.ops._C_cache_ops`` and ``torch.ops._C_cuda_utils`` (reference:
``csrc/torch_bindings.cpp``; loaded there by ``import vllm._C``, ``vllm/_custom_ops.py:11-14``).

There is no fallback: if the extension is missing, the import raises.
"""
from __future__ import annotations

import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libb200vllm.so")
EXT_PATH = os.path.join(_HERE, "_C.so")

_loaded = False


def load_extension() -> None:
    """Registers the torch ops. Raises (never falls back) when the build is missing."""
    global _loaded
    if _loaded:
        return
    if not (os.path.exists(LIB_PATH) and os.path.exists(EXT_PATH)):
        raise ImportError(
            "neural_magic_vllm_b200: the CUDA extension is not built "
            f"({LIB_PATH} / {EXT_PATH} missing). Run `python neural_magic_vllm_b200/build.py` "
            "(needs nvcc; no GPU required). There is no CPU or eager fallback.")
    import torch
    torch.ops.load_library(EXT_PATH)
    _loaded = True


def c_abi() -> ctypes.CDLL:
    """The raw C-ABI library (include/b200_vllm.h), for callers that bypass torch."""
    return ctypes.CDLL(LIB_PATH)


load_extension()

from . import _custom_ops as ops  # noqa: E402  (mirrors `from vllm import _custom_ops as ops`)

__all__ = ["ops", "load_extension", "c_abi", "LIB_PATH", "EXT_PATH"]
