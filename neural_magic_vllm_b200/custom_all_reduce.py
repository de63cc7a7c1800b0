"""One-shot all-reduce over NVLink peer memory -- host side.

Mirror of the reference's ``CustomAllreduce``
(``vllm/distributed/device_communicators/custom_all_reduce.py:38-272``): same constructor shape,
``should_custom_ar`` / ``custom_all_reduce`` / ``close`` with the same meaning (``custom_all_reduce``
returns the reduced tensor, or ``None`` when the tensor does not qualify and the caller must use
NCCL -- ``parallel_state.py`` of the reference does exactly that).

What differs, by design: there is no buffer registration (``register_buffer`` /
``register_graph_buffers``): the kernel copies its input into a per-rank IPC staging buffer itself,
so the same call works eagerly and under CUDA-graph capture, and the epoch that separates
consecutive calls lives on the device.  IPC handles are exchanged once, through
``torch.distributed.all_gather_object`` on the given group.

Calls ``libb200vllm.so`` through its C ABI (``include/b200_vllm.h``: ``b200_ar_*``,
``b200_all_reduce``); there is no fallback inside this class.
"""
from __future__ import annotations

import ctypes
from typing import List, Optional, Union

import torch
import torch.distributed as dist

from . import c_abi

_DTYPE = {torch.float16: 0, torch.bfloat16: 1, torch.float32: 2}


def _lib():
    lib = c_abi()
    lib.b200_last_error.restype = ctypes.c_char_p
    lib.b200_ar_alloc.argtypes = [ctypes.c_int64, ctypes.POINTER(ctypes.c_void_p), ctypes.c_char_p]
    lib.b200_ar_open.argtypes = [ctypes.c_char_p, ctypes.POINTER(ctypes.c_void_p)]
    lib.b200_ar_close.argtypes = [ctypes.c_void_p]
    lib.b200_ar_free.argtypes = [ctypes.c_void_p]
    lib.b200_all_reduce.argtypes = [ctypes.POINTER(ctypes.c_void_p), ctypes.c_int32, ctypes.c_int32,
                                    ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_int32,
                                    ctypes.c_int64, ctypes.c_void_p]
    lib.b200_all_reduce_add_rms_norm.argtypes = [ctypes.POINTER(ctypes.c_void_p), ctypes.c_int32, ctypes.c_int32,
                                                 ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                                 ctypes.c_float, ctypes.c_int32, ctypes.c_int32,
                                                 ctypes.c_int32, ctypes.c_int64, ctypes.c_void_p]
    return lib


class CustomAllreduce:
    _SUPPORTED_WORLD_SIZES = [2, 4, 6, 8]

    def __init__(self, group: Optional[dist.ProcessGroup] = None,
                 device: Union[int, str, torch.device, None] = None,
                 max_size: int = 2 * 1024 * 1024) -> None:
        # default max_size = the flag-in-data ("LL") range, where the kernel beats NCCL on B200
        # (profiles/r1c_all_reduce.md); larger max_size also enables the staged one-shot kernel
        self.disabled = True
        self.group = group
        self.rank = dist.get_rank(group)
        self.world_size = dist.get_world_size(group)
        if self.world_size == 1:
            return  # nothing to reduce (custom_all_reduce.py:66-68)
        if self.world_size not in self._SUPPORTED_WORLD_SIZES:
            return  # the caller keeps NCCL (custom_all_reduce.py:70-77)
        if device is None:
            device = torch.device("cuda", torch.cuda.current_device())
        elif isinstance(device, int):
            device = torch.device("cuda", device)
        elif isinstance(device, str):
            device = torch.device(device)
        self.device = device
        self.max_size = int(max_size)
        # A one-shot all-reduce moves (world - 1) x the message per rank, so the size up to which it
        # beats NCCL shrinks with the world size (the reference makes the same cut:
        # custom_all_reduce.cu:37-48 -- any size at world 2, < 512 KiB beyond).  Measured on B200
        # (profiles/r1c_all_reduce.md): world 2 wins through 2 MiB, world 4 through 512 KiB-1 MiB.
        # Round 2: from world 4 up, messages >= 128 KiB take the two-shot flag-in-data path inside the
        # same kernels (2 (W - 1) / W of the message per rank instead of W - 1 copies), which wins
        # through the 2 MiB the flag-in-data receive areas hold (world 8: 2 MiB 18.7 us vs NCCL 34.0;
        # profiles/r2c_all_reduce.md); above that the staged one-shot kernel loses to NCCL.
        self.one_shot_limit = self.max_size if self.world_size == 2 else 2 * 1024 * 1024
        self.one_shot_limit = min(self.one_shot_limit, self.max_size)
        self._lib = _lib()
        self._own = ctypes.c_void_p()
        handle = ctypes.create_string_buffer(64)
        with torch.cuda.device(device):
            self._check(self._lib.b200_ar_alloc(self.max_size, ctypes.byref(self._own), handle))
            handles: List[Optional[bytes]] = [None] * self.world_size
            dist.all_gather_object(handles, handle.raw, group=group)
            self._bases = (ctypes.c_void_p * self.world_size)()
            self._opened: List[ctypes.c_void_p] = []
            for r, h in enumerate(handles):
                if r == self.rank:
                    self._bases[r] = self._own.value
                else:
                    p = ctypes.c_void_p()
                    self._check(self._lib.b200_ar_open(h, ctypes.byref(p)))
                    self._opened.append(p)
                    self._bases[r] = p.value
        dist.barrier(group=group)  # every rank has mapped every buffer before the first call
        self.disabled = False

    def _check(self, rc: int) -> None:
        if rc != 0:
            raise RuntimeError(self._lib.b200_last_error().decode())

    def should_custom_ar(self, inp: torch.Tensor) -> bool:
        """custom_all_reduce.cu:37-48 of the reference: 16-byte multiple, within max_size,
        (weakly) contiguous, a supported dtype."""
        if self.disabled:
            return False
        nbytes = inp.numel() * inp.element_size()
        return (inp.is_cuda and inp.dtype in _DTYPE and nbytes % 16 == 0 and nbytes <= self.one_shot_limit
                and inp.is_contiguous() and inp.data_ptr() % 16 == 0)

    def all_reduce(self, inp: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        if out is None:
            out = torch.empty_like(inp)
        stream = torch.cuda.current_stream(inp.device).cuda_stream
        self._check(self._lib.b200_all_reduce(self._bases, self.rank, self.world_size, inp.data_ptr(),
                                              out.data_ptr(), inp.numel(), _DTYPE[inp.dtype],
                                              self.max_size, stream))
        return out

    def custom_all_reduce(self, input: torch.Tensor) -> Optional[torch.Tensor]:
        """The reduced tensor, or None when the caller must fall back to NCCL."""
        if self.disabled or not self.should_custom_ar(input):
            return None
        return self.all_reduce(input)

    def all_reduce_add_rms_norm(self, input: torch.Tensor, residual: torch.Tensor, weight: torch.Tensor,
                                epsilon: float) -> bool:
        """B200 extension: ``all_reduce(input)`` followed by ``fused_add_rms_norm(input, residual,
        weight, epsilon)`` in ONE kernel, in place (no reference op; the pair the reference issues
        after every row-parallel linear).  Returns False -- nothing done -- when the tensors do not
        qualify; the caller then issues the two ops."""
        if self.disabled or input.dim() != 2 or not self.should_custom_ar(input):
            return False
        tokens, hidden = input.shape
        row_bytes = hidden * input.element_size()
        # one CTA per row over a fixed 72-CTA grid: past ~2 rows per CTA the two-kernel path wins
        # (measured at TP2, batch 256: profiles/r1c_decode_tokens.md)
        if tokens > 144:
            return False
        # at world 6/8 a lone row (one CTA pushing to 7 peers) loses to the two-kernel path, which
        # spreads a small message over more CTAs; from a few dozen rows the fused kernel wins again
        # (TP8: batch 1/8 551/4082 vs 519/3906 tokens/s unfused vs fused, batch 64 21653 vs 22329)
        if self.world_size > 4 and tokens < 32:
            return False
        if (row_bytes % 8 or row_bytes > 32 * 1024 or not residual.is_contiguous() or not weight.is_contiguous()
                or residual.shape != input.shape or residual.dtype != input.dtype or weight.dtype != input.dtype
                or weight.numel() != hidden):
            return False
        stream = torch.cuda.current_stream(input.device).cuda_stream
        self._check(self._lib.b200_all_reduce_add_rms_norm(
            self._bases, self.rank, self.world_size, input.data_ptr(), residual.data_ptr(), weight.data_ptr(),
            float(epsilon), tokens, hidden, _DTYPE[input.dtype], self.max_size, stream))
        return True

    def close(self) -> None:
        if not self.disabled:
            self.disabled = True
            torch.cuda.synchronize(self.device)
            for p in self._opened:
                self._lib.b200_ar_close(p)
            self._lib.b200_ar_free(self._own)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
