#!/usr/bin/env python
"""Generates tests/golden/*.npz by running the REFERENCE's own Python utilities (imported by file
path from /root/reference; container only) on seeded inputs.  The committed fixtures pin
oracle/quant_ref.py and the CUDA kernels; nothing at test time reads /root/reference.

    python tests/golden/make_golden.py
"""
import importlib.util
import os
import sys
import types

import numpy as np
import torch

REF = os.environ.get("B200_REFERENCE", "/root/reference")
UTILS = os.path.join(REF, "vllm/model_executor/layers/quantization/utils")
OUT = os.path.dirname(os.path.abspath(__file__))


def _load_reference_utils():
    """Registers stub `vllm.*` packages so the reference utils import each other by their own
    absolute names, without importing the reference's (or the image's) `vllm` package."""
    pkgs = ["vllm", "vllm.model_executor", "vllm.model_executor.layers",
            "vllm.model_executor.layers.quantization", "vllm.model_executor.layers.quantization.utils",
            "vllm.platforms"]
    saved = {k: sys.modules.get(k) for k in pkgs}
    for p in pkgs:
        m = types.ModuleType(p)
        m.__path__ = []
        sys.modules[p] = m
    sys.modules["vllm.platforms"].current_platform = types.SimpleNamespace(
        get_device_capability=lambda: (10, 0))
    mods = {}
    for name in ["quant_utils", "format_24", "marlin_perms", "marlin_24_perms", "marlin_utils"]:
        full = f"vllm.model_executor.layers.quantization.utils.{name}"
        spec = importlib.util.spec_from_file_location(full, os.path.join(UTILS, name + ".py"))
        mod = importlib.util.module_from_spec(spec)
        sys.modules[full] = mod
        spec.loader.exec_module(mod)
        mods[name] = mod
    return mods, saved


def main():
    mods, _ = _load_reference_utils()
    qu, mu, f24 = mods["quant_utils"], mods["marlin_utils"], mods["format_24"]
    # inject_24 hard-codes .cuda() (marlin_utils.py:112): same statement on CPU
    def inject_24_cpu(w, size_k, size_n):
        mask = f24.mask_creator(w.t()).t().bool()
        return (mask * w).contiguous(), mask.contiguous()
    mu.inject_24 = inject_24_cpu

    # ---- dense Marlin: quantize / pack / permute (marlin_utils.py:72-107) ----
    cases = {}
    for (k, n, bits, group, act) in [(128, 64, 4, -1, False), (256, 128, 4, 128, False),
                                     (256, 256, 4, 64, False), (256, 128, 4, 32, True),
                                     (128, 128, 8, -1, False), (256, 64, 8, 128, False),
                                     (256, 128, 8, 64, True)]:
        torch.manual_seed(1000 + k + n + bits + (group % 97))
        w = torch.randn(k, n, dtype=torch.half)
        w_ref, mq, ms, g_idx, sort_idx, rand_perm = mu.marlin_quantize(w, bits, group, act)
        # the GPTQ-packed tensor gptq_marlin_repack consumes (test_marlin_gemm.py:62-115)
        gs = k if group == -1 else group
        _, q_w, s, g2, rp2 = (None, None, None, None, None)
        key = f"dense_k{k}_n{n}_b{bits}_g{group}_a{int(act)}"
        cases[key + "_w"] = w.numpy()
        cases[key + "_w_ref"] = w_ref.numpy()
        cases[key + "_marlin_q"] = mq.numpy()
        cases[key + "_marlin_s"] = ms.numpy()
        cases[key + "_g_idx"] = g_idx.numpy()
        cases[key + "_sort_idx"] = sort_idx.numpy()
        cases[key + "_rand_perm"] = rand_perm.numpy()
    # repack path: quantize -> (sort) -> gptq_pack ; expected = marlin_weights (test_marlin_gemm.py:76-107)
    for (k, n, bits, group, act) in [(128, 64, 4, 128, False), (256, 128, 4, 64, True),
                                     (128, 64, 8, 128, False), (256, 128, 8, 32, True)]:
        torch.manual_seed(2000 + k + n + bits)
        w = torch.randn(k, n, dtype=torch.half)
        w_ref, q_w, s, g_idx, rand_perm = qu.quantize_weights(w, bits, group, act)
        q_pack = qu.gptq_pack(q_w, bits, k, n)
        sort_idx = torch.empty(0, dtype=torch.int32)
        if act:
            q_w, g_idx, sort_idx = qu.sort_weights(q_w, g_idx)
        expect = mu.marlin_weights(q_w, k, n, bits, mods["marlin_perms"].marlin_perm[bits])
        key = f"repack_k{k}_n{n}_b{bits}_a{int(act)}"
        cases[key + "_gptq"] = q_pack.numpy()
        cases[key + "_sort_idx"] = sort_idx.numpy()
        cases[key + "_expect"] = expect.numpy()
    np.savez_compressed(os.path.join(OUT, "marlin_dense.npz"), **cases)

    # ---- 2:4 sparse Marlin (marlin_utils.py:167-204) ----
    cases = {}
    for (k, n, bits, group) in [(128, 128, 4, -1), (256, 128, 4, 128), (128, 256, 8, -1),
                                (256, 128, 8, 128)]:
        torch.manual_seed(3000 + k + n + bits)
        w = torch.randn(k, n, dtype=torch.half)
        w_ref, mq, meta, ms = mu.marlin_24_quantize(w, bits, group)
        key = f"s24_k{k}_n{n}_b{bits}_g{group}"
        cases[key + "_w"] = w.numpy()
        cases[key + "_w_ref"] = w_ref.numpy()
        cases[key + "_marlin_q"] = mq.numpy()
        cases[key + "_meta"] = meta.numpy()
        cases[key + "_marlin_s"] = ms.numpy()
    np.savez_compressed(os.path.join(OUT, "marlin_24.npz"), **cases)
    print("wrote", os.listdir(OUT))


if __name__ == "__main__":
    main()
