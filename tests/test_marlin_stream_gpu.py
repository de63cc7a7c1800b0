"""gptq_marlin_gemm through the stream-K persistent decode kernel (csrc/marlin_gemm_stream.cuh).

The dispatch heuristic sends only wide layers (>= num_sms / 2 tiles of 256 columns) to this kernel,
so the reference's parameter space (tests/kernels/test_marlin_gemm.py:126-178) is re-run here with
the kernel forced through the C-ABI test hook, plus the shapes that exercise what is specific to it:
ranges that split tiles between CTAs (publisher / fixer pieces), tiles with more than two
contributors, n tails (N % 256 != 0), single-unit grids, channel-wise scales, 8-bit, bf16,
determinism, the lock workspace handed back zeroed, and CUDA-graph replay.  Oracle: a @ w_ref with
w_ref from the restated marlin_quantize (oracle/quant_ref.py, pinned to the reference's golden
vectors in tests/test_oracle_cpu.py)."""
import ctypes
import os

import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DEV = "cuda"


@pytest.fixture(scope="module")
def ops():
    import neural_magic_vllm_b200 as pkg
    return pkg.ops


@pytest.fixture()
def stream_forced():
    lib = ctypes.CDLL(os.path.join(ROOT, "neural_magic_vllm_b200", "libb200vllm.so"))
    lib.b200_debug_set_marlin_decode_kernel.argtypes = [ctypes.c_int32]
    lib.b200_debug_set_marlin_decode_kernel(2)
    yield lib
    lib.b200_debug_set_marlin_decode_kernel(-1)


def _ws(n):
    return torch.zeros(n // 64 * 16, dtype=torch.int32, device=DEV)


def _case(ops, k, n, m, bits=4, group=128, dtype=torch.half, seed=0):
    from oracle import quant_ref as qr
    torch.manual_seed(seed)
    a = torch.randn(m, k, dtype=dtype, device=DEV)
    w = torch.randn(k, n, dtype=dtype)
    w_ref, mq, ms, g_idx, sort_idx, _ = qr.marlin_quantize(w, bits, group, False)
    ws = _ws(n)
    out = ops.gptq_marlin_gemm(a, mq.to(DEV), ms.to(dtype).to(DEV), g_idx.to(DEV), sort_idx.to(DEV), ws, bits,
                               m, n, k, True)
    assert int(ws.abs().max()) == 0, "lock workspace must come back zeroed (gptq_marlin.cu:329-332)"
    ref = torch.matmul(a.float(), w_ref.to(DEV).float())
    assert qr.compute_max_diff(out.float(), ref) < 0.04  # the reference's bar
    err = (out.float() - ref).abs()
    tol = 1e-2 * ref.abs() + 2e-3 * ref.abs().max() + 1e-3  # DESIGN.md section 4.3 (fp16 output rounding)
    if dtype == torch.bfloat16:
        tol = tol * 8
    assert bool((err <= tol).all()), f"max err {err.max()} vs scale {ref.abs().max()}"
    return out


# (k, n): one unit; one tile with several k blocks; several tiles; n tails of 64 / 128 / 192 columns;
# long k (a tile shared by many CTAs); wide (many tiles per CTA); k not a multiple of the group
SHAPES = [(64, 64), (128, 256), (512, 256), (1024, 320), (256, 896), (448, 1216), (8192, 256), (4096, 1024),
          (128, 9728), (192, 19200)]


@pytest.mark.parametrize("k,n", SHAPES)
@pytest.mark.parametrize("m", [1, 5, 16, 17, 33, 64])
def test_stream_kernel_shapes(ops, stream_forced, k, n, m):
    group = 64 if k % 128 else 128
    _case(ops, k, n, m, group=group)


@pytest.mark.parametrize("bits", [4, 8])
@pytest.mark.parametrize("group", [-1, 64, 128])
@pytest.mark.parametrize("dtype", [torch.half, torch.bfloat16])
@pytest.mark.parametrize("m", [3, 48])
def test_stream_kernel_formats(ops, stream_forced, bits, group, dtype, m):
    _case(ops, 1024, 768, m, bits=bits, group=group, dtype=dtype)


def test_stream_kernel_reference_matrix(ops, stream_forced):
    # the thinned reference matrix of tests/test_quant_gpu.py at the decode sizes, forced through
    # the stream-K kernel (act-order and groups of 32 are not its cases: they take the other kernels)
    for bits in (4, 8):
        for group in (-1, 64, 128):
            for n_chunk in (64, 128, 256):
                _case(ops, 128, n_chunk, 1, bits=bits, group=group)
            _case(ops, 128 * 8, 128 * 4, 1, bits=bits, group=group)
            _case(ops, 128 * 5, 64 * 7, 1, bits=bits, group=group)
    for mf, nf, kf in ((13, 17, 67), (26, 37, 13), (67, 13, 11)):
        if mf <= 64:
            _case(ops, 128 * kf, 256 * nf, mf)


def test_stream_kernel_identity_a_is_the_dequantised_tile(ops, stream_forced):
    # a = I_K: every output element is one dequantised weight, rounded once -> must equal w_ref
    # bit for bit (the in-GEMM dequant path: LDS -> lop3/hsub2/hfma2 -> hmul2 -> tcgen05.st -> MMA)
    from oracle import quant_ref as qr
    torch.manual_seed(1)
    for k, bits, group in ((64, 4, 64), (64, 8, -1), (64, 4, -1)):
        n = 512
        w = torch.randn(k, n, dtype=torch.half)
        w_ref, mq, ms, g_idx, sort_idx, _ = qr.marlin_quantize(w, bits, group, False)
        a = torch.eye(k, dtype=torch.half, device=DEV)
        out = ops.gptq_marlin_gemm(a, mq.to(DEV), ms.to(DEV), g_idx.to(DEV), sort_idx.to(DEV), _ws(n), bits, k, n,
                                   k, True)
        assert torch.equal(out.cpu(), w_ref), (k, bits, group)


def test_stream_kernel_wide_layer_default_dispatch(ops):
    # no hook: a layer with >= num_sms / 2 tiles takes the stream-K kernel by itself
    _case(ops, 256, 19200, 7)
    _case(ops, 256, 19200 + 64, 64)


def test_stream_kernel_deterministic_and_graph_replay(ops, stream_forced):
    from oracle import quant_ref as qr
    torch.manual_seed(5)
    k, n, m = 2048, 1024, 9
    a = torch.randn(m, k, dtype=torch.half, device=DEV)
    w = torch.randn(k, n, dtype=torch.half)
    _, mq, ms, g_idx, sort_idx, _ = qr.marlin_quantize(w, 4, 128, False)
    mq, ms, g_idx, sort_idx = mq.to(DEV), ms.to(DEV), g_idx.to(DEV), sort_idx.to(DEV)
    ws = _ws(n)
    first = ops.gptq_marlin_gemm(a, mq, ms, g_idx, sort_idx, ws, 4, m, n, k, True)
    for _ in range(5):
        again = ops.gptq_marlin_gemm(a, mq, ms, g_idx, sort_idx, ws, 4, m, n, k, True)
        assert torch.equal(first, again), "split-K fix-up order must not depend on arrival order"
    torch.cuda.synchronize()
    stream = torch.cuda.Stream()
    with torch.cuda.stream(stream):
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            outs = [ops.gptq_marlin_gemm(a, mq, ms, g_idx, sort_idx, ws, 4, m, n, k, True) for _ in range(4)]
    for o in outs:
        o.zero_()
    g.replay()
    g.replay()
    torch.cuda.synchronize()
    for o in outs:
        assert torch.equal(o, first)
    assert int(ws.abs().max()) == 0
