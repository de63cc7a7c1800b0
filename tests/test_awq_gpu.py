"""awq_dequantize / awq_gemm against the oracle restatement of the AWQ format
(oracle/quant_ref.py: awq_pack / awq_dequantize_ref) and against the reference's OWN kernels
(csrc/quantization/awq/gemm_kernels.cu compiled for sm_100 into oracle/_ref/_C_refcuda.so) on
identical inputs.  The reference holds no AWQ test or golden vector; the pin is its kernel.
Bars: dequantised weights bit-exact; GEMM rtol 1e-2 with an atol of two fp16 ulps of the largest
output (the reference accumulates its split-K slices in fp16)."""
import pytest
import torch

from oracle import quant_ref as qr
from oracle import ref_cuda

pytestmark = pytest.mark.gpu
DEV = "cuda"


@pytest.fixture(scope="module")
def ops():
    import neural_magic_vllm_b200 as pkg
    return pkg.ops


def _weights(k, n, g, seed):
    torch.manual_seed(seed)
    w = torch.randn(k, n) * 0.05
    return qr.awq_quantize(w, g)


@pytest.mark.parametrize("k,n,g", [(128, 64, 128), (256, 192, 32), (512, 128, 64), (4096, 512, 128), (384, 256, 96)])
def test_awq_dequantize_bit_exact(ops, k, n, g):
    w_ref, qw, s, qz = _weights(k, n, g, seed=k + n)
    out = ops.awq_dequantize(qw.to(DEV), s.to(DEV), qz.to(DEV), 0, 0, 0)
    assert out.dtype == torch.half and tuple(out.shape) == (k, n)
    assert torch.equal(out.cpu(), w_ref)


def test_awq_dequantize_extreme_nibbles(ops):
    # every (q, z) pair, including q - z = +-15, at a scale whose product needs the single rounding
    k, n, g = 256, 64, 32
    q = torch.arange(k * n).view(k, n) % 16
    z = (torch.arange((k // g) * n).view(k // g, n) // 3) % 16
    s = (torch.rand(k // g, n) * 0.013 + 1e-3).half()
    qw, qz = qr.awq_pack(q), qr.awq_pack(z)
    out = ops.awq_dequantize(qw.to(DEV), s.to(DEV), qz.to(DEV), 0, 0, 0)
    assert torch.equal(out.cpu(), qr.awq_dequantize_ref(qw, s, qz))


@pytest.mark.parametrize("m", [1, 7, 16, 33, 64, 100, 200])
@pytest.mark.parametrize("k,n,g", [(512, 256, 128), (1024, 192, 64), (4096, 1024, 128)])
def test_awq_gemm_vs_oracle(ops, m, k, n, g):
    w_ref, qw, s, qz = _weights(k, n, g, seed=m + k)
    torch.manual_seed(m)
    a = torch.randn(m, k, dtype=torch.half)
    out = ops.awq_gemm(a.to(DEV), qw.to(DEV), s.to(DEV), qz.to(DEV), 8)
    ref = a.float() @ w_ref.float()
    assert tuple(out.shape) == (m, n)
    torch.testing.assert_close(out.cpu().float(), ref, rtol=1e-2, atol=1e-3 * float(ref.abs().max()) + 1e-3)


def test_awq_gemm_identity_activations_are_the_dequantised_weights(ops):
    # A = I: the GEMM output is the weight tile the tensor core saw -- bit-exact against the oracle,
    # which pins the shared-memory operand layout (a misplaced k row or channel shows up here)
    k, n, g = 256, 256, 64
    w_ref, qw, s, qz = _weights(k, n, g, seed=5)
    a = torch.eye(k, dtype=torch.half)
    for m0 in (0, 128):
        out = ops.awq_gemm(a[m0:m0 + 128].contiguous().to(DEV), qw.to(DEV), s.to(DEV), qz.to(DEV), 1)
        assert torch.equal(out.cpu(), w_ref[m0:m0 + 128])


def test_awq_gemm_strided_activations_and_repeat(ops):
    k, n, g = 1024, 512, 128
    w_ref, qw, s, qz = _weights(k, n, g, seed=9)
    wide = torch.randn(24, k + 64, dtype=torch.half, device=DEV)
    a = wide[:, :k]
    qw, s, qz = qw.to(DEV), s.to(DEV), qz.to(DEV)
    o1 = ops.awq_gemm(a, qw, s, qz, 8)
    o2 = ops.awq_gemm(a, qw, s, qz, 8)
    assert torch.equal(o1, o2)  # deterministic: fixed split-K summation order
    ref = a.float().cpu() @ w_ref.float()
    torch.testing.assert_close(o1.cpu().float(), ref, rtol=1e-2, atol=1e-3 * float(ref.abs().max()) + 1e-3)


def test_awq_errors(ops):
    _, qw, s, qz = _weights(256, 64, 128, seed=1)
    a = torch.randn(4, 256, dtype=torch.half, device=DEV)
    with pytest.raises(RuntimeError):
        ops.awq_gemm(a[:, :128], qw.to(DEV), s.to(DEV), qz.to(DEV), 8)        # IC mismatch
    with pytest.raises(RuntimeError):
        ops.awq_gemm(a.float(), qw.to(DEV), s.to(DEV), qz.to(DEV), 8)         # fp16 only
    with pytest.raises(RuntimeError):
        ops.awq_dequantize(qw.to(DEV), s.to(DEV)[:, :32].contiguous(), qz.to(DEV), 0, 0, 0)  # scales shape


# ---- against the reference's own kernels ---------------------------------------------------------

@pytest.fixture(scope="module")
def ref():
    if not ref_cuda.available():
        pytest.skip(ref_cuda.why_unavailable())
    return ref_cuda.ops()


@pytest.mark.parametrize("k,n,g", [(512, 256, 128), (4096, 1024, 128), (256, 128, 32)])
def test_awq_dequantize_vs_reference_cuda(ops, ref, k, n, g):
    _, qw, s, qz = _weights(k, n, g, seed=3)
    qw, s, qz = qw.to(DEV), s.to(DEV), qz.to(DEV)
    assert torch.equal(ops.awq_dequantize(qw, s, qz, 0, 0, 0), ref.awq_dequantize(qw, s, qz, 0, 0, 0))


@pytest.mark.parametrize("m", [1, 16, 64])
@pytest.mark.parametrize("k,n,g", [(4096, 4096, 128), (4096, 1024, 64)])
def test_awq_gemm_vs_reference_cuda(ops, ref, m, k, n, g):
    _, qw, s, qz = _weights(k, n, g, seed=m)
    qw, s, qz = qw.to(DEV), s.to(DEV), qz.to(DEV)
    a = torch.randn(m, k, dtype=torch.half, device=DEV)
    theirs = ref.awq_gemm(a, qw, s, qz, 8)
    ours = ops.awq_gemm(a, qw, s, qz, 8)
    atol = 2 * 2.0 ** -10 * float(theirs.float().abs().max()) + 1e-3
    torch.testing.assert_close(ours.float(), theirs.float(), rtol=1e-2, atol=atol)
