import sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import neural_magic_vllm_b200 as pkg
from oracle import quant_ref as qr, ref_cuda
ops = pkg.ops
assert ref_cuda.available(), ref_cuda.why_unavailable()
ref = ref_cuda.ops()
DEV = "cuda"
e = torch.empty(0, dtype=torch.int32, device=DEV)
for (k, n) in [(4096, 4096), (4096, 28672)]:
    g = torch.Generator(device=DEV).manual_seed(0)
    w = torch.randn(k, n, dtype=torch.half, device=DEV, generator=g)
    w_ref, mq, ms, _, _, _ = qr.marlin_quantize(w, 4, 128, False)
    for m in (1, 16, 32, 48, 64, 128, 256, 2048):
        a = torch.randn(m, k, dtype=torch.half, device=DEV, generator=g)
        truth = a.float() @ w_ref.float()
        ws = torch.zeros(n // 64 * 16, dtype=torch.int32, device=DEV)
        o = ops.gptq_marlin_gemm(a, mq, ms, e, e, ws, 4, m, n, k, True).float()
        t = ref.gptq_marlin_gemm(a, mq, ms, e, e, ws, 4, m, n, k, True).float()
        t2 = ref.gptq_marlin_gemm(a, mq, ms, e, e, ws, 4, m, n, k, True).float()
        print(k, n, m, "ours-truth", float((o - truth).abs().max()), "ref-truth", float((t - truth).abs().max()),
              "ref-ref", float((t - t2).abs().max()), "ws", int(ws.abs().sum()), "max|truth|", float(truth.abs().max()), flush=True)
