import sys, torch
import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import neural_magic_vllm_b200 as pkg
ops = pkg.ops
g = torch.Generator(device="cuda").manual_seed(0)
def run(k, n, copies, m=16):
    qs = [torch.randint(-2**31, 2**31 - 1, (k, n // 8), dtype=torch.int32, device="cuda", generator=g) for _ in range(copies)]
    zs = [torch.randint(-2**31, 2**31 - 1, (k // 128, n // 8), dtype=torch.int32, device="cuda", generator=g) for _ in range(copies)]
    ss = [(torch.rand(k // 128, n, device="cuda", generator=g) * 0.02 + 0.001).half() for _ in range(copies)]
    a = torch.randn(m, k, dtype=torch.half, device="cuda")
    reps = 48 // copies
    for i in range(copies): ops.awq_gemm(a, qs[i], ss[i], zs[i], 8)
    torch.cuda.synchronize()
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gr):
        for _ in range(reps):
            for i in range(copies): ops.awq_gemm(a, qs[i], ss[i], zs[i], 8)
    gr.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5): gr.replay()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / (5 * reps * copies) * 1e3
for (k, n) in ((4096, 4096), (4096, 28672)):
    print(k, n, "L2-hot (1 copy): %.1f us" % run(k, n, 1), " rotating (48 copies): %.1f us" % run(k, n, 48 if n == 4096 else 6))
