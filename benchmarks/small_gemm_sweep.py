"""Split-K sweep of gptq_marlin_gemm on tensor-parallel shards (latency floor study).
Each configuration runs in its own process (B200_MARLIN_SPLITS is read once per process)."""
import json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SHAPES = {"qkv_tp8": (4096, 768), "o_tp8": (512, 4096), "gate_up_tp8": (4096, 3584), "down_tp8": (1792, 4096),
          "qkv_tp2": (4096, 3072), "o_tp2": (2048, 4096)}
if os.environ.get("SWEEP_SHAPES") == "tp1":  # the unsharded Llama-3-8B layers that take the split-K kernel
    SHAPES = {"qkv": (4096, 6144), "o": (4096, 4096), "down": (14336, 4096)}
if os.environ.get("SWEEP_SHAPES", "").startswith("{"):  # e.g. '{"a": [4096, 18944]}'
    SHAPES = {k: tuple(v) for k, v in json.loads(os.environ["SWEEP_SHAPES"]).items()}
MS = tuple(int(x) for x in os.environ.get("SWEEP_MS", "1,64").split(","))
CHILD = r'''
import sys, json, torch
sys.path.insert(0, %r)
import neural_magic_vllm_b200 as pkg
ops = pkg.ops
shapes = %r
g = torch.Generator(device="cuda").manual_seed(0)
for name, (k, n) in shapes.items():
    copies = 24
    qs = [torch.randint(-2**31, 2**31 - 1, (k // 16, n * 2), dtype=torch.int32, device="cuda", generator=g) for _ in range(copies)]
    ss = [(torch.rand(k // 128, n, device="cuda", generator=g) * 0.02 + 0.001).half() for _ in range(copies)]
    e = torch.empty(0, dtype=torch.int32, device="cuda")
    ws = torch.zeros(max(n // 64 * 16, 16), dtype=torch.int32, device="cuda")
    for m in %r:
        a = torch.randn(m, k, dtype=torch.half, device="cuda")
        for i in range(copies): ops.gptq_marlin_gemm(a, qs[i], ss[i], e, e, ws, 4, m, n, k, True)
        torch.cuda.synchronize()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr):
            for i in range(copies): ops.gptq_marlin_gemm(a, qs[i], ss[i], e, e, ws, 4, m, n, k, True)
        gr.replay(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10): gr.replay()
        e1.record(); torch.cuda.synchronize()
        print(json.dumps({"layer": name, "m": m, "us": e0.elapsed_time(e1) / (10 * copies) * 1e3}), flush=True)
'''
def main():
    res = {}
    for splits in tuple(int(x) for x in os.environ.get("SWEEP_SPLITS", "0,1,2,4,8").split(",")):
        env = dict(os.environ)
        if splits: env["B200_MARLIN_SPLITS"] = str(splits)
        out = subprocess.run([sys.executable, "-c", CHILD % (ROOT, SHAPES, MS)], env=env, capture_output=True, text=True).stdout
        for l in out.splitlines():
            if l.startswith("{"):
                d = json.loads(l); res.setdefault((d["layer"], d["m"]), {})[splits] = round(d["us"], 2)
    for k, v in res.items():
        print(k, v)
if __name__ == "__main__":
    main()
