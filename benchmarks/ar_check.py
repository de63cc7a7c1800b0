"""Multi-GPU check + latency of the one-shot NVLink all-reduce against NCCL (run under torchrun).

  python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 benchmarks/ar_check.py

Correctness: (a) every rank holds identical bits; (b) the result equals the fp32 sum of the ranks'
inputs in rank order, rounded once (reconstructed from an all_gather) -- bit-exact; (c) within
rtol 1e-2 / atol 1e-3 of NCCL's all_reduce; eagerly and from a replayed CUDA graph with varying
sizes back to back (the epoch / double-buffer protocol).  Then latency per call from a CUDA graph of
64 calls, one-shot vs NCCL."""
from __future__ import annotations

import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); local = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from neural_magic_vllm_b200.custom_all_reduce import CustomAllreduce
    car = CustomAllreduce(None, local, max_size=8 * 1024 * 1024)  # both kernels: LL <= 2 MiB < staged
    assert not car.disabled
    ok = True
    g = torch.Generator(device="cuda").manual_seed(100 + rank)
    sizes = [8, 4096, 8 * 4096, 1000 * 8, 64 * 4096, 256 * 4096, 1024 * 1024, 3 * 1024 * 1024]
    for dtype in (torch.half, torch.bfloat16, torch.float):
        for n in sizes:
            if n * torch.tensor([], dtype=dtype).element_size() > car.max_size:
                continue
            for it in range(3):
                x = torch.randn(n, device="cuda", generator=g).to(dtype)
                out = car.custom_all_reduce(x)
                assert out is not None
                gathered = [torch.empty_like(x) for _ in range(world)]
                dist.all_gather(gathered, x)
                acc = torch.zeros(n, device="cuda", dtype=torch.float32)
                for t in gathered:
                    acc += t.float()
                exact = acc.to(dtype)
                ref = x.clone(); dist.all_reduce(ref)
                outs = [torch.empty_like(out) for _ in range(world)]
                dist.all_gather(outs, out)
                same = all(torch.equal(o, outs[0]) for o in outs)
                bit = torch.equal(out, exact)
                close = torch.allclose(out.float(), ref.float(), rtol=1e-2, atol=1e-3 if dtype is not torch.bfloat16 else 2e-2)
                if not (same and bit and close):
                    ok = False
                    if rank == 0:
                        print(f"MISMATCH dtype={dtype} n={n} it={it} same={same} bit={bit} close={close}", flush=True)
    # not eligible -> None (caller falls back to NCCL)
    assert car.custom_all_reduce(torch.zeros(7, device="cuda", dtype=torch.half)) is None
    assert car.custom_all_reduce(torch.zeros(car.max_size, device="cuda", dtype=torch.half)) is None
    # graph: alternating sizes back to back, replayed
    xs = [torch.randn(n, device="cuda", generator=g).half() for n in (4096, 256 * 4096, 8, 64 * 4096, 4096)]
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for x in xs:
            car.custom_all_reduce(x)
        s.synchronize()
        gph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gph, stream=s):
            outs = [car.custom_all_reduce(x) for x in xs]
        for _ in range(20):
            gph.replay()
        s.synchronize()
    for x, o in zip(xs, outs):
        ref = x.clone(); dist.all_reduce(ref)
        if not torch.allclose(o.float(), ref.float(), rtol=1e-2, atol=1e-3):
            ok = False
            if rank == 0:
                print("GRAPH MISMATCH", x.numel(), flush=True)
    flag = torch.tensor([1 if ok else 0], device="cuda"); dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    ok = bool(flag.item())

    # latency
    rows = []
    for n in (4096, 8 * 4096, 64 * 4096, 256 * 4096, 1024 * 4096):
        x = torch.randn(n, device="cuda", generator=g).half()
        res = {}
        for name, fn in (("one_shot", lambda: car.custom_all_reduce(x)), ("nccl", lambda: dist.all_reduce(x))):
            with torch.cuda.stream(s):
                for _ in range(3):
                    fn()
                s.synchronize()
                gph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(gph, stream=s):
                    for _ in range(64):
                        fn()
                gph.replay(); s.synchronize()
                dist.barrier()
                st, en = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                st.record(s)
                for _ in range(10):
                    gph.replay()
                en.record(s)
                s.synchronize()
            t = torch.tensor([st.elapsed_time(en) * 1e3 / 640], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            res[name] = round(float(t.item()), 2)
        rows.append({"bytes": n * 2, "world": world, "us_one_shot": res["one_shot"], "us_nccl": res["nccl"]})
    if rank == 0:
        print(json.dumps({"all_reduce_check": "ok" if ok else "FAILED", "world": world}), flush=True)
        for r in rows:
            print(json.dumps(r), flush=True)
    torch.cuda.synchronize()
    dist.barrier()
    sys.stdout.flush()
    os._exit(0 if ok else 1)


if __name__ == "__main__":
    main()
