#!/usr/bin/env python
"""Headline benchmark: paged_attention_v2 at the Llama-3-8B decode shape (BASELINE.json configs[1]).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--kv fp16|fp8] [--impl b200|reference]

One "step" = one pass of the hot path over one batch: a `paged_attention_v2` call (split-KV
kernel + reduce) on 64 sequences x 4096 context, 32 query / 8 kv heads, head 128, block 16.
Prints ONE JSON line (contract in DESIGN.md section 6):
  value    whole-job algorithmic HBM GB/s, inputs resident in HBM, device-timed (CUDA events)
  e2e      same metric through the public op surface with HOST step inputs: per step
           H2D(query, block_tables, seq_lens from pinned memory) + op + D2H(out)
  roofline achieved GB/s of the op vs the measured HBM copy peak (MEASURED_PEAKS.json)
  cpu_baseline  the reference's own CPU kernel (oracle/_ref, csrc/cpu/attention.cpp) on host cores
`--impl reference` times only the reference CPU kernel on the same config.
N>1 (torchrun): every rank runs the same per-GPU workload on its own batch (the path shards by
sequence / head with no collective): "scaling": "weak".
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

NUM_SEQS, NUM_HEADS, NUM_KV_HEADS, HEAD, BLOCK, CTX = 64, 32, 8, 128, 16, 4096
PARTITION = 512


def algorithmic_bytes(kv_bytes_per_elem: int) -> int:
    """KV read once + q + out + block_tables + seq_lens."""
    kv = 2 * NUM_SEQS * CTX * NUM_KV_HEADS * HEAD * kv_bytes_per_elem
    q = NUM_SEQS * NUM_HEADS * HEAD * 2
    out = q
    bt = NUM_SEQS * (CTX // BLOCK) * 4
    sl = NUM_SEQS * 4
    return kv + q + out + bt + sl


def measured_peak_hbm():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured"
        except Exception:
            pass
    return 6650.0, "fallback"


class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.proc = index, None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-lms", "100", "-i", str(self.index)],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            out, _ = self.proc.communicate(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
            out, _ = self.proc.communicate()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in out.strip().splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for n, v in zip(names, f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None,
                "sm_max_mhz": max(mx) if mx else None, "samples": len(sm),
                "reasons": sorted(reasons)}


def cpu_reference_run(steps: int, warmup: int, seqs: int):
    """Times the reference's CPU paged_attention_v2 (csrc/cpu/attention.cpp:739) on bf16
    (the CPU backend has no fp16: cpu_types_x86.hpp:16-18) for `seqs` sequences of the C2 shape.
    Returns (GB/s algorithmic, seconds per call, cores, sample description)."""
    import torch
    from oracle import ref_cpu
    if not ref_cpu.available():
        return None, None, 0, ref_cpu.why_unavailable()
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    s = HEAD ** -0.5
    nb = seqs * (CTX // BLOCK)
    g = torch.Generator().manual_seed(0)
    q = torch.empty(seqs, NUM_HEADS, HEAD, dtype=torch.bfloat16).uniform_(-s, s, generator=g)
    kc = torch.empty(nb, NUM_KV_HEADS, HEAD // 8, BLOCK, 8, dtype=torch.bfloat16).uniform_(-s, s, generator=g)
    vc = torch.empty(nb, NUM_KV_HEADS, HEAD, BLOCK, dtype=torch.bfloat16).uniform_(-s, s, generator=g)
    bt = torch.randperm(nb, generator=g).to(torch.int32).view(seqs, CTX // BLOCK)
    sl = torch.full((seqs,), CTX, dtype=torch.int32)
    parts = CTX // PARTITION
    out = torch.empty_like(q)
    es = torch.empty(seqs, NUM_HEADS, parts)
    ml = torch.empty_like(es)
    tmp = torch.empty(seqs, NUM_HEADS, parts, HEAD, dtype=torch.bfloat16)

    def call():
        ref_cpu.paged_attention_v2(out, es, ml, tmp, q, kc, vc, NUM_KV_HEADS, s, bt, sl, BLOCK, CTX)

    for _ in range(max(warmup, 1)):
        call()
    t0 = time.perf_counter()
    for _ in range(steps):
        call()
    dt = (time.perf_counter() - t0) / steps
    nbytes = algorithmic_bytes(2) * seqs / NUM_SEQS
    return nbytes / dt / 1e9, dt, cores, (
        f"reference csrc/cpu paged_attention_v2, bf16 KV, {seqs} seqs x ctx {CTX} "
        f"({nbytes / 2**20:.0f} MiB/call), {steps} calls")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--kv", default="fp16", choices=["fp16", "fp8"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--profile", action="store_true",
                    help="ncu mode: warm-up + timed steps only (no e2e / clock soak / CPU baseline)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    config = {"workload": "paged_attention_v2 Llama-3-8B shape (configs[1])", "num_seqs": NUM_SEQS,
              "ctx": CTX, "q_heads": NUM_HEADS, "kv_heads": NUM_KV_HEADS, "head_size": HEAD,
              "block_size": BLOCK, "kv_cache": args.kv, "distinct_blocks": NUM_SEQS * CTX // BLOCK,
              "l2": "inputs larger than L2 (KV working set 1 GiB fp16 / 512 MiB fp8 vs 126 MB L2)",
              "parallelism": f"replica x{world} (independent batches, no collective)"}
    metric = "paged_attention_v2 algorithmic HBM GB/s (Llama-3-8B shape, batch 64, ctx 4096)"

    if args.impl == "reference":
        if rank != 0:
            return
        seqs = 16  # bounded sample: 256 MiB of bf16 KV per call
        steps = max(1, min(args.steps, 20))
        gbs, dt, cores, sample = cpu_reference_run(steps, min(args.warmup, 2), seqs)
        if gbs is None:
            print(json.dumps({"impl": "reference", "unavailable": sample}))
            return
        config["kv_cache"] = "bf16 (reference CPU backend has no fp16/fp8)"
        print(json.dumps({
            "impl": "reference", "metric": metric, "value": gbs, "unit": "GB/s", "n_gpus": args.gpus,
            "steps": steps, "warmup": min(args.warmup, 2), "ms_per_step": dt * 1e3 * NUM_SEQS / seqs,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
            "data": "synthetic", "config": config,
            "cpu_baseline": {"value": gbs, "unit": "GB/s", "cores": cores, "kind": "reference",
                             "sample": sample},
            "e2e": {"value": gbs, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}))
        return

    import torch
    import torch.distributed as dist
    assert torch.cuda.is_available(), "bench.py needs a GPU (there is no CPU fallback)"
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    import neural_magic_vllm_b200 as pkg
    ops = pkg.ops

    fp8 = args.kv == "fp8"
    kv_str = "fp8" if fp8 else "auto"
    esz = 1 if fp8 else 2
    x = 16 // esz
    s = HEAD ** -0.5
    torch.manual_seed(rank)
    nb = NUM_SEQS * CTX // BLOCK
    q = torch.empty(NUM_SEQS, NUM_HEADS, HEAD, dtype=torch.float16, device=dev).uniform_(-s, s)
    kc = torch.empty(nb, NUM_KV_HEADS, HEAD // x, BLOCK, x, dtype=torch.float16, device=dev).uniform_(-s, s)
    vc = torch.empty(nb, NUM_KV_HEADS, HEAD, BLOCK, dtype=torch.float16, device=dev).uniform_(-s, s)
    if fp8:
        kc8 = torch.empty(kc.shape, dtype=torch.uint8, device=dev)
        vc8 = torch.empty(vc.shape, dtype=torch.uint8, device=dev)
        ops.convert_fp8(kc8, kc, 1.0, "fp8")
        ops.convert_fp8(vc8, vc, 1.0, "fp8")
        kc, vc = kc8, vc8
    bt = torch.randperm(nb, device=dev).to(torch.int32).view(NUM_SEQS, CTX // BLOCK).contiguous()
    sl = torch.full((NUM_SEQS,), CTX, dtype=torch.int32, device=dev)
    parts = CTX // PARTITION
    out = torch.empty_like(q)
    es = torch.empty(NUM_SEQS, NUM_HEADS, parts, dtype=torch.float32, device=dev)
    ml = torch.empty_like(es)
    tmp = torch.empty(NUM_SEQS, NUM_HEADS, parts, HEAD, dtype=torch.float16, device=dev)

    def step():
        ops.paged_attention_v2(out, es, ml, tmp, q, kc, vc, NUM_KV_HEADS, s, bt, sl, BLOCK, CTX,
                               None, kv_str, 1.0)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(warmup):
        step()
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    # the op is ~0.2 ms: repeat the timed block so the clock sampler sees the GPU under load
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record()
    for _ in range(args.steps):
        step()
    ev1.record()
    barrier()
    ms_total = ev0.elapsed_time(ev1)

    if args.profile:
        if rank == 0:
            sampler.stop()
            print(json.dumps({"profile_mode": True, "ms_per_step": ms_total / args.steps}))
        return

    # ---- end to end: host step inputs -> device, op, result -> host ----
    q_h = q.cpu().pin_memory()
    bt_h = bt.cpu().pin_memory()
    sl_h = sl.cpu().pin_memory()
    out_h = torch.empty(out.shape, dtype=out.dtype).pin_memory()
    h2d = q_h.numel() * 2 + bt_h.numel() * 4 + sl_h.numel() * 4
    d2h = out_h.numel() * 2

    def e2e_step():
        q.copy_(q_h, non_blocking=True)
        bt.copy_(bt_h, non_blocking=True)
        sl.copy_(sl_h, non_blocking=True)
        step()
        out_h.copy_(out, non_blocking=True)
        torch.cuda.current_stream().synchronize()

    for _ in range(3):
        e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        e2e_step()
    barrier()
    e2e_s = time.perf_counter() - t0
    # the timed regions are tens of milliseconds: keep the same kernel running for ~1.5 s so that
    # the 100 ms nvidia-smi sampler sees the clocks under this load
    t_soak = time.perf_counter()
    while time.perf_counter() - t_soak < 1.5:
        for _ in range(50):
            step()
        torch.cuda.synchronize()
    clocks = sampler.stop() if rank == 0 else None
    if clocks is not None:
        clocks["window"] = "timed region + e2e region + 1.5 s continuation of the same op"

    t = torch.tensor([ms_total, e2e_s * 1e3], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total, e2e_ms = float(t[0]), float(t[1])
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    nbytes = algorithmic_bytes(esz)
    ms_step = ms_total / args.steps
    value = nbytes * world / (ms_step * 1e-3) / 1e9
    e2e_value = nbytes * world / (e2e_ms / args.steps * 1e-3) / 1e9
    peak, peak_kind = measured_peak_hbm()
    per_gpu = nbytes / (ms_step * 1e-3) / 1e9
    # dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed ncu --set full
    # capture of this workload (profiles/traffic.json names the source file); null when there is none
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            traffic = json.load(f).get("paged_attention_v2_" + ("fp8" if fp8 else "fp16"), {}).get("bytes")
    except Exception:
        traffic = None
    line = {
        "metric": metric, "value": value, "unit": "GB/s", "n_gpus": world, "steps": args.steps,
        "warmup": warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f16 (fp32 softmax/accumulate)" + (", fp8-e4m3 KV" if fp8 else ""),
        "data": "synthetic", "config": config,
        "roofline": {"bound": "hbm", "achieved": per_gpu, "peak": peak, "unit": "GB/s",
                     "frac": per_gpu / peak, "traffic": traffic, "peak_kind": f"of {peak_kind}",
                     "frac_of_8TBs_spec": per_gpu / 8000.0,
                     "kernel": "paged_attn_mma_kernel (+ reduce kernel: duration is the whole op)",
                     "algorithmic_bytes_per_launch": nbytes},
        "e2e": {"value": e2e_value, "unit": "GB/s", "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": d2h, "ms_per_step": e2e_ms / args.steps},
        "gpu_launches": 2 * args.steps,
        "tokens_attended_per_s": NUM_SEQS * CTX * world / (ms_step * 1e-3),
        "clocks": clocks,
    }
    if not args.no_cpu_baseline and world >= 1:
        gbs, dt, cores, sample = cpu_reference_run(3, 1, 8)
        line["cpu_baseline"] = ({"value": gbs, "unit": "GB/s", "cores": cores, "kind": "reference",
                                 "sample": sample} if gbs is not None else
                                {"value": None, "unit": "GB/s", "cores": 0, "kind": "reference",
                                 "sample": f"unavailable: {sample}"})
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
