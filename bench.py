#!/usr/bin/env python
"""Headline benchmark: Llama-3-8B W4A16 (gptq_marlin g128) decode tokens/s at TP = N (BASELINE.json
configs[2], metric (i)), with the configs[1] paged_attention_v2 microbench (metric (ii), fp16 and
fp8-e4m3 KV) nested in the same JSON line.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W

One "step" = one decode step of the whole model over one batch (headline: 64 sequences at context
1024): embedding gather, 32 x [RMSNorm, qkv gptq_marlin_gemm, rotary + reshape_and_cache,
paged_attention, o gptq_marlin_gemm, (all-reduce +) add + RMSNorm, gate_up gptq_marlin_gemm,
SiLU*mul, down gptq_marlin_gemm, (all-reduce +) add + RMSNorm], fp16 lm_head, greedy sampling --
replayed from one CUDA graph, as the reference replays decode (worker/model_runner.py).
N > 1 is tensor parallel: rank r owns q/kv heads and the column shards of
qkv / gate_up and the K shards of o / down; the two row-parallel partial sums per layer
(linear.py:792-793) go through this repo's one-shot NVLink all-reduce (fused with add + RMSNorm
where the message qualifies) or NCCL above its size limit.  "scaling": "strong".

ONE JSON line (contract in DESIGN.md section 6):
  value      whole-job tokens/s at batch 64, inputs resident in HBM, device-timed (CUDA events)
  e2e        same, through the graph with HOST step inputs: per step H2D(tokens, positions,
             slot_mapping, block_tables, seq_lens from pinned memory) + step + D2H(sampled ids)
  roofline   (weight + KV + lm_head bytes per GPU) / step time vs the measured HBM copy peak,
             plus the dominant kernel (gate_up gptq_marlin_gemm) timed alone, live
  batches    tokens/s at batch 1 and 256 as well
  attention  configs[1]: paged_attention_v2 64 x 4096, fp16 and fp8 KV, GB/s + roofline each
  checks     output checked against the oracle once, outside the timed regions; at N > 1 the
             all-reduce bit-identity check and a TP-N vs TP-1 hidden-state comparison
  cpu_baseline  the reference's CPU path for the same step on the host cores (N = 1 only)
`--impl reference` times only that CPU path (compiled csrc/cpu attention + cache ops, torch
dequant-matmul linears) on a bounded sample of the same step.
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

HEADLINE_BATCH, EXTRA_BATCHES, DECODE_CTX = 64, (1, 256), 1024
# configs[1] attention microbench
NUM_SEQS, NUM_HEADS, NUM_KV_HEADS, HEAD, BLOCK, CTX = 64, 32, 8, 128, 16, 4096
PARTITION = 512
METRIC = "decode tokens/s Llama-3-8B W4A16 gptq_marlin g128 (whole job at TP=n_gpus; /n_gpus = per GPU)"


def attention_bytes(kv_bytes_per_elem: int) -> int:
    """KV read once + q + out + block_tables + seq_lens."""
    kv = 2 * NUM_SEQS * CTX * NUM_KV_HEADS * HEAD * kv_bytes_per_elem
    q = NUM_SEQS * NUM_HEADS * HEAD * 2
    return kv + 2 * q + NUM_SEQS * (CTX // BLOCK) * 4 + NUM_SEQS * 4


def measured_peak_hbm():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


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
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in out.strip().splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
                pw.append(float(f[3]))
            except ValueError:
                continue
            for n, v in zip(names, f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        # "under load" = samples drawing well above idle power
        load = [s for s, p in zip(sm, pw) if p > 300.0] or sm
        return {"sm_mhz": statistics.median(load) if load else None,
                "sm_max_mhz": max(mx) if mx else None, "samples": len(sm),
                "samples_under_load": len(load), "reasons": sorted(reasons)}


def cpu_reference(steps: int, warmup: int):
    """The CPU arm (oracle/cpu_decode_ref.py): bounded sample of the headline step."""
    from oracle import cpu_decode_ref
    return cpu_decode_ref.run(batch=HEADLINE_BATCH, ctx=DECODE_CTX, sample_layers=2,
                              steps=max(1, min(steps, 3)), warmup=max(1, min(warmup, 1)))


class CountingOps:
    """Proxy over the op surface that counts kernel launches of THIS repo's ops during one eager
    step (lower bound: a split-K GEMM with its reduce kernel is counted once)."""
    LAUNCHES = {"paged_attention_v2": 2}

    def __init__(self, ops):
        self._ops, self.count = ops, 0

    def __getattr__(self, name):
        fn = getattr(self._ops, name)

        def wrapped(*a, **k):
            self.count += self.LAUNCHES.get(name, 1)
            return fn(*a, **k)
        return wrapped


def graph_step(dec, torch):
    """Warm the step eagerly, then capture it into one CUDA graph; -> (graph, replay)."""
    with torch.no_grad():
        for _ in range(2):
            dec.step()
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            dec.step()
    return g


def time_steps(run, steps, warmup, barrier, torch):
    """W untimed + exactly K timed calls of `run`, CUDA events on the launching stream, barrier +
    synchronize on both sides.  -> ms per step on this rank."""
    for _ in range(warmup):
        run()
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(steps):
        run()
    ev1.record()
    barrier()
    return ev0.elapsed_time(ev1) / steps


def decode_bytes(dec, batch, ctx):
    """Algorithmic HBM bytes of one decode step on ONE GPU: every packed
    weight + scale of the rank's shards once, the rank's KV heads of every sequence once, the
    rank's lm_head shard once.  Activations (batch x 4096 fp16 per op) are not counted."""
    lin = dec.linear_bytes()
    kv = 2 * batch * (ctx + 1) * dec.layers[0].kvh * HEAD * 2 * len(dec.layers)
    head = dec.lm_head.numel() * 2 if dec.lm_head is not None else 0
    return lin, kv, head


def run_decode(ops, torch, dist, world, rank, car, batch, steps, warmup, barrier, e2e: bool):
    """One batch size: build the TP shard of the decoder, capture, time.  -> dict."""
    from benchmarks.decode_bench import Decoder
    dec = Decoder(ops, world, rank, batch, DECODE_CTX, lm_head=True, glue="fused", car=car)
    counter = CountingOps(ops)
    dec.ops = counter
    for L in dec.layers:
        for lin in (L.qkv, L.o, L.gate_up, L.down):
            lin.ops = counter
    with torch.no_grad():
        dec.step()
    launches = counter.count
    dec.ops = ops
    for L in dec.layers:
        for lin in (L.qkv, L.o, L.gate_up, L.down):
            lin.ops = ops
    g = graph_step(dec, torch)
    ms = time_steps(g.replay, steps, warmup, barrier, torch)
    lin, kv, head = decode_bytes(dec, batch, DECODE_CTX)
    res = {"batch": batch, "ms_per_step": ms, "launches_per_step": launches,
           "bytes_per_gpu": lin + kv + head, "linear_bytes_per_gpu": lin, "kv_bytes_per_gpu": kv,
           "lm_head_bytes_per_gpu": head,
           "attention_op": "paged_attention_v1" if dec.use_v1 else "paged_attention_v2"}
    if e2e:
        # host-resident step inputs (what the reference's model_runner prepares every step:
        # worker/model_runner.py _prepare_model_input) in pinned memory; sampled ids back to host
        host = {k: getattr(dec, k).cpu().pin_memory()
                for k in ("tokens", "positions", "slots", "block_tables", "seq_lens")}
        next_h = torch.empty(batch, dtype=torch.int64).pin_memory()
        h2d = sum(t.numel() * t.element_size() for t in host.values())
        d2h = next_h.numel() * 8
        stream = torch.cuda.current_stream()

        def e2e_step():
            for k, t in host.items():
                getattr(dec, k).copy_(t, non_blocking=True)
            g.replay()
            next_h.copy_(dec.next_tokens, non_blocking=True)
            stream.synchronize()

        for _ in range(3):
            e2e_step()
        barrier()
        t0 = time.perf_counter()
        for _ in range(steps):
            e2e_step()
        barrier()
        res.update({"e2e_ms_per_step": (time.perf_counter() - t0) / steps * 1e3,
                    "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h})
    res["_dec"], res["_graph"] = dec, g
    return res


def run_attention(ops, torch, dev, fp8: bool, steps, warmup, barrier, peak):
    """configs[1]: paged_attention_v2, 64 seqs x ctx 4096, 32 q / 8 kv heads, D 128, block 16, a
    random permutation of 16384 distinct blocks (KV working set 1 GiB / 512 MiB > 126 MB L2).
    The output of 2 sequences is checked against the oracle once, outside the timed region."""
    from oracle import attention_ref as ar
    kv_str, esz = ("fp8", 1) if fp8 else ("auto", 2)
    x = 16 // esz
    s = HEAD ** -0.5
    g = torch.Generator(device=dev).manual_seed(0)
    nb = NUM_SEQS * CTX // BLOCK
    q = torch.empty(NUM_SEQS, NUM_HEADS, HEAD, dtype=torch.float16, device=dev).uniform_(-s, s, generator=g)
    kc = torch.empty(nb, NUM_KV_HEADS, HEAD // x, BLOCK, x, dtype=torch.float16, device=dev).uniform_(-s, s, generator=g)
    vc = torch.empty(nb, NUM_KV_HEADS, HEAD, BLOCK, dtype=torch.float16, device=dev).uniform_(-s, s, generator=g)
    if fp8:
        kc8 = torch.empty(kc.shape, dtype=torch.uint8, device=dev)
        vc8 = torch.empty(vc.shape, dtype=torch.uint8, device=dev)
        ops.convert_fp8(kc8, kc, 1.0, "fp8")
        ops.convert_fp8(vc8, vc, 1.0, "fp8")
        kc, vc = kc8, vc8
    bt = torch.randperm(nb, device=dev, generator=g).to(torch.int32).view(NUM_SEQS, CTX // BLOCK).contiguous()
    sl = torch.full((NUM_SEQS,), CTX, dtype=torch.int32, device=dev)
    parts = CTX // PARTITION
    out = torch.empty_like(q)
    es = torch.empty(NUM_SEQS, NUM_HEADS, parts, dtype=torch.float32, device=dev)
    ml = torch.empty_like(es)
    tmp = torch.empty(NUM_SEQS, NUM_HEADS, parts, HEAD, dtype=torch.float16, device=dev)

    def step():
        ops.paged_attention_v2(out, es, ml, tmp, q, kc, vc, NUM_KV_HEADS, s, bt, sl, BLOCK, CTX,
                               None, kv_str, 1.0)

    ms = time_steps(step, steps, warmup, barrier, torch)
    # oracle check (2 sequences; their blocks compacted into a small CPU cache)
    sample = [0, NUM_SEQS - 1]
    blocks = bt[sample].reshape(-1).long()
    bt_s = torch.arange(len(sample) * (CTX // BLOCK), dtype=torch.int32).view(len(sample), -1)
    ref = ar.paged_attention_ref(q[sample].cpu(), kc[blocks].cpu(), vc[blocks].cpu(), bt_s, sl[sample].cpu(),
                                 s, None, kv_str, 1.0)
    atol = 1e-2 if fp8 else 1e-3
    torch.testing.assert_close(out[sample].cpu().float(), ref.float(), atol=atol, rtol=1e-2)
    # end to end through the op with host step inputs
    q_h, bt_h, sl_h = q.cpu().pin_memory(), bt.cpu().pin_memory(), sl.cpu().pin_memory()
    out_h = torch.empty(out.shape, dtype=out.dtype).pin_memory()
    stream = torch.cuda.current_stream()

    def e2e_step():
        q.copy_(q_h, non_blocking=True)
        bt.copy_(bt_h, non_blocking=True)
        sl.copy_(sl_h, non_blocking=True)
        step()
        out_h.copy_(out, non_blocking=True)
        stream.synchronize()

    for _ in range(3):
        e2e_step()
    t0 = time.perf_counter()
    for _ in range(steps):
        e2e_step()
    e2e_ms = (time.perf_counter() - t0) / steps * 1e3
    nbytes = attention_bytes(esz)
    gbs = nbytes / (ms * 1e-3) / 1e9
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            traffic = json.load(f).get("paged_attention_v2_" + ("fp8" if fp8 else "fp16"), {}).get("bytes")
    except Exception:
        pass
    return {"metric": "paged_attention_v2 algorithmic HBM GB/s (Llama-3-8B shape, batch 64, ctx 4096)",
            "kv_cache": "fp8_e4m3" if fp8 else "fp16", "value": gbs, "unit": "GB/s", "ms_per_step": ms,
            "steps": steps, "oracle_check": f"ok (2 sequences, rtol 1e-2 atol {atol})",
            "roofline": {"bound": "hbm", "achieved": gbs, "peak": peak, "unit": "GB/s", "frac": gbs / peak,
                         "frac_of_8TBs_spec": gbs / 8000.0, "traffic": traffic,
                         "algorithmic_bytes_per_launch": nbytes,
                         "kernel": "paged_attn_mma_kernel (+ reduce kernel: duration is the whole op)"},
            "e2e": {"value": nbytes / (e2e_ms * 1e-3) / 1e9, "unit": "GB/s", "ms_per_step": e2e_ms,
                    "h2d_bytes_per_step": q_h.numel() * 2 + bt_h.numel() * 4 + sl_h.numel() * 4,
                    "d2h_bytes_per_step": out_h.numel() * 2}}


def time_dominant_kernel(dec, batch, torch, steps, peak):
    """The kernel with the largest share of the headline step: the gate_up gptq_marlin_gemm
    (marlin_gemm_tmem_kernel at M <= 64).  Timed alone with CUDA events, rotating over the 32
    layers' weights (32 x 59 MB > L2) so that every launch streams its weights from HBM."""
    x = torch.randn(batch, 4096, dtype=torch.half, device="cuda")
    lins = [L.gate_up for L in dec.layers]
    with torch.no_grad():
        for lin in lins:
            lin(x)
        torch.cuda.synchronize()
        reps = max(1, steps // 8)
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record()
        for _ in range(reps):
            for lin in lins:
                lin(x)
        ev1.record()
        torch.cuda.synchronize()
    us = ev0.elapsed_time(ev1) * 1e3 / (reps * len(lins))
    lin = lins[0]
    nbytes = lin.bytes() + batch * (lin.k + lin.n) * 2
    gbs = nbytes / us / 1e3
    return {"kernel": "gptq_marlin_gemm gate_up (marlin_gemm_tmem_kernel + split-K sum where split)",
            "shape_mkn": [batch, lin.k, lin.n], "us_per_launch": us, "launches_timed": reps * len(lins),
            "algorithmic_bytes_per_launch": nbytes, "achieved": gbs, "peak": peak, "unit": "GB/s",
            "frac": gbs / peak, "share_of_step": None}


def step_traffic(kern):
    """dram bytes per launch of the dominant kernel from the committed ncu --set full capture, if one
    exists for it (profiles/traffic.json), else None."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            t = json.load(f)
    except OSError:
        return None
    key = "step_attention_b64" if kern["kernel"].startswith("paged_attn") else "step_gate_up_m64"
    return t.get(key, {}).get("bytes")


def time_attention_in_step(dec, ops, batch, torch, steps, peak):
    """The decode attention of the headline step, timed alone with CUDA events: the same op the
    step calls (paged_attention_v1 / v2 by the reference's rule), rotating over the layers' KV
    caches (268 MB per layer per call at batch 64 x ctx 1024: every launch streams from HBM)."""
    L0 = dec.layers[0]
    q = torch.randn(batch, L0.h, HEAD, dtype=torch.half, device="cuda")
    out = torch.empty_like(q)

    def call(L):
        if dec.use_v1:
            ops.paged_attention_v1(out, q, L.kc, L.vc, L.kvh, HEAD ** -0.5, dec.block_tables, dec.seq_lens,
                                   BLOCK, dec.ctx + 1, None, "auto", 1.0)
        else:
            ops.paged_attention_v2(out, dec.es, dec.ml, dec.tmp, q, L.kc, L.vc, L.kvh, HEAD ** -0.5,
                                   dec.block_tables, dec.seq_lens, BLOCK, dec.ctx + 1, None, "auto", 1.0)

    with torch.no_grad():
        for L in dec.layers:
            call(L)
        torch.cuda.synchronize()
        reps = max(1, steps // 8)
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record()
        for _ in range(reps):
            for L in dec.layers:
                call(L)
        ev1.record()
        torch.cuda.synchronize()
    us = ev0.elapsed_time(ev1) * 1e3 / (reps * len(dec.layers))
    nbytes = 2 * int(dec.seq_lens.sum().item()) * L0.kvh * HEAD * 2 + 2 * batch * L0.h * HEAD * 2
    gbs = nbytes / us / 1e3
    return {"kernel": "paged_attn_mma_kernel (%s: duration is the whole op)" %
                      ("paged_attention_v1" if dec.use_v1 else "paged_attention_v2 incl. its reduce kernel"),
            "shape": {"num_seqs": batch, "ctx": dec.ctx + 1, "heads": L0.h, "kv_heads": L0.kvh, "head_size": HEAD},
            "us_per_launch": us, "launches_timed": reps * len(dec.layers),
            "algorithmic_bytes_per_launch": nbytes, "achieved": gbs, "peak": peak, "unit": "GB/s",
            "frac": gbs / peak, "share_of_step": None}


def check_all_reduce(car, torch, dist, world, rank):
    """One-shot all-reduce: bit-identical on every rank and equal to the fp32 sum of the ranks'
    inputs rounded once (benchmarks/ar_check.py holds the full matrix of sizes / dtypes)."""
    g = torch.Generator(device="cuda").manual_seed(100 + rank)
    n_ok = 0
    for n in (4096, 64 * 4096, 8):
        for dtype in (torch.half, torch.bfloat16):
            x = torch.randn(n, device="cuda", generator=g).to(dtype)
            out = car.all_reduce(x)
            gathered = [torch.empty_like(x) for _ in range(world)]
            dist.all_gather(gathered, x)
            acc = torch.zeros(n, device="cuda", dtype=torch.float32)
            for t in gathered:
                acc += t.float()
            assert torch.equal(out, acc.to(dtype)), f"one-shot all-reduce mismatch n={n} {dtype}"
            outs = [torch.empty_like(out) for _ in range(world)]
            dist.all_gather(outs, out)
            assert all(torch.equal(o, outs[0]) for o in outs), "ranks disagree"
            n_ok += 1
    return f"ok ({n_ok} size x dtype cases bit-exact vs the fp32 sum, identical on all ranks)"


def check_tp_vs_tp1(ops, torch, world, rank, car):
    """TP-N hidden state (2 layers, batch 64, through the fused all-reduce + add + RMSNorm) against
    the same 2 layers at TP-1 computed locally from the same full weights."""
    from benchmarks.decode_bench import Decoder
    with torch.no_grad():
        d_tp = Decoder(ops, world, rank, 64, 128, lm_head=False, car=car, num_layers=2)
        h_tp = d_tp.step().float().clone()
        del d_tp
        d_1 = Decoder(ops, 1, 0, 64, 128, lm_head=False, car=None, num_layers=2)
        h_1 = d_1.step().float().clone()
        del d_1
    torch.cuda.synchronize()
    err = float((h_tp - h_1).abs().max() / h_1.abs().max())
    assert err < 2e-2, f"TP-{world} hidden state differs from TP-1: rel-to-max error {err}"
    return {"max_err_rel_to_max": err, "layers": 2, "batch": 64, "bound": 2e-2}


def check_vs_reference_cuda(dec, batch, torch):
    """gate_up gptq_marlin_gemm against the reference's OWN CUDA kernel (oracle/_ref/_C_refcuda.so,
    recompiled for sm_100) on identical inputs; also times that kernel the same way as ours."""
    try:
        from oracle import ref_cuda
        if not ref_cuda.available():
            return {"unavailable": ref_cuda.why_unavailable()}
    except Exception as e:  # the checker is optional on the bench path
        return {"unavailable": repr(e)}
    lins = [L.gate_up for L in dec.layers]
    out = {}
    with torch.no_grad():
        ws = torch.zeros(lins[0].n // 64 * 16, dtype=torch.int32, device="cuda")

        def ref_call(lin, x):
            return torch.ops._C_refcuda.gptq_marlin_gemm(x, lin.q, lin.s, lin.e, lin.e, ws, 4, x.shape[0],
                                                         lin.n, lin.k, True)
        # parity at M = 16, where the reference kernel recompiled for sm_100 is deterministic (from
        # M = 48-64 up two identical calls of it disagree on B200: tests/test_vs_ref_cuda_gpu.py)
        x16 = torch.randn(16, 4096, dtype=torch.half, device="cuda")
        ours, ref = lins[0](x16), ref_call(lins[0], x16)
        atol = 2 * 2.0 ** -10 * float(ref.float().abs().max()) + 1e-3  # the reference reduces k-slices in fp16
        torch.testing.assert_close(ours.float(), ref.float(), rtol=1e-2, atol=atol)
        out["gate_up_m16_vs_reference_cuda"] = f"ok (rtol 1e-2 atol {atol:.2g})"
        x = torch.randn(batch, 4096, dtype=torch.half, device="cuda")
        r1, r2 = ref_call(lins[0], x), ref_call(lins[0], x)
        out["reference_cuda_deterministic_at_headline_m"] = bool(torch.equal(r1, r2))
        torch.cuda.synchronize()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record()
        for lin in lins:
            ref_call(lin, x)
        ev1.record()
        torch.cuda.synchronize()
    out["reference_cuda_gate_up_us_per_launch"] = ev0.elapsed_time(ev1) * 1e3 / len(lins)
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-attention", action="store_true")
    ap.add_argument("--no-extra-batches", action="store_true")
    ap.add_argument("--ar", choices=["one_shot", "nccl"], default="one_shot")
    ap.add_argument("--profile", action="store_true",
                    help="ncu mode: headline batch only, warm-up + timed steps, nothing else")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    config = {"workload": "Llama-3-8B W4A16 gptq_marlin g128 decode step (configs[2]): 32 layers, hidden 4096, "
                          "32 q / 8 kv heads, D 128, inter 14336, vocab 128256, fp16 lm_head, greedy sampling",
              "batch": HEADLINE_BATCH, "ctx": DECODE_CTX, "block_size": BLOCK, "cuda_graph": True,
              "parallelism": f"tp{world}" if world > 1 else "single GPU",
              "l2": "inputs larger than L2 (weights + KV read per step: GBs vs 126 MB L2)",
              "excluded": "scheduler, tokenizer"}

    if args.impl == "reference":
        if rank != 0:
            return
        r = cpu_reference(args.steps, args.warmup)
        if "unavailable" in r:
            print(json.dumps({"impl": "reference", "unavailable": r["unavailable"]}))
            return
        config["parallelism"] = f"host CPU, {r['cores']} threads"
        print(json.dumps({
            "impl": "reference", "metric": METRIC, "value": r["tokens_per_s"], "unit": "tokens/s",
            "n_gpus": args.gpus, "steps": max(1, min(args.steps, 3)), "warmup": 1,
            "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": r["matmul_dtype"] + " (bf16 KV)", "data": "synthetic", "config": config,
            "cpu_baseline": {"value": r["tokens_per_s"], "unit": "tokens/s", "cores": r["cores"],
                             "kind": "port", "sample": r["sample"]},
            "e2e": {"value": r["tokens_per_s"], "unit": "tokens/s", "h2d_bytes_per_step": 0,
                    "d2h_bytes_per_step": 0},
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
    warmup = max(args.warmup, 3)
    peak, peak_kind = measured_peak_hbm()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    car = None
    checks = {}
    if world > 1 and args.ar == "one_shot":
        from neural_magic_vllm_b200.custom_all_reduce import CustomAllreduce
        car = CustomAllreduce(None, local_rank)
        if car.disabled:
            car = None
    if world > 1 and not args.profile:
        if car is not None:
            checks["ar_check"] = check_all_reduce(car, torch, dist, world, rank)
        checks["tp_vs_tp1"] = check_tp_vs_tp1(ops, torch, world, rank, car)
        barrier()

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    head = run_decode(ops, torch, dist, world, rank, car, HEADLINE_BATCH, args.steps, warmup, barrier,
                      e2e=not args.profile)
    if args.profile:
        if rank == 0:
            sampler.stop()
            print(json.dumps({"profile_mode": True, "ms_per_step": head["ms_per_step"]}))
        return
    dec = head.pop("_dec")
    graph = head.pop("_graph")
    dom = time_dominant_kernel(dec, HEADLINE_BATCH, torch, args.steps, peak)
    dom_attn = time_attention_in_step(dec, ops, HEADLINE_BATCH, torch, args.steps, peak)
    if rank == 0:
        checks.update(check_vs_reference_cuda(dec, HEADLINE_BATCH, torch))
    # keep the headline graph replaying ~1.5 s so the 100 ms nvidia-smi sampler sees it under load
    # (the graph holds collectives at N > 1: every rank must replay it the SAME number of times)
    t_ms = torch.tensor([head["ms_per_step"]], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t_ms, op=dist.ReduceOp.MAX)
    for _ in range(int(1500.0 / max(float(t_ms[0]), 0.05)) + 1):
        graph.replay()
    torch.cuda.synchronize()
    clocks = sampler.stop() if rank == 0 else None
    if clocks is not None:
        clocks["window"] = "timed region + e2e region + 1.5 s continuation of the headline step"
    del dec, graph
    torch.cuda.empty_cache()

    extra = []
    if not args.no_extra_batches:
        for b in EXTRA_BATCHES:
            r = run_decode(ops, torch, dist, world, rank, car, b, min(args.steps, 50), warmup, barrier, e2e=False)
            r.pop("_dec"), r.pop("_graph")
            extra.append(r)
            torch.cuda.empty_cache()

    attention = []
    if not args.no_attention and rank == 0:
        for fp8 in (False, True):
            attention.append(run_attention(ops, torch, dev, fp8, min(args.steps, 100), warmup,
                                           torch.cuda.synchronize, peak))
            torch.cuda.empty_cache()
    barrier()

    keys = ["ms_per_step", "e2e_ms_per_step"]
    t = torch.tensor([head[k] for k in keys] + [r["ms_per_step"] for r in extra], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    vals = t.tolist()
    head["ms_per_step"], head["e2e_ms_per_step"] = vals[0], vals[1]
    for r, v in zip(extra, vals[2:]):
        r["ms_per_step"] = v
    if rank == 0:
        ms = head["ms_per_step"]
        per_gpu_gbs = head["bytes_per_gpu"] / (ms * 1e-3) / 1e9
        dom["share_of_step"] = dom["us_per_launch"] * 32 / (ms * 1e3)
        dom_attn["share_of_step"] = dom_attn["us_per_launch"] * 32 / (ms * 1e3)
        # the dominant kernel is whichever of the two has the larger share of the step
        top, second = (dom_attn, dom) if dom_attn["share_of_step"] >= dom["share_of_step"] else (dom, dom_attn)
        line = {
            "metric": METRIC, "value": HEADLINE_BATCH / (ms * 1e-3), "unit": "tokens/s", "n_gpus": world,
            "steps": args.steps, "warmup": warmup, "ms_per_step": ms, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "w4a16: int4 weights, f16 activations, fp32 accumulate",
            "data": "synthetic: random-init packed int4 weights + scales, random KV, random token ids",
            "config": config,
            "tokens_per_s_per_gpu": HEADLINE_BATCH / (ms * 1e-3) / world,
            "roofline": {"bound": "hbm", "achieved": top["achieved"], "peak": peak, "unit": "GB/s",
                         "frac": top["frac"], "traffic": step_traffic(top), "peak_kind": peak_kind,
                         "what": "dominant kernel of the step, timed alone with CUDA events inside this run: "
                                 "algorithmic bytes per launch / average launch duration",
                         "dominant_kernel": top, "second_kernel": second,
                         "step": {"what": "whole decode step on one GPU: (packed weights + scales + KV + lm_head "
                                          "shard bytes) / step time",
                                  "achieved": per_gpu_gbs, "frac": per_gpu_gbs / peak,
                                  "algorithmic_bytes_per_step_per_gpu": head["bytes_per_gpu"],
                                  "linear_bytes": head["linear_bytes_per_gpu"], "kv_bytes": head["kv_bytes_per_gpu"],
                                  "lm_head_bytes": head["lm_head_bytes_per_gpu"]}},
            "e2e": {"value": HEADLINE_BATCH / (head["e2e_ms_per_step"] * 1e-3), "unit": "tokens/s",
                    "h2d_bytes_per_step": head["h2d_bytes_per_step"], "d2h_bytes_per_step": head["d2h_bytes_per_step"],
                    "ms_per_step": head["e2e_ms_per_step"]},
            "gpu_launches": head["launches_per_step"] * args.steps,
            "launches_per_step": head["launches_per_step"],
            "all_reduce": (("one_shot_nvlink (fused with add+RMSNorm where eligible), NCCL above its limit"
                            if car is not None else "nccl") if world > 1 else None),
            "batches": [{"batch": r["batch"], "tokens_per_s": r["batch"] / (r["ms_per_step"] * 1e-3),
                         "ms_per_step": r["ms_per_step"],
                         "hbm_frac": r["bytes_per_gpu"] / (r["ms_per_step"] * 1e-3) / 1e9 / peak,
                         "attention_op": r["attention_op"]} for r in extra],
            "attention": attention, "checks": checks, "clocks": clocks,
        }
        if not args.no_cpu_baseline and world == 1:
            r = cpu_reference(2, 1)
            line["cpu_baseline"] = ({"value": r["tokens_per_s"], "unit": "tokens/s", "cores": r["cores"],
                                     "kind": "port", "sample": r["sample"]} if "unavailable" not in r else
                                    {"value": None, "unit": "tokens/s", "cores": 0, "kind": "port",
                                     "sample": "unavailable: " + str(r["unavailable"])})
        print(json.dumps(line), flush=True)
    if world > 1:
        # orderly teardown: graphs that hold captured collectives go first, then the IPC buffers
        # (after every rank has stopped using its peers'), then NCCL
        torch.cuda.synchronize()
        dist.barrier()
        if car is not None:
            car.close()
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
