"""Llama-3-8B W4A16 (gptq_marlin, group 128) decode tokens/s — BASELINE.json metric (i), SURVEY.md §8(d).

A minimal decoder (32 layers, hidden 4096, 32 q / 8 kv heads, D 128, inter 14336, vocab 128256) whose
hot ops are this repo's kernels through the reference's op surface: gptq_marlin_gemm for the four
linears, reshape_and_cache + paged_attention_v1/v2 over the block-table KV cache, and (default,
--glue fused) this repo's rms_norm / fused_add_rms_norm / rotary_embedding_and_cache / silu_and_mul
kernels for everything in between; --glue torch keeps those as plain torch
ops (the round-1a configuration) for an A/B.  The fp16 lm_head and greedy argmax are torch ops.
Scheduler, sampler and tokenizer are excluded (stated).

Tensor parallel: launched under torchrun, rank r owns q heads / kv heads / the
column shards of qkv and gate_up and the row (K) shards of o and down; the two partial sums per
layer (linear.py:792-793 of the reference) are reduced inside the captured CUDA graph by this repo's
one-shot NVLink all-reduce -- fused with the add + RMSNorm behind it where the message qualifies --
and by NCCL above its size limit (--ar nccl: NCCL only; --no-fuse-ar-norm: two kernels).

Random-init quantized weights (any int4 pattern is a valid weight), context filled with random KV.

  python benchmarks/decode_bench.py --batches 1,8,64,256 --ctx 1024 --steps 50
  python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 benchmarks/decode_bench.py
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch
import torch.distributed as dist
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

HIDDEN, HEADS, KV_HEADS, HEAD, INTER, VOCAB, LAYERS = 4096, 32, 8, 128, 14336, 128256, 32
BLOCK = 16


class QLinear:
    """One gptq_marlin W4A16 g128 linear shard with random packed weights.

    The FULL layer [k, n] is drawn from `gen` (seeded identically on every rank) in the Marlin
    layout and then sharded the way the reference shards it (linear.py:277-333, 490-560, 700-760):
    column-parallel layers keep, for every logical sub-matrix in `col_parts` (q|k|v, gate|up), the
    rank's slice of its columns; row-parallel layers keep the rank's K rows (and its scale groups).
    A Marlin tile is 16 k x 64 n (gptq_marlin_repack.cu:32-256), so column slices at multiples of 64
    and row slices at multiples of 16 are slices of the packed tensor too: [k/16, 2n] int32 and
    [k/128, n] scales (scale permutation works inside 64-column chunks, gptq_marlin.py:28-57).
    TP-N therefore computes the same function as TP-1 up to summation order."""

    def __init__(self, ops, k, n, gen, tp=1, rank=0, shard="none", col_parts=None):
        self.ops = ops
        q = torch.randint(-2**31, 2**31 - 1, (k // 16, n * 2), dtype=torch.int32, device="cuda", generator=gen)
        s = (torch.rand(k // 128, n, device="cuda", generator=gen) * 0.004 + 0.0005).half()
        if tp > 1 and shard == "col":
            qs, ss, off = [], [], 0
            for w in (col_parts or [n]):
                lo = off + rank * (w // tp)
                assert (w // tp) % 64 == 0
                qs.append(q[:, 2 * lo: 2 * (lo + w // tp)])
                ss.append(s[:, lo: lo + w // tp])
                off += w
            q, s = torch.cat(qs, 1).contiguous(), torch.cat(ss, 1).contiguous()
        elif tp > 1 and shard == "row":
            kk = k // tp
            assert kk % 128 == 0
            q = q[rank * kk // 16: (rank + 1) * kk // 16].contiguous()
            s = s[rank * kk // 128: (rank + 1) * kk // 128].contiguous()
        self.q, self.s = q, s
        self.k, self.n = q.shape[0] * 16, q.shape[1] // 2
        n = self.n
        self.e = torch.empty(0, dtype=torch.int32, device="cuda")
        self.ws = torch.zeros(max(n // 64, 1) * 16, dtype=torch.int32, device="cuda")

    def bytes(self):
        return self.q.numel() * 4 + self.s.numel() * 2

    def __call__(self, x):
        return self.ops.gptq_marlin_gemm(x, self.q, self.s, self.e, self.e, self.ws, 4, x.shape[0], self.n,
                                         self.k, True)

    def add_norm(self, x, res, w):
        """gptq_marlin_gemm + fused_add_rms_norm(out, res, w) as one operator (TP = 1)."""
        return self.ops.gptq_marlin_gemm_add_rms_norm(x, self.q, self.s, self.e, self.e, self.ws, 4, x.shape[0],
                                                      self.n, self.k, True, res, w, 1e-5)


class Layer:
    def __init__(self, ops, tp, rank, gen, num_blocks):
        assert KV_HEADS % tp == 0, "kv heads are the sharding unit (8 / tp)"
        self.h, self.kvh = HEADS // tp, KV_HEADS // tp
        self.qkv = QLinear(ops, HIDDEN, (HEADS + 2 * KV_HEADS) * HEAD, gen, tp, rank, "col",
                           [HEADS * HEAD, KV_HEADS * HEAD, KV_HEADS * HEAD])
        self.o = QLinear(ops, HEADS * HEAD, HIDDEN, gen, tp, rank, "row")
        self.gate_up = QLinear(ops, HIDDEN, 2 * INTER, gen, tp, rank, "col", [INTER, INTER])
        self.down = QLinear(ops, INTER, HIDDEN, gen, tp, rank, "row")
        self.ln1 = (torch.rand(HIDDEN, device="cuda", generator=gen) + 0.5).half()
        self.ln2 = (torch.rand(HIDDEN, device="cuda", generator=gen) + 0.5).half()
        s = HEAD ** -0.5
        # the full 8-kv-head cache content is drawn per layer, the rank keeps its heads
        # (cache_engine.py:39: [num_blocks, kv_heads / tp, ...] per rank)
        kc = torch.empty(num_blocks, KV_HEADS, HEAD // 8, BLOCK, 8, dtype=torch.half, device="cuda")
        vc = torch.empty(num_blocks, KV_HEADS, HEAD, BLOCK, dtype=torch.half, device="cuda")
        kc.uniform_(-s, s, generator=gen)
        vc.uniform_(-s, s, generator=gen)
        lo = rank * self.kvh
        self.kc = kc[:, lo: lo + self.kvh].contiguous() if tp > 1 else kc
        self.vc = vc[:, lo: lo + self.kvh].contiguous() if tp > 1 else vc

    def linear_bytes(self):
        return self.qkv.bytes() + self.o.bytes() + self.gate_up.bytes() + self.down.bytes()


def rope(x, cos, sin):  # neox style, x [B, H, D]
    x1, x2 = x[..., : HEAD // 2], x[..., HEAD // 2:]
    return torch.cat((x1 * cos - x2 * sin, x2 * cos + x1 * sin), dim=-1)


class Decoder:
    def __init__(self, ops, tp, rank, batch, ctx, lm_head=True, glue="fused", car=None, num_layers=LAYERS):
        self.ops, self.tp, self.rank, self.batch, self.ctx = ops, tp, rank, batch, ctx
        self.car = car  # one-shot NVLink all-reduce (None: NCCL)
        self.fuse_ar_norm = True
        # TP = 1: row-parallel GEMM + fused_add_rms_norm as one operator.  Off by default: measured
        # neutral to slightly slower in the step (one CTA per token row sums the split-K planes, where the
        # separate sum kernel spreads over the whole GPU and PDL already hides its launch)
        self.fuse_gemm_norm = False
        self.step = self.step_fused if glue == "fused" else self.step_torch
        # one seed for every rank: each rank draws the full tensors and keeps its shard
        gen = torch.Generator(device="cuda").manual_seed(1234)
        blocks_per_seq = (ctx + BLOCK) // BLOCK
        self.num_blocks = batch * blocks_per_seq
        self.embed = (torch.randn(4096, HIDDEN, device="cuda", generator=gen) * 0.02).half()
        perm = torch.randperm(self.num_blocks, device="cuda", generator=gen).to(torch.int32)
        self.tokens = torch.randint(0, 4096, (batch,), device="cuda", generator=gen)
        self.norm = (torch.rand(HIDDEN, device="cuda", generator=gen) + 0.5).half()
        self.layers = [Layer(ops, tp, rank, gen, self.num_blocks) for _ in range(num_layers)]
        self.lm_head = None
        if lm_head:
            full = (torch.randn(VOCAB, HIDDEN, device="cuda", generator=gen) * 0.02).half()
            vs = VOCAB // tp  # vocab-parallel lm_head (vocab_parallel_embedding.py of the reference)
            self.lm_head = full[rank * vs: (rank + 1) * vs].contiguous() if tp > 1 else full
            del full
        self.block_tables = perm.view(batch, blocks_per_seq).contiguous()
        self.seq_lens = torch.full((batch,), ctx + 1, dtype=torch.int32, device="cuda")
        last = self.block_tables[:, ctx // BLOCK].to(torch.int64)
        self.slots = last * BLOCK + ctx % BLOCK
        ang = (ctx * 500000.0 ** (-torch.arange(0, HEAD, 2, device="cuda").float() / HEAD))
        self.cos, self.sin = ang.cos().half(), ang.sin().half()
        # reference layout of the rotary table: [max_position, rot_dim] = (cos | sin) per position
        inv = 500000.0 ** (-torch.arange(0, HEAD, 2, device="cuda").float() / HEAD)
        tab = torch.arange(ctx + 1, device="cuda").float()[:, None] * inv[None]
        self.cos_sin_cache = torch.cat((tab.cos(), tab.sin()), -1).half()
        self.positions = torch.full((batch,), ctx, dtype=torch.int64, device="cuda")
        self.next_tokens = torch.zeros(batch, dtype=torch.int64, device="cuda")
        h = self.layers[0].h
        self.parts = (ctx + 1 + 511) // 512
        self.use_v1 = self.parts == 1 or batch * h > 512  # paged_attn.py:118-119 of the reference
        self.tmp = torch.empty(batch, h, self.parts, HEAD, dtype=torch.half, device="cuda")
        self.es = torch.empty(batch, h, self.parts, dtype=torch.float32, device="cuda")
        self.ml = torch.empty_like(self.es)

    def sample(self, hn):
        """Greedy sampling over the vocab-parallel lm_head: every rank takes (max, argmax) of its
        vocab shard, the [tp, batch, 2] candidates are all-gathered (NCCL, inside the graph) and
        every rank picks the global winner -- the same token the reference's gather-to-rank-0 of the
        full logits + argmax yields (logits_processor.py:59-66 + sampler greedy)."""
        logits = hn @ self.lm_head.t()
        if self.tp == 1:
            nxt = logits.argmax(-1)
        else:
            val, idx = logits.float().max(-1)
            cand = torch.stack((val, (idx + self.rank * (VOCAB // self.tp)).float()), -1)  # ids < 2^24: exact
            allc = torch.empty(self.tp, self.batch, 2, dtype=torch.float32, device="cuda")
            dist.all_gather_into_tensor(allc, cand)
            best = allc[..., 0].argmax(0)
            nxt = allc[..., 1].gather(0, best[None]).squeeze(0).long()
        self.next_tokens.copy_(nxt)
        self.tokens.copy_(nxt % 4096)  # the harness embeds through a 4096-row table (stated)

    def all_reduce(self, t):
        """Row-parallel partial sums (linear.py:792-793): one-shot NVLink kernel when the message
        qualifies, NCCL otherwise -- the same split the reference makes (parallel_state.py)."""
        if self.car is not None:
            out = self.car.custom_all_reduce(t)
            if out is not None:
                return out
        dist.all_reduce(t)
        return t

    def reduce_add_norm(self, t, res, w):
        """all_reduce(t); fused_add_rms_norm(t, res, w) -- one kernel when the message qualifies."""
        if self.car is not None and self.fuse_ar_norm and self.car.all_reduce_add_rms_norm(t, res, w, 1e-5):
            return t
        t = self.all_reduce(t)
        self.ops.fused_add_rms_norm(t, res, w, 1e-5)
        return t

    def step_fused(self):
        """Every op between the embedding gather and the lm_head is one of this repo's kernels."""
        ops, tp = self.ops, self.tp
        x = self.embed[self.tokens]
        res = None
        fuse = tp == 1 and self.fuse_gemm_norm
        for li, L in enumerate(self.layers):
            if res is None:
                res, hn = x, torch.empty_like(x)
                ops.rms_norm(hn, x, L.ln1, 1e-5)
            elif tp > 1:
                hn = x = self.reduce_add_norm(x, res, L.ln1)  # previous layer's down_proj partial sums
            elif fuse:
                hn = x  # already normalised by the previous layer's down_proj + norm operator
            else:
                ops.fused_add_rms_norm(x, res, L.ln1, 1e-5)
                hn = x
            qkv = L.qkv(hn)
            q, k, v = qkv.split([L.h * HEAD, L.kvh * HEAD, L.kvh * HEAD], dim=-1)
            ops.rotary_embedding_and_cache(self.positions, q, k, v, HEAD, self.cos_sin_cache, True, L.kc, L.vc,
                                           self.slots, "auto", 1.0)
            q = q.view(-1, L.h, HEAD)
            out = torch.empty(self.batch, L.h, HEAD, dtype=torch.half, device="cuda")
            if self.use_v1:
                ops.paged_attention_v1(out, q, L.kc, L.vc, L.kvh, HEAD ** -0.5, self.block_tables, self.seq_lens,
                                       BLOCK, self.ctx + 1, None, "auto", 1.0)
            else:
                ops.paged_attention_v2(out, self.es, self.ml, self.tmp, q, L.kc, L.vc, L.kvh, HEAD ** -0.5,
                                       self.block_tables, self.seq_lens, BLOCK, self.ctx + 1, None, "auto", 1.0)
            if fuse:
                o = L.o.add_norm(out.view(-1, L.h * HEAD), res, L.ln2)
            else:
                o = L.o(out.view(-1, L.h * HEAD))
                if tp > 1:
                    o = self.reduce_add_norm(o, res, L.ln2)
                else:
                    ops.fused_add_rms_norm(o, res, L.ln2, 1e-5)
            gu = L.gate_up(o)
            act = torch.empty(self.batch, INTER // tp, dtype=torch.half, device="cuda")
            ops.silu_and_mul(act, gu)
            if fuse:  # the norm behind down_proj is the next layer's input norm (or the final norm)
                nw = self.layers[li + 1].ln1 if li + 1 < len(self.layers) else self.norm
                x = L.down.add_norm(act, res, nw)
            else:
                x = L.down(act)  # TP: partial sums, reduced together with the next norm
        if tp > 1:
            x = self.reduce_add_norm(x, res, self.norm)
        elif not fuse:
            ops.fused_add_rms_norm(x, res, self.norm, 1e-5)
        hn = x
        if self.lm_head is not None:
            self.sample(hn)
        return hn

    def step_torch(self):
        ops, tp = self.ops, self.tp
        x = self.embed[self.tokens]
        res = x
        for L in self.layers:
            hn = F.rms_norm(res, (HIDDEN,), L.ln1, 1e-5)
            qkv = L.qkv(hn)
            q, k, v = qkv.split([L.h * HEAD, L.kvh * HEAD, L.kvh * HEAD], dim=-1)
            q = rope(q.view(-1, L.h, HEAD), self.cos, self.sin)
            k = rope(k.view(-1, L.kvh, HEAD), self.cos, self.sin)
            v = v.reshape(-1, L.kvh, HEAD)
            ops.reshape_and_cache(k, v, L.kc, L.vc, self.slots, "auto", 1.0)
            out = torch.empty_like(q)
            if self.use_v1:
                ops.paged_attention_v1(out, q, L.kc, L.vc, L.kvh, HEAD ** -0.5, self.block_tables, self.seq_lens,
                                       BLOCK, self.ctx + 1, None, "auto", 1.0)
            else:
                ops.paged_attention_v2(out, self.es, self.ml, self.tmp, q, L.kc, L.vc, L.kvh, HEAD ** -0.5,
                                       self.block_tables, self.seq_lens, BLOCK, self.ctx + 1, None, "auto", 1.0)
            o = L.o(out.view(-1, L.h * HEAD))
            if tp > 1:
                dist.all_reduce(o)
            res = res + o
            hn = F.rms_norm(res, (HIDDEN,), L.ln2, 1e-5)
            gu = L.gate_up(hn)
            act = F.silu(gu[:, : INTER // tp]) * gu[:, INTER // tp:]
            d = L.down(act)
            if tp > 1:
                dist.all_reduce(d)
            res = res + d
        hn = F.rms_norm(res, (HIDDEN,), self.norm, 1e-5)
        if self.lm_head is not None:
            self.sample(hn)
        return hn

    def linear_bytes(self):
        return sum(L.linear_bytes() for L in self.layers)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batches", default="1,8,64,256")
    ap.add_argument("--ctx", type=int, default=1024)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--no-lm-head", action="store_true")
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--glue", choices=["fused", "torch"], default="fused")
    ap.add_argument("--fuse-gemm-norm", action="store_true",
                    help="TP = 1: run the row-parallel GEMMs and the fused_add_rms_norm behind them as one operator")
    ap.add_argument("--no-fuse-ar-norm", action="store_true",
                    help="keep all_reduce and fused_add_rms_norm as two kernels")
    ap.add_argument("--ar", choices=["one_shot", "nccl"], default="one_shot",
                    help="TP all-reduce: this repo's one-shot NVLink kernel (NCCL above its max_size) or NCCL only")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    import neural_magic_vllm_b200 as pkg
    ops = pkg.ops
    hbm = 6545.6
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            hbm = float(json.load(f)["hbm_gbs"])
    except Exception:
        pass
    car = None
    if world > 1 and args.ar == "one_shot":
        from neural_magic_vllm_b200.custom_all_reduce import CustomAllreduce
        car = CustomAllreduce(None, local)
        if car.disabled:
            car = None
    for batch in [int(b) for b in args.batches.split(",")]:
        dec = Decoder(ops, world, rank, batch, args.ctx, lm_head=not args.no_lm_head, glue=args.glue, car=car)
        dec.fuse_ar_norm = not args.no_fuse_ar_norm
        dec.fuse_gemm_norm = args.fuse_gemm_norm
        with torch.no_grad():
            for _ in range(2):
                dec.step()
            torch.cuda.synchronize()
            graph = None
            if not args.no_graph:
                graph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(graph):
                    dec.step()
            run = graph.replay if graph is not None else dec.step
            for _ in range(args.warmup):
                run()
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize()
            st, en = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            st.record()
            for _ in range(args.steps):
                run()
            en.record()
            torch.cuda.synchronize()
            ms = st.elapsed_time(en) / args.steps
        t = torch.tensor([ms], device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
        if rank == 0:
            lin = dec.linear_bytes()
            kv = 2 * batch * (args.ctx + 1) * dec.layers[0].kvh * HEAD * 2 * LAYERS
            head = dec.lm_head.numel() * 2 if dec.lm_head is not None else 0
            byts = lin + kv + head
            print(json.dumps({
                "metric": "decode tokens/s/GPU Llama-3-8B W4A16 gptq_marlin g128",
                "tp": world, "batch": batch, "ctx": args.ctx, "ms_per_step": ms,
                "tokens_per_s": batch / ms * 1e3, "tokens_per_s_per_gpu": batch / ms * 1e3 / world,
                "cuda_graph": graph is not None, "lm_head": dec.lm_head is not None,
                "attention": "paged_attention_v1" if dec.use_v1 else "paged_attention_v2", "glue": args.glue,
                "all_reduce": ("one_shot_nvlink" if car is not None else "nccl") if world > 1 else None,
                "ar_norm_fused": bool(car is not None and dec.fuse_ar_norm),
                "gemm_norm_fused": bool(world == 1 and dec.fuse_gemm_norm and args.glue == "fused"),
                "hbm_bytes_per_step_per_gpu": byts, "linear_bytes_per_gpu": lin, "kv_bytes_per_gpu": kv,
                "hbm_GBps": byts / ms / 1e6, "hbm_frac": byts / ms / 1e6 / hbm,
                "excluded": "scheduler, tokenizer; sampler = greedy argmax over the vocab-parallel lm_head",
                "data": "synthetic: random-init packed int4 weights, random KV"}), flush=True)
        del dec, graph
        torch.cuda.empty_cache()
    # leave without tearing NCCL down: destroy_process_group() after graph-captured collectives
    # hung at exit on the B200 box (every number above is already printed and flushed)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    sys.stdout.flush()
    os._exit(0)


if __name__ == "__main__":
    main()
