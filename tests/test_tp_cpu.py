"""Host-side tensor-parallel logic on CPU (gloo, world_size 2): the sharding the N>1 path uses
 — attention heads and column-parallel weights split per rank, row-parallel partial
sums all-reduced — must reproduce the unsharded oracle.  The arithmetic here is the ORACLE's (this is
a test of the partitioning, not of the kernels)."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import attention_ref as ar
from oracle import quant_ref as qr


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.manual_seed(0)  # identical inputs on every rank
        heads, kv_heads, d, hidden = 8, 4, 64, 256
        q, kc, vc, bt, sl = ar.make_paged_inputs(3, heads, kv_heads, d, 16, [5, 40, 17], torch.float32, seed=0)
        full = ar.paged_attention_ref(q, kc, vc, bt, sl, d ** -0.5, None, "auto", 1.0)
        # rank r owns q heads [r*H/tp, ...) and kv heads [r*KVH/tp, ...) of the SAME block table
        h0, h1 = rank * heads // world, (rank + 1) * heads // world
        k0, k1 = rank * kv_heads // world, (rank + 1) * kv_heads // world
        part = ar.paged_attention_ref(q[:, h0:h1].contiguous(), kc[:, k0:k1].contiguous(),
                                      vc[:, k0:k1].contiguous(), bt, sl, d ** -0.5, None, "auto", 1.0)
        assert torch.allclose(part, full[:, h0:h1], atol=1e-6)
        # row-parallel o-proj: W4 group-128 weight sharded along K (scales sharded along K too)
        w = torch.randn(heads * d, hidden, dtype=torch.half)
        w_ref, _, s, _, _ = qr.quantize_weights(w, 4, 128, False)
        x = full.reshape(3, heads * d)
        ks, ke = rank * heads * d // world, (rank + 1) * heads * d // world
        y = x[:, ks:ke].float() @ w_ref[ks:ke].float()
        dist.all_reduce(y)
        ref = x.float() @ w_ref.float()
        assert torch.allclose(y, ref, rtol=1e-4, atol=1e-4)
        assert s.shape[0] == heads * d // 128  # one scale row per K group: shards cleanly with K
        ret[rank] = True
    finally:
        dist.destroy_process_group()


def test_tp_shard_gloo():
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    procs = [ctx.Process(target=_worker, args=(r, world, port, ret)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert all(ret.get(r) for r in range(world))
