"""One-shot NVLink all-reduce (csrc/custom_all_reduce.cu, custom_all_reduce.py).

CPU: the host-side contract of the reference's CustomAllreduce that needs no device (world size 1 and
unsupported world sizes disable it and `custom_all_reduce` returns None = "use NCCL";
custom_all_reduce.py:66-77, 241-262 of the reference), over a real gloo group of 2 processes.
GPU (needs >= 2 devices; skipped on a 1-GPU box): benchmarks/ar_check.py under torchrun --
bit-identical results on all ranks, bit-exact against the rank-order fp32 sum, within tolerance of
NCCL, eager and graph-replayed with sizes alternating back to back."""
import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

_WORKER = r"""
import os, sys
sys.path.insert(0, os.environ["B200_ROOT"])
import torch, torch.distributed as dist
dist.init_process_group("gloo")
from neural_magic_vllm_b200.custom_all_reduce import CustomAllreduce
# world size 2 on a machine without CUDA: construction must fail loudly (no silent CPU path) ...
if not torch.cuda.is_available():
    try:
        CustomAllreduce(None, "cuda:0")
        raise SystemExit("CustomAllreduce built without a CUDA device")
    except (RuntimeError, AssertionError):
        pass
# ... while a 1-rank group is simply disabled, as in the reference
g1 = dist.new_group([0]); g2 = dist.new_group([1])
mine = g1 if dist.get_rank() == 0 else g2
car = CustomAllreduce(mine, "cuda:0")
assert car.disabled and car.custom_all_reduce(torch.zeros(8)) is None and not car.should_custom_ar(torch.zeros(8))
dist.barrier()
print("ok")
"""


def test_custom_allreduce_host_contract_gloo(tmp_path):
    script = tmp_path / "w.py"
    script.write_text(_WORKER)
    env = dict(os.environ, B200_ROOT=ROOT, CUDA_VISIBLE_DEVICES="")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                        "--master-addr", "127.0.0.1", "--master-port", "29577", str(script)],
                       capture_output=True, text=True, timeout=300, env=env)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert r.stdout.count("ok") == 2


@pytest.mark.gpu
def test_one_shot_all_reduce_two_gpus():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (run: gpurun --gpus 2 -- python -m pytest tests/test_custom_ar.py -m gpu)")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                        "--master-addr", "127.0.0.1", "--master-port", "29578",
                        os.path.join(ROOT, "benchmarks", "ar_check.py")],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert '"all_reduce_check": "ok"' in r.stdout


@pytest.mark.gpu
def test_two_shot_all_reduce_two_gpus():
    """The two-shot path (taken from world 4 up by default) forced at world 2: the same check matrix --
    bit-identical on every rank, equal to the fp32 rank-order sum rounded once, plain and fused with
    add + RMSNorm, eager and under graph replay."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (run: gpurun --gpus 2 -- python -m pytest tests/test_custom_ar.py -m gpu)")
    env = dict(os.environ, B200_AR_TWO_SHOT="1")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                        "--master-addr", "127.0.0.1", "--master-port", "29579",
                        os.path.join(ROOT, "benchmarks", "ar_check.py")],
                       capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert '"all_reduce_check": "ok"' in r.stdout
