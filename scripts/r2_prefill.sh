#!/bin/bash
# prefill attention: parity tests, then the bench table (one line per shape)
timeout 600 python -m pytest tests/test_prefill_attention_gpu.py -x -q -m gpu 2>&1 | tail -3
timeout 600 python benchmarks/prefill_attention_bench.py > gpurun_out/prefill_attention.jsonl 2>&1
grep "^{" gpurun_out/prefill_attention.jsonl | python -c '
import sys, json
for l in sys.stdin:
    d = json.loads(l)
    print(d["dtype"], d["batch"], d["seqlen"], "us", round(d["us"], 1), "TF", round(d["TFLOPs"]), "frac", round(d["tensor_frac"], 3), "sdpa", round(d["torch_sdpa_us"], 1), "fa2", round(d["flash_attn_lib_us"], 1))'
