#!/bin/bash
# gptq_marlin_24_gemm quick check: golden bit-exact + m sweep tests, then the bench table
timeout 600 python -m pytest tests/test_quant_gpu.py -x -q -m gpu -k "marlin_24" 2>&1 | tail -2
timeout 600 python benchmarks/gemm_bench.py --op marlin24 --ms 1,16,64 > gpurun_out/marlin24.jsonl 2> gpurun_out/marlin24.err; echo "bench rc=$?"
grep "^{" gpurun_out/marlin24.jsonl | python -c '
import sys, json
for l in sys.stdin:
    d = json.loads(l); print(d["layer"], d["m"], "us", round(d["us"], 1), "hbm_frac", round(d.get("hbm_frac", 0), 3))'
