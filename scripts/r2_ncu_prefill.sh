#!/bin/bash
# one ncu --set full capture of the prefill attention kernel (1 x 8192, bf16), after a plain run exits 0
set -e
timeout 300 python benchmarks/prefill_attention_bench.py --shape 1x8192 --iters 3
timeout 900 ncu --set full --clock-control none --import-source on -k regex:varlen_attention -s 3 -c 1 \
  -o gpurun_out/prefill_attention_8k python benchmarks/prefill_attention_bench.py --shape 1x8192 --iters 3 > gpurun_out/ncu_prefill.log 2>&1
tail -3 gpurun_out/ncu_prefill.log
