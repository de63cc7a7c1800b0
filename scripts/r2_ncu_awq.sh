#!/bin/bash
# one ncu --set full capture of awq_gemm_kernel (gate_up, M = 16) after the plain run exits 0
set -e
timeout 300 python benchmarks/gemm_bench.py --op awq --ms 16 --layers gate_up --iters 5
timeout 900 ncu --set full --clock-control none --import-source on -k regex:awq_gemm_kernel -s 8 -c 1 \
  -o gpurun_out/awq_gate_up_m16 python benchmarks/gemm_bench.py --op awq --ms 16 --layers gate_up --iters 5 --no-graph > gpurun_out/ncu_awq.log 2>&1
tail -2 gpurun_out/ncu_awq.log
