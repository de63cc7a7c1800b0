#!/bin/bash
# one ncu --set full capture of the step's dominant kernel: paged attention at batch 64 x ctx 1024
set -e
timeout 300 python benchmarks/attn_shape_bench.py --batch 64 --ctx 1025
timeout 900 ncu --set full --clock-control none --import-source on -k regex:paged_attn_mma_kernel -s 4 -c 1 \
  -o gpurun_out/r2c_step_attention python benchmarks/attn_shape_bench.py --batch 64 --ctx 1025 > gpurun_out/r2c_ncu_attention.log 2>&1
tail -2 gpurun_out/r2c_ncu_attention.log
