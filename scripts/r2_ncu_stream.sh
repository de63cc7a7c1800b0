#!/bin/bash
# ncu --set full of the stream-K Marlin decode kernel, gate_up M=16 (run plain first: profiling recipe)
mkdir -p gpurun_out/ncu
CMD="python benchmarks/gemm_bench.py --op marlin --ms ${M:-16} --layers ${LAYER:-gate_up} --iters 4 --no-graph"
timeout 300 $CMD > gpurun_out/ncu/plain.log 2>&1 &&
timeout 600 ncu --set full --import-source on --clock-control none -k regex:${KERN:-stream_kernel} -s 6 -c 1 -f -o gpurun_out/ncu/${OUT:-stream_m16_gateup} $CMD > gpurun_out/ncu/ncu.log 2>&1
echo "rc=$?"; tail -3 gpurun_out/ncu/ncu.log
