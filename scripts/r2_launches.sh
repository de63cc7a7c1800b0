#!/bin/bash
# per-kernel device time of one decode step (ncu launch list; cold-cache, serialised: compare SHARES)
mkdir -p gpurun_out/launches
for B in ${BATCHES:-1 64}; do
  CMD="python benchmarks/decode_bench.py --batches $B --steps 2 --warmup 1 --no-graph"
  timeout 300 $CMD > gpurun_out/launches/plain_b$B.log 2>&1 &&
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 1400 -c 700 --csv --log-file gpurun_out/launches/decode_b$B.csv $CMD > gpurun_out/launches/ncu_b$B.log 2>&1
  echo "b=$B rc=$?"
done
