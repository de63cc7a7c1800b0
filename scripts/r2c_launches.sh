#!/bin/bash
# ncu launch list (device time per kernel; cold-cache, serialised: compare SHARES) of the bench's
# headline step: `bench.py --profile` = headline batch only, eager (no graph), warm-up + timed steps
set -e
CMD="python bench.py --profile --steps 2 --warmup 1"
timeout 600 $CMD > gpurun_out/r2c_profile_plain.log 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv \
  --log-file gpurun_out/r2c_bench_launches.csv $CMD > gpurun_out/r2c_profile_ncu.log 2>&1
tail -2 gpurun_out/r2c_profile_ncu.log
