#!/bin/bash
# gptq_marlin_24_gemm: split-K factor sweep (B200_M24_SPLITS is read once per process)
for sp in 0 1 2 4 6 8 12; do
  echo "== splits $sp (0 = heuristic)"
  B200_M24_SPLITS=$sp timeout 300 python benchmarks/gemm_bench.py --op marlin24 --ms 1,64 2>/dev/null | grep "^{" | python -c '
import sys, json
print(" ".join("%s/%d:%.1f" % (d["layer"], d["m"], d["us"]) for d in map(json.loads, sys.stdin)))'
done
