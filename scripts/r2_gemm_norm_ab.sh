#!/bin/bash
# gptq_marlin_gemm_add_rms_norm: parity against the unfused pair, then the decode step with / without it
timeout 900 python -m pytest tests/test_glue_gpu.py -x -q -m gpu -k "add_rms_norm" 2>&1 | tail -3
for flag in "--fuse-gemm-norm" ""; do
  echo "== decode_bench $flag"
  timeout 400 python benchmarks/decode_bench.py --batches 1,16,64,256 --steps 30 $flag 2>/dev/null | grep "^{" | python -c '
import sys, json
print(" ".join("b%d:%.3fms" % (d["batch"], d["ms_per_step"]) for d in map(json.loads, sys.stdin) if "batch" in d))'
done
