#!/bin/bash
# paged attention: 3-warp CTAs (four per SM) where that saves a wave; batch sweep, then the decode step
for cfg in "B200_ATTN_V1_SPLIT=0 B200_ATTN_WARPS=4" "B200_ATTN_V1_SPLIT=0 B200_ATTN_WARPS3_MAX_WAVES=9" "B200_ATTN_V1_SPLIT=0"; do
  echo "== $cfg"
  for b in 64 128 160 192 256 384; do
    env $cfg timeout 200 python benchmarks/attn_shape_bench.py --batch $b 2>/dev/null | head -1 | python -c '
import sys, json
d = json.loads(sys.stdin.readline()); print("  b%d %.1f us (%.2f)" % (d["batch"], d["us"], d["hbm_frac"]), end="")'
  done
  echo
done
for cfg in "B200_ATTN_V1_SPLIT=0" "B200_ATTN_WARPS=4"; do
  echo "== decode_bench $cfg"
  env $cfg timeout 400 python benchmarks/decode_bench.py --batches 32,64,128,256 --steps 30 2>/dev/null | grep "^{" | python -c '
import sys, json
print(" ".join("b%d:%.3fms" % (d["batch"], d["ms_per_step"]) for d in map(json.loads, sys.stdin) if "batch" in d))'
done
