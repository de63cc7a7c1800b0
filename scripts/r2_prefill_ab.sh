#!/bin/bash
# prefill attention: parity + the 8k / 8x2k shapes for each polynomial-exp2 share (B200_FA_POLY of 16 pairs)
for p in 0 4 6 8; do
  echo "== B200_FA_POLY=$p"
  B200_FA_POLY=$p timeout 600 python -m pytest tests/test_prefill_attention_gpu.py -x -q -m gpu 2>&1 | tail -1
  for sh in 1x8192 8x2048 1x2048; do
    B200_FA_POLY=$p timeout 300 python benchmarks/prefill_attention_bench.py --shape $sh | python -c '
import sys, json
for l in sys.stdin:
    if l.startswith("{"):
        d = json.loads(l); print(d["batch"], d["seqlen"], "us", round(d["us"], 1), "TF", round(d["TFLOPs"]))'
  done
done
