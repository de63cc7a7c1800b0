#!/bin/bash
# decode tokens/s with the default kernel choice vs the first-generation kernel everywhere
mkdir -p gpurun_out/dec
for V in 0 1 2; do
  B200_MARLIN_DECODE_V1=$V timeout 400 python benchmarks/decode_bench.py --batches ${BATCHES:-1,16,64,256} --steps 30 > gpurun_out/dec/decode_choice$V.jsonl 2> gpurun_out/dec/decode_choice$V.err; echo "choice $V rc=$?"
  grep "^{" gpurun_out/dec/decode_choice$V.jsonl | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print({k:(round(v,2) if isinstance(v,float) else v) for k,v in d.items() if k in ('batch','tokens_per_s','ms_per_step','tp')})"
done
