#!/bin/bash
# standalone paged attention at decode-step shapes: batch sweep, then ring-depth knob at batch 64
for b in 8 16 32 64 96 128 256; do timeout 200 python benchmarks/attn_shape_bench.py --batch $b; done > gpurun_out/attn_shape.jsonl 2>&1
for st in 2 3 4 6; do echo "stages $st" >> gpurun_out/attn_shape.jsonl; B200_ATTN_STAGES=$st timeout 200 python benchmarks/attn_shape_bench.py --batch 64 >> gpurun_out/attn_shape.jsonl 2>&1; done
python - <<'PY'
import json
for l in open("gpurun_out/attn_shape.jsonl"):
    if l.startswith("{"):
        d = json.loads(l); print(d["op"][-2:], "b", d["batch"], "us %.1f" % d["us"], "frac %.2f" % d["hbm_frac"])
    else:
        print(l.strip()[:100])
PY
