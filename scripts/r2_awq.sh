#!/bin/bash
# awq_gemm / awq_dequantize: parity tests, then the Llama-3-8B layer bench beside the reference's kernel
timeout 600 python -m pytest tests/test_awq_gpu.py -x -q -m gpu 2>&1 | tail -2
timeout 600 python benchmarks/gemm_bench.py --op awq --awq-ref --ms ${MS:-1,16,64,128} > gpurun_out/awq.jsonl 2> gpurun_out/awq.err; echo "bench rc=$?"
python - <<'PY'
import json
for l in open("gpurun_out/awq.jsonl"):
    if l.startswith("{"):
        d = json.loads(l)
        print(d["layer"], d["m"], "us %.1f" % d["us"], "hbm %.3f" % d["hbm_frac"], "ref %.1f" % d.get("reference_cuda_us", 0))
PY
