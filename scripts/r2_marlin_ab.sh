#!/bin/bash
# A/B of the two Marlin decode kernels (B200_MARLIN_DECODE_V1=1 selects the first-generation one)
mkdir -p gpurun_out/ab
timeout 600 python -m pytest tests/test_quant_gpu.py tests/test_baseline_shapes_gpu.py -x -q -m gpu -k "marlin and not 24" > gpurun_out/ab/pytest_marlin.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/ab/pytest_marlin.log
timeout 300 python benchmarks/gemm_bench.py --op marlin --ms 1,16,64 > gpurun_out/ab/gemm_v2.jsonl 2> gpurun_out/ab/gemm_v2.err; echo "v2 rc=$?"
B200_MARLIN_DECODE_V1=1 timeout 300 python benchmarks/gemm_bench.py --op marlin --ms 1,16,64 > gpurun_out/ab/gemm_v1.jsonl 2> gpurun_out/ab/gemm_v1.err; echo "v1 rc=$?"
python - <<'PY'
import json
for f in ("v2","v1"):
    print(f)
    for l in open(f"gpurun_out/ab/gemm_{f}.jsonl"):
        if l.startswith("{"):
            d=json.loads(l); print(d["layer"], d["m"], round(d["us"],1), round(d["hbm_frac"],3))
PY
