#!/bin/bash
# round-2 evidence run: full GPU test suite, smoke, GEMM microbench table
mkdir -p gpurun_out/full
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/full/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/full/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/full/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/full/smoke.log
timeout 900 python benchmarks/gemm_bench.py --op all > gpurun_out/full/gemm_all.jsonl 2> gpurun_out/full/gemm_all.err; echo "gemm rc=$?"
python - <<'PY'
import json
for l in open("gpurun_out/full/gemm_all.jsonl"):
    if l.startswith("{"):
        d=json.loads(l); print(d["op"][:14], d.get("shard",""), d["layer"], d["m"], f'{d["k"]}x{d["n"]}', round(d["us"],1), round(d["hbm_frac"],3), round(d["tensor_frac"],3))
PY
