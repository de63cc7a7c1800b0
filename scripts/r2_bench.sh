#!/bin/bash
# round-2 record: full GPU test suite, smoke, bench.py (N=1), reference arm, prefill attention bench
mkdir -p gpurun_out/r2b
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r2b/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2b/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2b/smoke.log 2>&1; echo "smoke rc=$?"
timeout 1200 python bench.py > gpurun_out/r2b/bench_n1.json 2> gpurun_out/r2b/bench_n1.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2b/bench_reference.json 2> gpurun_out/r2b/bench_reference.err; echo "ref rc=$?"
timeout 300 python benchmarks/prefill_attention_bench.py > gpurun_out/r2b/prefill_attention.jsonl 2>&1; echo "fa rc=$?"
timeout 300 python benchmarks/decode_bench.py --batches 1,8,16,64,256 --steps 30 > gpurun_out/r2b/decode_tp1.jsonl 2>/dev/null; echo "dec rc=$?"
