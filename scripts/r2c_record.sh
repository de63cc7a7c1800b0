#!/bin/bash
# round-2 final record (r2c): full GPU suite, smoke, bench.py N=1 + reference arm, the micro benches
O=gpurun_out/r2c; mkdir -p $O
timeout 1800 python -m pytest tests -x -q -m gpu > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 $O/smoke.log
timeout 1200 python bench.py > $O/bench_n1.json 2> $O/bench_n1.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_reference.json 2> $O/bench_reference.err; echo "ref rc=$?"
timeout 300 python benchmarks/prefill_attention_bench.py > $O/prefill_attention.jsonl 2>&1; echo "fa rc=$?"
timeout 600 python benchmarks/gemm_bench.py --op marlin24 --ms 1,16,64 > $O/marlin24.jsonl 2>/dev/null; echo "m24 rc=$?"
timeout 600 python benchmarks/gemm_bench.py --op awq --awq-ref --ms 1,16,64,256 > $O/awq.jsonl 2>/dev/null; echo "awq rc=$?"
timeout 600 python benchmarks/gemm_bench.py --op marlin --ms 1,16,64,2048 > $O/marlin.jsonl 2>/dev/null; echo "marlin rc=$?"
timeout 400 python benchmarks/decode_bench.py --batches 1,8,16,32,64,256 --steps 30 > $O/decode_tp1.jsonl 2>/dev/null; echo "dec rc=$?"
timeout 900 python benchmarks/vs_ref_cuda.py > $O/vs_ref_cuda.jsonl 2>/dev/null; echo "vsref rc=$?"
grep "^{" $O/bench_n1.json | cut -c1-400
