set -x
timeout 600 python -m pytest tests/test_quant_gpu.py -m gpu -q --timeout 120 -k "marlin_24" 2>&1 | tail -25
timeout 600 python -m pytest tests/test_quant_gpu.py -m gpu -q --timeout 120 -k "marlin and not marlin_24" 2>&1 | tail -3
timeout 300 python benchmarks/gemm_bench.py --op marlin --ms 1,16,64 > gpurun_out/gemm_marlin_r1h.jsonl 2> gpurun_out/gemm_marlin_r1h.err
tail -2 gpurun_out/gemm_marlin_r1h.err
