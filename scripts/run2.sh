set -x
timeout 600 python -m pytest tests/test_quant_gpu.py -m gpu -q --timeout 120 -x -k "marlin" 2>&1 | tail -25
timeout 300 python benchmarks/gemm_bench.py --op marlin --ms 1,16,64,256,2048 > gpurun_out/gemm_marlin_r1e.jsonl 2> gpurun_out/gemm_marlin_r1e.err
tail -2 gpurun_out/gemm_marlin_r1e.err
