set -x
timeout 300 python benchmarks/gemm_bench.py --op marlin --ms 1,16,64,256,2048 > gpurun_out/gemm_marlin_graph.jsonl 2> gpurun_out/gemm_marlin_graph.err
timeout 300 ncu --set full --import-source on --clock-control none -k regex:marlin_gemm_kernel -s 6 -c 1 -o gpurun_out/marlin_m16_gateup -f python benchmarks/gemm_bench.py --op marlin --ms 16 --layers gate_up --iters 4 --no-graph > gpurun_out/ncu_marlin.log 2>&1
timeout 300 ncu --set full --import-source on --clock-control none -k regex:marlin_gemm_kernel -s 6 -c 1 -o gpurun_out/marlin_m2048_gateup -f python benchmarks/gemm_bench.py --op marlin --ms 2048 --layers gate_up --iters 4 --no-graph >> gpurun_out/ncu_marlin.log 2>&1
tail -3 gpurun_out/ncu_marlin.log
