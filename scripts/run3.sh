timeout 300 ncu --set full --import-source on --clock-control none -k regex:marlin_gemm -s 6 -c 1 -o gpurun_out/marlin_m16_gateup_d -f python benchmarks/gemm_bench.py --op marlin --ms 16 --layers gate_up --iters 4 --no-graph > gpurun_out/ncu_marlin_d.log 2>&1
tail -3 gpurun_out/ncu_marlin_d.log
