mkdir -p gpurun_out/f
timeout 900 python -m pytest tests -m gpu -q --timeout 120 > gpurun_out/f/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/f/pytest_gpu.log; tail -3 gpurun_out/f/pytest_gpu.log
timeout 300 python __graft_entry__.py --smoke > gpurun_out/f/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/f/smoke.log
timeout 400 python bench.py > gpurun_out/f/bench_fp16.json 2> gpurun_out/f/bench_fp16.err; echo "bench rc=$?"
timeout 400 python bench.py --kv fp8 --no-cpu-baseline > gpurun_out/f/bench_fp8.json 2> gpurun_out/f/bench_fp8.err
timeout 300 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/f/bench_ref.json 2> gpurun_out/f/bench_ref.err
timeout 600 python benchmarks/gemm_bench.py --op all > gpurun_out/f/gemm_all.jsonl 2> gpurun_out/f/gemm_all.err; echo "gemm rc=$?"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/f/launches.csv python bench.py --steps 20 --warmup 3 --profile > gpurun_out/f/ncu_launch.log 2>&1
timeout 300 ncu --set full --import-source on --clock-control none -k regex:marlin_gemm -s 6 -c 1 -o gpurun_out/f/marlin_m16_gateup -f python benchmarks/gemm_bench.py --op marlin --ms 16 --layers gate_up --iters 4 --no-graph > gpurun_out/f/ncu1.log 2>&1
timeout 300 ncu --set full --import-source on --clock-control none -k regex:marlin24 -s 6 -c 1 -o gpurun_out/f/marlin24_m16_gateup -f python benchmarks/gemm_bench.py --op marlin24 --ms 16 --layers gate_up --iters 4 --no-graph > gpurun_out/f/ncu2.log 2>&1
timeout 300 ncu --set full --import-source on --clock-control none -k regex:scaled_mm -s 4 -c 1 -o gpurun_out/f/scaled_mm_m2048 -f python benchmarks/gemm_bench.py --op scaled_mm --ms 2048 --layers gate_up --iters 4 --no-graph > gpurun_out/f/ncu3.log 2>&1
ls gpurun_out/f
