mkdir -p gpurun_out/f2
timeout 1500 python -m pytest tests -m gpu -q --timeout 300 > gpurun_out/f2/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/f2/pytest_gpu.log; tail -3 gpurun_out/f2/pytest_gpu.log
timeout 300 python __graft_entry__.py --smoke > gpurun_out/f2/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/f2/smoke.log
timeout 400 python bench.py > gpurun_out/f2/bench_fp16.json 2> gpurun_out/f2/bench_fp16.err; echo "bench rc=$?"
timeout 400 python bench.py --kv fp8 --no-cpu-baseline > gpurun_out/f2/bench_fp8.json 2> gpurun_out/f2/bench_fp8.err
timeout 300 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/f2/bench_ref.json 2> gpurun_out/f2/bench_ref.err
timeout 300 python benchmarks/glue_bench.py > gpurun_out/f2/glue_bench.jsonl 2> gpurun_out/f2/glue_bench.err; echo "glue rc=$?"
timeout 600 python benchmarks/gemm_bench.py --op all > gpurun_out/f2/gemm_all.jsonl 2> gpurun_out/f2/gemm_all.err; echo "gemm rc=$?"
timeout 300 python benchmarks/decode_bench.py --batches 1,8,64,256 --steps 30 > gpurun_out/f2/decode_tp1_fused_pdl.jsonl 2> gpurun_out/f2/decode_tp1.err; echo "decode rc=$?"
B200_PDL=0 timeout 300 python benchmarks/decode_bench.py --batches 1,8,64,256 --steps 30 > gpurun_out/f2/decode_tp1_fused_nopdl.jsonl 2>> gpurun_out/f2/decode_tp1.err
timeout 300 python benchmarks/decode_bench.py --batches 1,8,64,256 --steps 30 --glue torch > gpurun_out/f2/decode_tp1_torchglue.jsonl 2>> gpurun_out/f2/decode_tp1.err
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/f2/launches.csv python bench.py --steps 20 --warmup 3 --profile > gpurun_out/f2/ncu_launch.log 2>&1
timeout 300 ncu --set full --import-source on --clock-control none -k regex:rms_norm_kernel -s 2 -c 1 -o gpurun_out/f2/glue_add_rms -f python benchmarks/glue_bench.py --tokens 16384 --iters 2 > gpurun_out/f2/ncu1.log 2>&1
timeout 300 ncu --set full --import-source on --clock-control none -k regex:act_kernel -s 1 -c 1 -o gpurun_out/f2/glue_silu -f python benchmarks/glue_bench.py --tokens 16384 --iters 2 > gpurun_out/f2/ncu2.log 2>&1
timeout 300 ncu --set full --import-source on --clock-control none -k regex:rotary_cache_kernel -s 1 -c 1 -o gpurun_out/f2/glue_rotary_cache -f python benchmarks/glue_bench.py --tokens 16384 --iters 2 > gpurun_out/f2/ncu3.log 2>&1
timeout 300 ncu --set full --import-source on --clock-control none -k regex:marlin_gemm -s 6 -c 1 -o gpurun_out/f2/marlin_m16_gateup -f python benchmarks/gemm_bench.py --op marlin --ms 16 --layers gate_up --iters 4 --no-graph > gpurun_out/f2/ncu4.log 2>&1
ls gpurun_out/f2
