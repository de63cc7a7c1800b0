mkdir -p gpurun_out/f3
timeout 1500 python -m pytest tests -m gpu -q --timeout 300 > gpurun_out/f3/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/f3/pytest_gpu.log; tail -3 gpurun_out/f3/pytest_gpu.log
timeout 300 python __graft_entry__.py --smoke > gpurun_out/f3/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/f3/smoke.log | cut -c1-120
timeout 400 python bench.py > gpurun_out/f3/bench_fp16.json 2> gpurun_out/f3/bench_fp16.err; echo "bench rc=$?"; cut -c1-200 gpurun_out/f3/bench_fp16.json
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/f3/bench_ref.json 2> gpurun_out/f3/bench_ref.err; echo "ref rc=$?"
timeout 600 python benchmarks/gemm_bench.py --op all > gpurun_out/f2/gemm_all.jsonl 2> gpurun_out/f2/gemm_all.err; echo "gemm rc=$?"
timeout 300 python benchmarks/decode_bench.py --batches 1,8,64,256 --steps 30 > gpurun_out/f2/decode_tp1_fused_pdl.jsonl 2> gpurun_out/f2/decode_tp1.err; echo "decode rc=$?"; grep "^{" gpurun_out/f2/decode_tp1_fused_pdl.jsonl | cut -c80-170
timeout 300 ncu --set full --import-source on --clock-control none -k regex:marlin_gemm -s 6 -c 1 -o gpurun_out/f2/marlin_m16_gateup -f python benchmarks/gemm_bench.py --op marlin --ms 16 --layers gate_up --iters 4 --no-graph > gpurun_out/f2/ncu4.log 2>&1
