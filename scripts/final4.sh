mkdir -p gpurun_out/f2
timeout 600 python benchmarks/gemm_bench.py --op all > gpurun_out/f2/gemm_all.jsonl 2> gpurun_out/f2/gemm_all.err; echo "gemm rc=$?"
timeout 300 python benchmarks/decode_bench.py --batches 1,8,64,256 --steps 30 > gpurun_out/f2/decode_tp1_fused_pdl.jsonl 2> gpurun_out/f2/decode_tp1.err; echo "decode rc=$?"; grep "^{" gpurun_out/f2/decode_tp1_fused_pdl.jsonl | cut -c80-170
timeout 300 ncu --set full --import-source on --clock-control none -k regex:marlin_gemm -s 6 -c 1 -o gpurun_out/f2/marlin_m16_gateup -f python benchmarks/gemm_bench.py --op marlin --ms 16 --layers gate_up --iters 4 --no-graph > gpurun_out/f2/ncu4.log 2>&1; echo "ncu rc=$?"
