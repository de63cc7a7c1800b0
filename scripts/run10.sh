timeout 1200 python -m pytest tests/test_glue_gpu.py tests/test_attention_gpu.py -m gpu -x -q > gpurun_out/glue_attn_pytest.log 2>&1
echo "pytest rc=$?"; tail -4 gpurun_out/glue_attn_pytest.log
timeout 300 python benchmarks/decode_bench.py --batches 1,8,64,256 --steps 30 --glue fused > gpurun_out/decode_tp1_r1e_pdl.jsonl 2> gpurun_out/decode_tp1_r1e_pdl.err
echo "pdl rc=$?"; grep "^{" gpurun_out/decode_tp1_r1e_pdl.jsonl | cut -c1-200
timeout 300 python benchmarks/glue_bench.py > gpurun_out/glue_bench_r1.jsonl 2> gpurun_out/glue_bench_r1.err
echo "glue bench rc=$?"; cat gpurun_out/glue_bench_r1.jsonl | cut -c1-200; tail -3 gpurun_out/glue_bench_r1.err
timeout 300 python bench.py --steps 50 --warmup 5 > gpurun_out/bench_r1e.json 2> gpurun_out/bench_r1e.err
echo "bench rc=$?"; cut -c1-600 gpurun_out/bench_r1e.json
