timeout 1500 python -m pytest tests/test_glue_gpu.py tests/test_quant_gpu.py -m gpu -q > gpurun_out/glue_quant_pytest.log 2>&1
echo "pytest rc=$?"; tail -6 gpurun_out/glue_quant_pytest.log
timeout 300 python benchmarks/decode_bench.py --batches 1,8,64,256 --steps 30 --glue fused > gpurun_out/decode_tp1_r1f.jsonl 2> gpurun_out/decode_tp1_r1f.err
echo "decode rc=$?"; grep "^{" gpurun_out/decode_tp1_r1f.jsonl | cut -c1-200
timeout 300 python benchmarks/glue_bench.py --tokens 1,64,4096 > gpurun_out/glue_bench_r1.jsonl 2> gpurun_out/glue_bench_r1.err
echo "glue bench rc=$?"; grep rotary_embedding_and_cache gpurun_out/glue_bench_r1.jsonl | cut -c1-200
