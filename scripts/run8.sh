timeout 900 python -m pytest tests/test_glue_gpu.py tests/test_attention_gpu.py -m gpu -x -q > gpurun_out/glue_pytest.log 2>&1
echo "pytest rc=$?"; tail -5 gpurun_out/glue_pytest.log
timeout 300 python benchmarks/decode_bench.py --batches 1,8,64,256 --steps 30 --glue fused > gpurun_out/decode_tp1_r1c_fused.jsonl 2> gpurun_out/decode_tp1_r1c_fused.err
echo "fused rc=$?"; grep "^{" gpurun_out/decode_tp1_r1c_fused.jsonl | cut -c1-200
timeout 300 python benchmarks/decode_bench.py --batches 1,8,64,256 --steps 30 --glue torch > gpurun_out/decode_tp1_r1c_torch.jsonl 2> gpurun_out/decode_tp1_r1c_torch.err
echo "torch rc=$?"; grep "^{" gpurun_out/decode_tp1_r1c_torch.jsonl | cut -c1-200
