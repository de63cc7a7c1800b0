timeout 1500 python -m pytest tests/test_glue_gpu.py tests/test_quant_gpu.py -m gpu -x -q > gpurun_out/glue_quant_pytest.log 2>&1
echo "pytest rc=$?"; tail -4 gpurun_out/glue_quant_pytest.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 700 -c 340 --csv --log-file gpurun_out/decode_b1_launches.csv python benchmarks/decode_bench.py --batches 1 --steps 2 --warmup 1 --no-graph > gpurun_out/ncu_decode.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/ncu_decode.log | cut -c1-200
