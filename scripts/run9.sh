timeout 900 python -m pytest tests/test_glue_gpu.py -m gpu -x -q > gpurun_out/glue_pytest.log 2>&1
echo "pytest rc=$?"; tail -5 gpurun_out/glue_pytest.log
timeout 600 python -m pytest tests/test_quant_gpu.py -m gpu -x -q -k "marlin" > gpurun_out/marlin_pytest.log 2>&1
echo "pytest marlin rc=$?"; tail -3 gpurun_out/marlin_pytest.log
timeout 300 python benchmarks/decode_bench.py --batches 1,8,64,256 --steps 30 --glue fused > gpurun_out/decode_tp1_r1d_pdl.jsonl 2> gpurun_out/decode_tp1_r1d_pdl.err
echo "pdl rc=$?"; grep "^{" gpurun_out/decode_tp1_r1d_pdl.jsonl | cut -c1-200
B200_PDL=0 timeout 300 python benchmarks/decode_bench.py --batches 1,8,64,256 --steps 30 --glue fused > gpurun_out/decode_tp1_r1d_nopdl.jsonl 2> gpurun_out/decode_tp1_r1d_nopdl.err
echo "nopdl rc=$?"; grep "^{" gpurun_out/decode_tp1_r1d_nopdl.jsonl | cut -c1-200
timeout 300 python benchmarks/gemm_bench.py --op marlin > gpurun_out/gemm_marlin_r1i.jsonl 2> gpurun_out/gemm_marlin_r1i.err
echo "gemm rc=$?"; grep "^{" gpurun_out/gemm_marlin_r1i.jsonl | cut -c1-220 | head -12
