timeout 600 python benchmarks/decode_bench.py --batches 1,8,64,256 --steps 30 > gpurun_out/decode_tp1_r1.jsonl 2> gpurun_out/decode_tp1_r1.err
echo rc=$?; tail -5 gpurun_out/decode_tp1_r1.err; cat gpurun_out/decode_tp1_r1.jsonl | cut -c1-330
