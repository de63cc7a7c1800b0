N=${1:-2}
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 benchmarks/decode_bench.py --batches 1,8,64,256 --steps 30 > gpurun_out/decode_tp${N}_r1.jsonl 2> gpurun_out/decode_tp${N}_r1.err
echo rc=$?; tail -5 gpurun_out/decode_tp${N}_r1.err; cut -c1-250 gpurun_out/decode_tp${N}_r1.jsonl
