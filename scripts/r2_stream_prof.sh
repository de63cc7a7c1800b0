#!/bin/bash
# rebuild with cycle accounting ON THE BOX (the shipped .so stays clean), then print the breakdown
mkdir -p gpurun_out/prof
B200_EXTRA_NVCC_FLAGS=-DB200_STREAM_PROF python - <<'PY' > gpurun_out/prof/build.log 2>&1
import os, importlib.util
os.utime("neural_magic_vllm_b200/csrc/marlin_gemm.cu")
spec = importlib.util.spec_from_file_location("b", "neural_magic_vllm_b200/build.py"); m = importlib.util.module_from_spec(spec); spec.loader.exec_module(m); m.build(verbose=True)
PY
echo "build rc=$?"
for M in ${MS:-16}; do for L in ${LAYERS:-gate_up o}; do echo "== M=$M layer=$L"; timeout 120 python benchmarks/stream_prof.py --m $M --layer $L 2>&1 | tee gpurun_out/prof/prof_${L}_m$M.txt; done; done
