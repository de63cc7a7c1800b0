#!/bin/bash
# incremental in-tree build of libb200vllm.so + _C.so (nvcc cross-compiles without a GPU)
cd "$(dirname "$0")/.." && python neural_magic_vllm_b200/build.py
