#!/bin/bash
# incremental in-tree build of libb200vllm.so + _C.so (nvcc cross-compiles without a GPU);
# exits non-zero when a compile or link step fails
cd "$(dirname "$0")/.." && python neural_magic_vllm_b200/build.py > /tmp/b200_build.log 2>&1
rc=$?
grep -i "error" /tmp/b200_build.log | head -20
tail -1 /tmp/b200_build.log
exit $rc
