#!/bin/bash
# usage: scripts/gpu_kernels.sh "<-k expression>"   (each expression in its own process)
mkdir -p gpurun_out; : > gpurun_out/kernels.log
for t in "$@"; do
  echo "=== $t" >> gpurun_out/kernels.log
  timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -q -s -k "$t" --timeout 300 2>&1 | tail -30 >> gpurun_out/kernels.log
done
grep -E "===|passed|failed|probe|assert|Error" gpurun_out/kernels.log | head -60
