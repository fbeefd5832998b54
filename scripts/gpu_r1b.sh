#!/bin/bash
# round-1 re-entry check: conv kernels (halo probe first), all kernel tests, parity, bench
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,driver_version,memory.total --format=csv > gpurun_out/gpu.txt
: > gpurun_out/kernels.log
for t in test_conv_tc_halo_descriptor_probe test_conv_tc test_conv_direct "not conv"; do
  echo "=== $t" >> gpurun_out/kernels.log
  timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -q -s -k "$t" --timeout 300 2>&1 | tail -40 >> gpurun_out/kernels.log
done
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -s --timeout 600 2>&1 | tail -40 > gpurun_out/parity.log
timeout 1200 python bench.py --steps 1 --warmup 1 > gpurun_out/bench1.json 2> gpurun_out/bench1.err
grep -E "===|passed|failed|probe|Error" gpurun_out/kernels.log | head -40
tail -5 gpurun_out/parity.log
tail -c 2500 gpurun_out/bench1.json; tail -3 gpurun_out/bench1.err
