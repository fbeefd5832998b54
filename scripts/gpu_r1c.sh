#!/bin/bash
# conv v3 (tile kernel, 16 epilogue warps, fused deterministic statistics): tests, microbench, parity, bench
mkdir -p gpurun_out; : > gpurun_out/kernels.log
for t in "conv_tc or gn_finalize" "not conv_tc and not gn_finalize"; do
  echo "=== $t" >> gpurun_out/kernels.log
  timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -q -x -k "$t" --timeout 300 2>&1 | tail -30 >> gpurun_out/kernels.log
done
grep -E "===|passed|failed|Error|assert" gpurun_out/kernels.log | head -30
python scripts/bench_conv.py 32 10 > gpurun_out/bench_conv_v3.log 2>&1; cat gpurun_out/bench_conv_v3.log
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -s --timeout 600 2>&1 | grep -E "rel-L2|passed|failed|Error|assert" > gpurun_out/parity.log; cat gpurun_out/parity.log
timeout 900 python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/bench1.json 2> gpurun_out/bench1.err; tail -c 1500 gpurun_out/bench1.json; tail -3 gpurun_out/bench1.err
