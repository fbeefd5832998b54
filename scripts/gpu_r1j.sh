#!/bin/bash
# r1j: MMA warp probes next step's barriers under a full queue; timing experiments DITSEP_CONV_DBG=1 (alternate accumulators), 2 (no tap shift)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k "conv_tc" 2>&1 | tail -3
python scripts/bench_conv.py 32 10 > gpurun_out/r1j_bench_conv.log 2>&1; cat gpurun_out/r1j_bench_conv.log
for d in 1 2 3; do echo "DBG=$d"; DITSEP_CONV_DBG=$d python scripts/bench_conv.py 32 10 256x256 2>&1 | tee gpurun_out/r1j_bench_conv_dbg$d.log; done
timeout 300 python scripts/trace_conv.py 64 64 256 32 0 > gpurun_out/r1j_trace_64_64.txt 2>&1; tail -12 gpurun_out/r1j_trace_64_64.txt
