#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -q -x -k "conv_tc or gn_finalize" --timeout 300 2>&1 | tail -5
python scripts/bench_conv.py 32 10 > gpurun_out/bench_conv_v3.log 2>&1; cat gpurun_out/bench_conv_v3.log
