#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -q -x -k "groupnorm or fir or gn_finalize" --timeout 300 2>&1 | tail -3
python scripts/bench_hbm.py 32 10 > gpurun_out/bench_hbm.log 2>&1; cat gpurun_out/bench_hbm.log
