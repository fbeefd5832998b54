#!/bin/bash
# r1r: narrow 1x1 conv fast path, fir_down at 4 CTAs per SM
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q --timeout 600 2>&1 | tail -15 > gpurun_out/r1r_pytest.log; tail -3 gpurun_out/r1r_pytest.log
DITSEP_BENCH_DUMP=gpurun_out/r1r_shapes.txt timeout 900 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r1r_bench.json 2> gpurun_out/r1r_bench.err; tail -c 700 gpurun_out/r1r_bench.json; tail -2 gpurun_out/r1r_bench.err
grep "conv_direct\|fir_resample:256\|fir_resample:128x128x64" gpurun_out/r1r_shapes.txt
