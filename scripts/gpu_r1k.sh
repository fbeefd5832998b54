#!/bin/bash
# r1k: fused res-block shortcut (extra 1x1 tap in TMEM), per-pixel split_hilo, MMA warp barrier prefetch
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q --timeout 900 2>&1 | tail -15 > gpurun_out/r1k_pytest.log; tail -5 gpurun_out/r1k_pytest.log
DITSEP_BENCH_DUMP=gpurun_out/r1k_shapes.txt timeout 900 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r1k_bench.json 2> gpurun_out/r1k_bench.err; tail -c 1500 gpurun_out/r1k_bench.json; tail -2 gpurun_out/r1k_bench.err
head -30 gpurun_out/r1k_shapes.txt
