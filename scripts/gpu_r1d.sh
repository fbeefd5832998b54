#!/bin/bash
# CUDA-graphed score step: parity + bench (B=32) + batch-1 latency
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -s --timeout 600 2>&1 | grep -E "rel-L2|passed|failed|Error|assert" > gpurun_out/parity.log; cat gpurun_out/parity.log
timeout 900 python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/bench1.json 2> gpurun_out/bench1.err; tail -c 1400 gpurun_out/bench1.json; tail -3 gpurun_out/bench1.err
timeout 900 python bench.py --steps 2 --warmup 1 --no-cpu-baseline --batch 1 > gpurun_out/bench_b1.json 2> gpurun_out/bench_b1.err; tail -c 1400 gpurun_out/bench_b1.json; tail -3 gpurun_out/bench_b1.err
