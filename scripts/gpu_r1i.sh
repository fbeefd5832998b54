#!/bin/bash
# r1i: first layer on tensor cores (hi/lo split), conv per-tile timeline (clock64 trace), bench
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q --timeout 900 2>&1 | tail -15 > gpurun_out/r1i_pytest.log; tail -4 gpurun_out/r1i_pytest.log
timeout 300 python -m pytest tests/test_gpu_kernels.py -m gpu -q -s -k hilo 2>&1 | grep -E "rel-L2|passed|failed" > gpurun_out/r1i_hilo.log; cat gpurun_out/r1i_hilo.log
timeout 300 python scripts/trace_conv.py 64 64 256 32 0 > gpurun_out/r1i_trace_64_64.txt 2>&1; tail -4 gpurun_out/r1i_trace_64_64.txt
timeout 300 python scripts/trace_conv.py 128 64 256 32 0 > gpurun_out/r1i_trace_128_64.txt 2>&1; tail -2 gpurun_out/r1i_trace_128_64.txt
timeout 300 python scripts/trace_conv.py 64 64 256 32 1 > gpurun_out/r1i_trace_64_64_res.txt 2>&1; tail -2 gpurun_out/r1i_trace_64_64_res.txt
timeout 900 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r1i_bench.json 2> gpurun_out/r1i_bench.err; tail -c 1200 gpurun_out/r1i_bench.json; tail -2 gpurun_out/r1i_bench.err
