#!/bin/bash
# r1l: software-pipelined MMA unit loop (tail | commits | next head in one elect region, barrier probes under a queued head)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q --timeout 900 2>&1 | tail -15 > gpurun_out/r1l_pytest.log; tail -4 gpurun_out/r1l_pytest.log
for c in 0 64 128; do timeout 120 python scripts/trace_conv.py 64 64 256 32 0 $c > gpurun_out/r1l_trace_sc$c.txt 2>&1; head -1 gpurun_out/r1l_trace_sc$c.txt; tail -1 gpurun_out/r1l_trace_sc$c.txt; done
timeout 120 python scripts/trace_conv.py 128 64 256 32 0 128 > gpurun_out/r1l_trace_128_sc128.txt 2>&1; head -1 gpurun_out/r1l_trace_128_sc128.txt; tail -1 gpurun_out/r1l_trace_128_sc128.txt
python scripts/bench_conv.py 32 10 > gpurun_out/r1l_bench_conv.log 2>&1; cat gpurun_out/r1l_bench_conv.log
DITSEP_BENCH_DUMP=gpurun_out/r1l_shapes.txt timeout 900 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r1l_bench.json 2> gpurun_out/r1l_bench.err; tail -c 1300 gpurun_out/r1l_bench.json; tail -2 gpurun_out/r1l_bench.err
head -16 gpurun_out/r1l_shapes.txt
