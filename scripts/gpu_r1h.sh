#!/bin/bash
# r1h: slim conv epilogue, strip FIR kernels, parallel gn_finalize, fast SiLU -- tests, microbenchmarks, bench
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q --timeout 900 2>&1 | tail -15 > gpurun_out/r1h_pytest.log; tail -5 gpurun_out/r1h_pytest.log
python scripts/bench_conv.py 32 10 > gpurun_out/r1h_bench_conv.log 2>&1; cat gpurun_out/r1h_bench_conv.log
python scripts/bench_hbm.py 32 > gpurun_out/r1h_bench_hbm.log 2>&1; cat gpurun_out/r1h_bench_hbm.log
timeout 900 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r1h_bench.json 2> gpurun_out/r1h_bench.err; tail -c 2500 gpurun_out/r1h_bench.json; tail -3 gpurun_out/r1h_bench.err
