#!/bin/bash
# r1zo: end-of-session state check exactly as the driver runs it: pytest -m gpu, smoke(), default bench, reference arm
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q --timeout 900 2>&1 | tail -4 > gpurun_out/r1zo_pytest.log; tail -2 gpurun_out/r1zo_pytest.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | grep -v Warn | tail -6 | tee gpurun_out/r1zo_smoke.log
( time timeout 1200 python bench.py > gpurun_out/r1zo_bench.json 2> gpurun_out/r1zo_bench.err ) 2>&1 | grep real; tail -c 1500 gpurun_out/r1zo_bench.json
( time timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r1zo_bench_ref.json 2> gpurun_out/r1zo_bench_ref.err ) 2>&1 | grep real; tail -c 600 gpurun_out/r1zo_bench_ref.json
