#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -s --timeout 600 2>&1 | grep -E "^score step|^sampler|passed|failed" > gpurun_out/parity2.log
timeout 1500 python bench.py --steps 1 --warmup 1 > gpurun_out/bench1.json 2> gpurun_out/bench1.err
tail -c 3000 gpurun_out/bench1.json
