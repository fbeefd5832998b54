#!/bin/bash
# r1m: GroupNorm + SiLU of the conv input fused into the conv's shared-memory operand path (4 transform warps)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q --timeout 900 2>&1 | tail -15 > gpurun_out/r1m_pytest.log; tail -4 gpurun_out/r1m_pytest.log
python scripts/bench_conv.py 32 10 256x256 > gpurun_out/r1m_bench_conv.log 2>&1; cat gpurun_out/r1m_bench_conv.log
DITSEP_BENCH_DUMP=gpurun_out/r1m_shapes.txt timeout 900 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r1m_bench.json 2> gpurun_out/r1m_bench.err; tail -c 1300 gpurun_out/r1m_bench.json; tail -2 gpurun_out/r1m_bench.err
head -24 gpurun_out/r1m_shapes.txt
DITSEP_FUSE_GN=0 timeout 900 python bench.py --steps 2 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('unfused GN:', d['config']['ms_per_score_step'])"
