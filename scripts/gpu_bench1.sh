#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python bench.py --steps 1 --warmup 1 > gpurun_out/bench1.json 2> gpurun_out/bench1.err
tail -c 4000 gpurun_out/bench1.json; tail -3 gpurun_out/bench1.err
python scripts/profile_step.py 8 2 > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/launches_r1a.csv python scripts/profile_step.py 8 2 > gpurun_out/ncu.log 2>&1
tail -2 gpurun_out/ncu.log
