#!/bin/bash
# ncu --set full of the tile kernel on the dominant shape (B=8 256x256 64->64 3x3, bias_b + statistics)
mkdir -p gpurun_out
python scripts/bench_conv.py 8 2 "256x256:64->64:k3:b" > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:conv_tc_tile -s 3 -c 1 -o gpurun_out/prof_tile python scripts/bench_conv.py 8 2 "256x256:64->64:k3:b" > gpurun_out/ncu_tile.log 2>&1
tail -3 gpurun_out/ncu_tile.log; cat gpurun_out/plain.log
