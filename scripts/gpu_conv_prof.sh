#!/bin/bash
mkdir -p gpurun_out
for m in tap halo; do
  echo "== $m" >> gpurun_out/bench_conv.log
  DITSEP_CONV_MODE=$m python scripts/bench_conv.py 32 10 >> gpurun_out/bench_conv.log 2>&1
done
cat gpurun_out/bench_conv.log
DITSEP_CONV_MODE=halo python scripts/bench_conv.py 8 2 "256x256:64->64:k3:b" > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:conv_tc_halo -s 3 -c 1 -o gpurun_out/prof_halo python scripts/bench_conv.py 8 2 "256x256:64->64:k3:b" > gpurun_out/ncu_halo.log 2>&1
tail -3 gpurun_out/ncu_halo.log
