#!/bin/bash
# diagnosis of conv_ts_kernel: debug toggles (DITSEP_TS_DEBUG) on the dominant shape, then one ncu --set full capture
mkdir -p gpurun_out
for d in 0 1 3 4 5 7 8; do
  echo "== DITSEP_TS_DEBUG=$d"
  DITSEP_TS_DEBUG=$d timeout 300 python scripts/bench_conv.py 32 10 "256x256:64->64:k3" 2>&1 | grep "256x256"
done
timeout 600 ncu --set full --clock-control none --import-source on -k regex:conv_ts -s 3 -c 1 -o gpurun_out/r2q_ts_full -f python scripts/bench_conv.py 32 2 "256x256:64->64:k3" > gpurun_out/r2q_ts_ncu.log 2>&1
tail -3 gpurun_out/r2q_ts_ncu.log
