#!/bin/bash
# r1zq: ncu launch list of whole score steps + full capture of the 256x256 conv launches for the end-of-round build
mkdir -p gpurun_out
export DITSEP_GRAPH=0
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2600 --csv --log-file gpurun_out/r1zq_launches_B32.csv python scripts/profile_step.py 32 2 > gpurun_out/r1zq_ncu1.log 2>&1; tail -1 gpurun_out/r1zq_ncu1.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:conv_tc_tile_kernel -s 0 -c 4 -o gpurun_out/r1zq_full -f python scripts/profile_step.py 32 1 > gpurun_out/r1zq_ncu2.log 2>&1; tail -1 gpurun_out/r1zq_ncu2.log
