#!/bin/bash
# r1g: diagnose what r1f did not fix (fir_down at 11% of HBM, gn_finalize, conv_direct) + conv_tc after the epilogue change
mkdir -p gpurun_out
export DITSEP_GRAPH=0
python scripts/profile_step.py 32 1 > gpurun_out/r1g_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"fir_down|gn_finalize_srcs|conv_direct|conv_tc_tile" -c 12 -o gpurun_out/r1g_full python scripts/profile_step.py 32 1 > gpurun_out/r1g_ncu.log 2>&1
tail -3 gpurun_out/r1g_plain.log; tail -3 gpurun_out/r1g_ncu.log; ls -la gpurun_out | grep r1g
