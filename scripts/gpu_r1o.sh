#!/bin/bash
# r1o: ncu launch list of whole score steps + full capture of the 256x256 conv launches; 2-GPU bench through torchrun
mkdir -p gpurun_out
export DITSEP_GRAPH=0
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2600 --csv --log-file gpurun_out/r1o_launches_B32.csv python scripts/profile_step.py 32 2 > gpurun_out/r1o_ncu1.log 2>&1; tail -1 gpurun_out/r1o_ncu1.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:conv_tc_tile_kernel -s 0 -c 4 -o gpurun_out/r1o_full -f python scripts/profile_step.py 32 1 > gpurun_out/r1o_ncu2.log 2>&1; tail -1 gpurun_out/r1o_ncu2.log
unset DITSEP_GRAPH
if [ "$(nvidia-smi -L | wc -l)" -ge 2 ]; then
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r1o_bench_2gpu.json 2> gpurun_out/r1o_bench_2gpu.err; tail -c 700 gpurun_out/r1o_bench_2gpu.json; tail -3 gpurun_out/r1o_bench_2gpu.err
fi
