#!/bin/bash
# r1n: state check (fused GN opt-in), ncu launch list + full capture of the dominant conv kernel
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q --timeout 900 2>&1 | tail -15 > gpurun_out/r1n_pytest.log; tail -3 gpurun_out/r1n_pytest.log
DITSEP_BENCH_DUMP=gpurun_out/r1n_shapes.txt timeout 900 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r1n_bench.json 2> gpurun_out/r1n_bench.err; tail -c 900 gpurun_out/r1n_bench.json; tail -2 gpurun_out/r1n_bench.err
python scripts/bench_hbm.py 32 > gpurun_out/r1n_bench_hbm.log 2>&1; tail -12 gpurun_out/r1n_bench_hbm.log
timeout 600 python scripts/profile_step.py 32 2 > gpurun_out/r1n_plain.log 2>&1; tail -1 gpurun_out/r1n_plain.log
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/r1n_launches_B32.csv python scripts/profile_step.py 32 2 > gpurun_out/r1n_ncu1.log 2>&1; tail -2 gpurun_out/r1n_ncu1.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:conv_tc_tile_kernel -s 40 -c 6 -o gpurun_out/r1n_full -f python scripts/profile_step.py 32 2 > gpurun_out/r1n_ncu2.log 2>&1; tail -2 gpurun_out/r1n_ncu2.log
ls -la gpurun_out/r1n_full.ncu-rep
