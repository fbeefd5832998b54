#!/bin/bash
# round-1 state check: full gpu test suite, smoke, default bench (with CPU baseline), ncu launch list + full capture
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q --timeout 900 2>&1 | tail -15 > gpurun_out/r1e_pytest.log; cat gpurun_out/r1e_pytest.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -8 > gpurun_out/r1e_smoke.log; cat gpurun_out/r1e_smoke.log
timeout 900 python bench.py > gpurun_out/r1e_bench.json 2> gpurun_out/r1e_bench.err; tail -c 3000 gpurun_out/r1e_bench.json; tail -3 gpurun_out/r1e_bench.err
timeout 600 python bench.py --impl reference --steps 1 --warmup 1 > gpurun_out/r1e_bench_ref.json 2> gpurun_out/r1e_bench_ref.err; cat gpurun_out/r1e_bench_ref.json
export DITSEP_GRAPH=0
python scripts/profile_step.py 32 2 > gpurun_out/r1e_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 470 -c 520 --csv --log-file gpurun_out/r1e_launches_B32.csv python scripts/profile_step.py 32 2 > gpurun_out/r1e_ncu1.log 2>&1
python scripts/profile_step.py 32 1 > gpurun_out/r1e_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"conv_tc_tile|gn_apply|gn_finalize|fir_down" -s 2 -c 36 -o gpurun_out/r1e_full python scripts/profile_step.py 32 1 > gpurun_out/r1e_ncu2.log 2>&1
ls -la gpurun_out | tail -20
