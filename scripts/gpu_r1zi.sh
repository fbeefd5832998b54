#!/bin/bash
# r1zi: one-MUFU SiLU in the bf16 FIR kernels
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q --timeout 900 2>&1 | tail -6 > gpurun_out/r1zi_pytest.log; tail -3 gpurun_out/r1zi_pytest.log
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -s -k "score_step_vs_reference" 2>&1 | grep -E "rel-L2" | head -8
DITSEP_BENCH_DUMP=gpurun_out/r1zi_shapes.txt timeout 900 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r1zi_bench.json 2> gpurun_out/r1zi_bench.err; python -c "import json; d=json.loads(open('gpurun_out/r1zi_bench.json').read()); print(d['config']['ms_per_score_step'], d['value'], d['gpu_launches'], d['roofline']['ms_by_kernel_class'])"; tail -2 gpurun_out/r1zi_bench.err
timeout 600 python bench.py --batch 1 --steps 2 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('B=1:', d['config']['ms_per_score_step'], d['value'])"
grep fir_resample gpurun_out/r1zi_shapes.txt | head -6
