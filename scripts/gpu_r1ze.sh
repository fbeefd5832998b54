#!/bin/bash
# r1ze: temb_dense with the embedding staged in shared memory
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q --timeout 900 2>&1 | tail -6 > gpurun_out/r1ze_pytest.log; tail -3 gpurun_out/r1ze_pytest.log
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -s -k "score_step_vs_reference" 2>&1 | grep -E "rel-L2" | head -8
DITSEP_BENCH_DUMP=gpurun_out/r1ze_shapes.txt timeout 900 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r1ze_bench.json 2> gpurun_out/r1ze_bench.err; python -c "import json; d=json.loads(open('gpurun_out/r1ze_bench.json').read()); print(d['config']['ms_per_score_step'], d['value'], d['gpu_launches'], d['roofline']['ms_by_kernel_class'])"; tail -2 gpurun_out/r1ze_bench.err
timeout 600 python bench.py --batch 1 --steps 2 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('B=1:', d['config']['ms_per_score_step'], d['value'])"
