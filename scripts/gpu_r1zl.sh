#!/bin/bash
# r1zl: programmatic dependent launch extended to gn_finalize, gn_apply, gn_small and the small-image conv kernel
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q --timeout 900 2>&1 | tail -6 > gpurun_out/r1zl_pytest.log; tail -3 gpurun_out/r1zl_pytest.log
for pdl in 0 1 0 1; do
DITSEP_PDL=$pdl timeout 900 python bench.py --steps 2 --warmup 3 --no-cpu-baseline 2>/dev/null > gpurun_out/r1zl_bench_pdl$pdl.json; python -c "import json; d=json.loads(open('gpurun_out/r1zl_bench_pdl$pdl.json').read()); print('PDL=$pdl', d['config']['ms_per_score_step'], d['value'], d['e2e']['value'])"
done
for pdl in 0 1; do
DITSEP_PDL=$pdl timeout 600 python bench.py --batch 1 --steps 2 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('PDL=$pdl B=1:', d['config']['ms_per_score_step'], d['value'])"
done
