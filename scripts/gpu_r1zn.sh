#!/bin/bash
# r1zn: A/B of the programmatic-dependent-launch levels on one box (0 = off, 1 = conv + GroupNorm kernels, 2 = + FIR, attention, gn_partial)
mkdir -p gpurun_out
for pdl in 0 1 2 0 1 2; do
DITSEP_PDL=$pdl timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline 2>/dev/null > gpurun_out/r1zn_bench_pdl$pdl.json; python -c "import json; d=json.loads(open('gpurun_out/r1zn_bench_pdl$pdl.json').read()); print('PDL=$pdl', d['config']['ms_per_score_step'], d['value'], d['e2e']['value'])"
done
for pdl in 1 2; do
DITSEP_PDL=$pdl timeout 600 python bench.py --batch 1 --steps 2 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('PDL=$pdl B=1:', d['config']['ms_per_score_step'], d['value'])"
done
