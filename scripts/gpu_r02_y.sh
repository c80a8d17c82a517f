#!/bin/bash
# round 2, call Y: the round-end check once more on the final tree (pytest -x -m gpu, smoke, default bench)
set -u
mkdir -p gpurun_out
O=gpurun_out/r02y
echo "== pytest -x -q -m gpu"; timeout -k 10 1500 python -m pytest tests/ -x -q -m gpu > ${O}_tests.log 2>&1; echo "rc=$?" >> ${O}_tests.log; tail -4 ${O}_tests.log | cut -c1-300
echo "== smoke"; timeout -k 10 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > ${O}_smoke.log 2>&1; echo "rc=$?" >> ${O}_smoke.log; tail -3 ${O}_smoke.log | cut -c1-300
echo "== default bench"; S=$(date +%s); timeout -k 10 900 python bench.py > ${O}_default.json 2> ${O}_default.err; echo "rc=$? wall=$(( $(date +%s) - S ))s"; python -c "
import json;d=json.loads(open('${O}_default.json').read().strip().splitlines()[-1]);r=d['roofline'];print(d['ms_per_step'],d['value'],r['frac'],r['traffic'],d['e2e']['ms_per_step'],d['residual_only']['frac_hbm'],d['fast_mode']['ms_per_step'],d['cpu_baseline']['value'],d['gpu_launches'])"
