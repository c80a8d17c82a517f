#!/bin/bash
# round 2, call T: the default bench line on the final tree (with the fast_mode key), launch list of the same command
set -u
mkdir -p gpurun_out
O=gpurun_out/r02t
echo "== default bench"; S=$(date +%s); timeout -k 10 900 python bench.py > ${O}_default.json 2> ${O}_default.err; echo "rc=$? wall=$(( $(date +%s) - S ))s"; tail -c 300 ${O}_default.err; python -c "
import json;d=json.loads(open('${O}_default.json').read().strip().splitlines()[-1]);r=d['roofline'];print(d['ms_per_step'],d['value'],r['frac'],r['traffic'],d['e2e']['ms_per_step'],d['residual_only']['frac_hbm'],d['fast_mode'],d['cpu_baseline']['value'],d['gpu_launches'],d['clocks'])"
echo "== launch list (ncu, small size)"; export MB2_BENCH_ALLOW_SHORT=1
timeout -k 10 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file ${O}_launches_n16.csv python bench.py --n 16 --steps 2 --warmup 3 --no-cpu --no-e2e > ${O}_ncu.log 2>&1; echo "ncu rc=$?"; wc -l ${O}_launches_n16.csv
