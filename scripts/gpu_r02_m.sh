#!/bin/bash
# round 2, call M: the allocation-free entry point alone in a fresh process, the GMRES adapter inside the reference's time
# loop, the whole GPU suite (the no-alloc test runs last), c4 / c5 at full single-GPU size with and without TMEM parking
set -u
mkdir -p gpurun_out
export MB2_BENCH_ALLOW_SHORT=1
O=gpurun_out/r02m
echo "== no-alloc test alone"; timeout -k 10 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -s > ${O}_noalloc_alone.log 2>&1; echo "rc=$?" >> ${O}_noalloc_alone.log; tail -6 ${O}_noalloc_alone.log | cut -c1-300
echo "== GMRES adapter time loop"; timeout -k 10 600 python -m pytest tests/test_gpu_adapter.py -m gpu -q -s -k "config1" > ${O}_gmres.log 2>&1; echo "rc=$?" >> ${O}_gmres.log; grep -E "passed|failed|FAILED|rc=|Newton|linear|reference  |max \|x|assert|Error" ${O}_gmres.log | cut -c1-600 | tail -16
echo "== full suite"; timeout -k 10 1500 python -m pytest tests -m gpu -q > ${O}_tests_full.log 2>&1; echo "rc=$?" >> ${O}_tests_full.log; tail -8 ${O}_tests_full.log | cut -c1-300
for park in 1 0; do
  echo "== c4 n=100 park=$park"; timeout -k 10 600 python bench.py --workload c4 --steps 3 --warmup 3 --no-cpu --no-e2e --no-res-only --park $park > ${O}_c4_n100_park$park.json 2> ${O}_c4_n100_park$park.err; echo "rc=$?"; tail -c 400 ${O}_c4_n100_park$park.err
  python -c "
import json;d=json.loads(open('${O}_c4_n100_park$park.json').read().strip().splitlines()[-1]);r=d['roofline'];print(d['ms_per_step'],r['kernel_ms'],r.get('frac'),r.get('scatter_ms'))"
done
echo "== c5 n=100"; timeout -k 10 900 python bench.py --workload c5 --steps 3 --warmup 3 > ${O}_c5_n100.json 2> ${O}_c5_n100.err; echo "rc=$?"; tail -c 400 ${O}_c5_n100.err; python -c "
import json;d=json.loads(open('${O}_c5_n100.json').read().strip().splitlines()[-1]);print(d['ms_per_step'],d['roofline'].get('frac'),d['extra'],d['e2e'],d['cpu_baseline'])"
du -sh gpurun_out
