#!/bin/bash
# round 2, call J: residual-only kernel v2 with register prefetch of the gather and rotating R warp
set -u
mkdir -p gpurun_out
export MB2_BENCH_ALLOW_SHORT=1
O=gpurun_out/r02j
echo "== tests"; timeout -k 10 1200 python -m pytest tests/test_gpu_parity.py tests/test_gpu_j2.py tests/test_gpu_newton.py tests/test_gpu_partition.py -m gpu -q -x > ${O}_tests.log 2>&1; echo "rc=$?" >> ${O}_tests.log; grep -E "passed|failed|FAILED|rc=|Error" ${O}_tests.log | cut -c1-220 | tail -12
for rv in 0 2 1; do
  echo "== c3 n=48 res_variant=$rv"; timeout -k 10 600 python bench.py --n 48 --steps 10 --warmup 3 --no-cpu --no-e2e --scaling weak --res-variant $rv > ${O}_c3_n48_res$rv.json 2> ${O}_c3_n48_res$rv.err; echo "rc=$?"; tail -c 600 ${O}_c3_n48_res$rv.err
  python -c "
import json;d=json.loads(open('${O}_c3_n48_res$rv.json').read().strip().splitlines()[-1]);print(d['ms_per_step'],d['roofline']['frac'],d['residual_only'])"
done
cap() {  # name, kernel regex, skip, command...
  local name=$1 k=$2 skip=$3; shift 3
  timeout -k 10 900 ncu --set full --clock-control none --import-source on -k regex:$k -s $skip -c 1 -f -o ${O}_$name "$@" > ${O}_ncu_$name.log 2>&1; echo "ncu $name rc=$?"
  ncu -i ${O}_$name.ncu-rep --page details > ${O}_${name}_details.txt 2>&1
  ncu -i ${O}_$name.ncu-rep --page raw --csv > ${O}_${name}_raw.csv 2>&1
  ncu -i ${O}_$name.ncu-rep --page source --csv --print-source cuda,sass > ${O}_${name}_source.csv 2>&1
  rm -f ${O}_$name.ncu-rep
}
cap res2 element_res2_kernel 1 python bench.py --n 32 --steps 2 --warmup 3 --no-cpu --no-e2e --scaling weak
du -sh gpurun_out
