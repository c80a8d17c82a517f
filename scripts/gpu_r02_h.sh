#!/bin/bash
# round 2, call H: TMEM accumulator parking A/B (J2 p=2 kernel, p=3 split kernel), ncu of both + the residual-only kernel
set -u
mkdir -p gpurun_out
export MB2_BENCH_ALLOW_SHORT=1
O=gpurun_out/r02h
echo "== tests"; timeout -k 10 1200 python -m pytest tests/test_gpu_j2.py tests/test_gpu_multipatch.py tests/test_gpu_parity.py tests/test_gpu_partition.py -m gpu -q > ${O}_tests.log 2>&1; echo "rc=$?" >> ${O}_tests.log; grep -E "passed|failed|FAILED|rc=|Error" ${O}_tests.log | cut -c1-220 | tail -12
for park in 1 0; do
  echo "== c5 n=16 park=$park"; timeout -k 10 600 python bench.py --workload c5 --n 16 --steps 5 --warmup 3 --no-cpu --park $park > ${O}_c5_n16_park$park.json 2> ${O}_c5_n16_park$park.err; echo "rc=$?"; tail -c 600 ${O}_c5_n16_park$park.err
  python -c "
import json;d=json.loads(open('${O}_c5_n16_park$park.json').read().strip().splitlines()[-1]);r=d['roofline'];print(d['ms_per_step'],r['kernel_ms'],r.get('frac'),r.get('achieved'),r.get('peak'),d['e2e'])"
  echo "== c5 n=48 park=$park"; timeout -k 10 600 python bench.py --workload c5 --n 48 --steps 3 --warmup 3 --no-cpu --no-e2e --park $park > ${O}_c5_n48_park$park.json 2> ${O}_c5_n48_park$park.err; echo "rc=$?"; tail -c 600 ${O}_c5_n48_park$park.err
  python -c "
import json;d=json.loads(open('${O}_c5_n48_park$park.json').read().strip().splitlines()[-1]);r=d['roofline'];print(d['ms_per_step'],r['kernel_ms'],r.get('frac'),r.get('achieved'),r.get('peak'))"
  echo "== c4 n=32 park=$park"; timeout -k 10 600 python bench.py --workload c4 --n 32 --steps 5 --warmup 3 --no-cpu --no-e2e --park $park > ${O}_c4_n32_park$park.json 2> ${O}_c4_n32_park$park.err; echo "rc=$?"; tail -c 600 ${O}_c4_n32_park$park.err
  python -c "
import json;d=json.loads(open('${O}_c4_n32_park$park.json').read().strip().splitlines()[-1]);r=d['roofline'];print(d['ms_per_step'],r['kernel_ms'],r.get('frac'))"
done
cap() {  # name, kernel regex, skip, command...
  local name=$1 k=$2 skip=$3; shift 3
  timeout -k 10 900 ncu --set full --clock-control none --import-source on -k regex:$k -s $skip -c 1 -f -o ${O}_$name "$@" > ${O}_ncu_$name.log 2>&1; echo "ncu $name rc=$?"
  ncu -i ${O}_$name.ncu-rep --page details > ${O}_${name}_details.txt 2>&1
  ncu -i ${O}_$name.ncu-rep --page raw --csv > ${O}_${name}_raw.csv 2>&1
  ncu -i ${O}_$name.ncu-rep --page source --csv --print-source cuda,sass > ${O}_${name}_source.csv 2>&1
  rm -f ${O}_$name.ncu-rep
}
cap j2park element_fd_kernel 3 python bench.py --workload c5 --n 16 --steps 2 --warmup 3 --no-cpu --no-e2e
cap split element_fd_split_kernel 3 python bench.py --workload c4 --n 16 --steps 2 --warmup 3 --no-cpu --no-e2e
cap res element_res_kernel 1 python bench.py --n 32 --steps 2 --warmup 3 --no-cpu --no-e2e --scaling weak
du -sh gpurun_out
