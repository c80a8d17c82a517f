#!/bin/bash
# round 2, call X: residual-only kernel for p = 3 (element_res3_kernel): parity, A/B against the generic kernel at 32^3 / 2 and 64^3 / 2
set -u
mkdir -p gpurun_out
export MB2_BENCH_ALLOW_SHORT=1
O=gpurun_out/r02x
timeout -k 10 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_j2.py tests/test_gpu_multipatch.py tests/test_gpu_partition.py -m gpu -q -x > ${O}_tests.log 2>&1; echo "rc=$?" >> ${O}_tests.log; tail -4 ${O}_tests.log | cut -c1-300
for rv in 0 1; do
  echo "== c4 n=64 res_variant=$rv"; timeout -k 10 600 python bench.py --workload c4 --size 64 --steps 3 --warmup 3 --no-cpu --no-e2e --res-variant $rv > ${O}_c4_n64_res$rv.json 2> ${O}_c4_n64_res$rv.err; echo "rc=$?"; tail -c 300 ${O}_c4_n64_res$rv.err
  python -c "
import json;d=json.loads(open('${O}_c4_n64_res$rv.json').read().strip().splitlines()[-1]);print(d['ms_per_step'],d['residual_only'])"
done
cap() {  # name, kernel regex, skip, command...
  local name=$1 k=$2 skip=$3; shift 3
  timeout -k 10 600 ncu --set full --clock-control none --import-source on -k regex:$k -s $skip -c 1 -f -o ${O}_$name "$@" > ${O}_ncu_$name.log 2>&1; echo "ncu $name rc=$?"
  ncu -i ${O}_$name.ncu-rep --page details > ${O}_${name}_details.txt 2>&1
  ncu -i ${O}_$name.ncu-rep --page raw --csv > ${O}_${name}_raw.csv 2>&1
  ncu -i ${O}_$name.ncu-rep --page source --csv --print-source cuda,sass > ${O}_${name}_source.csv 2>&1
  rm -f ${O}_$name.ncu-rep
}
cap res3 element_res3_kernel 1 python bench.py --workload c4 --size 32 --steps 2 --warmup 3 --no-cpu --no-e2e
du -sh gpurun_out
