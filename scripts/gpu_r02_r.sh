#!/bin/bash
# round 2, call P: fast mode third form: tests, A/B against the T-buffer form at 48^3, ncu, full C3
set -u
mkdir -p gpurun_out
export MB2_BENCH_ALLOW_SHORT=1
O=gpurun_out/r02r
echo "== fast tests"; timeout -k 10 600 python -m pytest tests/test_gpu_fast.py -m gpu -q -s > ${O}_tests.log 2>&1; echo "rc=$?" >> ${O}_tests.log; grep -E "passed|failed|FAILED|rc=|Error|assert" ${O}_tests.log | cut -c1-300 | tail -8
for g in 0; do
  echo "== c3 n=48 analytic fast_gemm=$g"; timeout -k 10 300 python bench.py --n 48 --steps 10 --warmup 3 --no-cpu --no-e2e --no-res-only --scaling weak --tangent analytic --fast-gemm $g > ${O}_fast_n48_g$g.json 2> ${O}_fast_n48_g$g.err; echo "rc=$?"; tail -c 300 ${O}_fast_n48_g$g.err
  python -c "
import json;d=json.loads(open('${O}_fast_n48_g$g.json').read().strip().splitlines()[-1]);r=d['roofline'];print(d['ms_per_step'],r['kernel_ms'],r['frac'],r['achieved'],r['peak'])"
done
echo "== C3 analytic"; timeout -k 10 400 python bench.py --tangent analytic --steps 5 --warmup 3 --no-cpu --no-res-only > ${O}_c3_fast.json 2> ${O}_c3_fast.err; echo "rc=$?"; tail -c 300 ${O}_c3_fast.err; python -c "
import json;d=json.loads(open('${O}_c3_fast.json').read().strip().splitlines()[-1]);r=d['roofline'];print(d['ms_per_step'],d['value'],r['kernel_ms'],r['frac'],r.get('scatter_ms'),d['e2e']['ms_per_step'],d['e2e_host_csr']['ms_per_step'])"
cap() {  # name, kernel regex, skip, command...
  local name=$1 k=$2 skip=$3; shift 3
  timeout -k 10 600 ncu --set full --clock-control none --import-source on -k regex:$k -s $skip -c 1 -f -o ${O}_$name "$@" > ${O}_ncu_$name.log 2>&1; echo "ncu $name rc=$?"
  ncu -i ${O}_$name.ncu-rep --page details > ${O}_${name}_details.txt 2>&1
  ncu -i ${O}_$name.ncu-rep --page raw --csv > ${O}_${name}_raw.csv 2>&1
  ncu -i ${O}_$name.ncu-rep --page source --csv --print-source cuda,sass > ${O}_${name}_source.csv 2>&1
  rm -f ${O}_$name.ncu-rep
}
cap fastv3 element_fast_p2v3_kernel 3 python bench.py --n 16 --steps 2 --warmup 3 --no-cpu --no-e2e --no-res-only --scaling weak --tangent analytic
du -sh gpurun_out
