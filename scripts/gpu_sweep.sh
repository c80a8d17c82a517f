#!/bin/bash
# usage: gpu_sweep.sh <tag> "<variants>" "<ncu variants>" [n]
# Element-kernel A/B on one B200: parity of every variant, timing sweep, ncu --set full of the listed variants
# (reports are exported to text on the box; the .ncu-rep files are deleted: gpurun_out/ is capped at 64 MiB).
set -u
TAG=$1; VARS=$2; NCUV=$3; N=${4:-48}
mkdir -p gpurun_out
export MB2_BENCH_ALLOW_SHORT=1
echo "== variant parity"; timeout -k 10 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "variant or mass_resident" > gpurun_out/${TAG}_variants.log 2>&1; echo "rc=$?" >> gpurun_out/${TAG}_variants.log; tail -3 gpurun_out/${TAG}_variants.log
for v in $VARS; do
  timeout -k 10 200 python bench.py --n $N --steps 5 --warmup 3 --no-e2e --no-cpu --no-res-only --variant $v > gpurun_out/${TAG}_sweep_v$v.json 2> gpurun_out/${TAG}_sweep_v$v.err; rc=$?
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/${TAG}_sweep_v$v.json").read().strip().splitlines()[-1])
    print("variant $v rc=$rc ms/step %.3f kernel_ms %.3f frac %.4f sm_mhz %s" % (d["ms_per_step"], d["roofline"]["kernel_ms"], d["roofline"]["frac"], d["clocks"]["sm_mhz"]))
except Exception as ex:
    print("variant $v rc=$rc: no line", ex)
PY
done
for v in $NCUV; do
  echo "== ncu variant $v"
  CMD="python bench.py --n 16 --steps 2 --warmup 3 --no-e2e --no-cpu --no-res-only --variant $v"
  timeout -k 10 600 ncu --set full --clock-control none --import-source on -k regex:element_fd -s 3 -c 1 -f -o gpurun_out/${TAG}_elem_v$v $CMD > gpurun_out/${TAG}_ncu_v$v.log 2>&1; echo "rc=$?"
  ncu -i gpurun_out/${TAG}_elem_v$v.ncu-rep --page details > gpurun_out/${TAG}_elem_v${v}_details.txt 2>&1
  ncu -i gpurun_out/${TAG}_elem_v$v.ncu-rep --page raw --csv > gpurun_out/${TAG}_elem_v${v}_raw.csv 2>&1
  ncu -i gpurun_out/${TAG}_elem_v$v.ncu-rep --page source --csv --print-source cuda,sass > gpurun_out/${TAG}_elem_v${v}_source.csv 2>&1
  rm -f gpurun_out/${TAG}_elem_v$v.ncu-rep
done
du -sh gpurun_out
