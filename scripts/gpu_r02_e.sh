#!/bin/bash
set -u
mkdir -p gpurun_out
export MB2_BENCH_ALLOW_SHORT=1
echo "== curved + fast tests"; timeout -k 10 900 python -m pytest tests/test_gpu_curved.py tests/test_gpu_fast.py tests/test_gpu_adapter.py -m gpu -q -s -k "curved or fast or dmma or tensor or newton or annulus or cylinder or square_nurbs" > gpurun_out/r02e_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r02e_tests.log; grep -E "passed|failed|FAILED|rc=|bit-equal|Error" gpurun_out/r02e_tests.log | cut -c1-220 | tail -30
for cfg in "analytic 0" "analytic 1"; do
  set -- $cfg
  timeout -k 10 300 python bench.py --n 48 --steps 5 --warmup 3 --no-e2e --no-cpu --no-res-only --tangent $1 --fast-gemm $2 > gpurun_out/r02e_sweep_$1_$2.json 2> gpurun_out/r02e_sweep_$1_$2.err; rc=$?
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/r02e_sweep_$1_$2.json").read().strip().splitlines()[-1]); r=d["roofline"]
    print("$1 gemm=$2 rc=$rc ms/step %.3f kernel_ms %.3f scatter_ms %.3f achieved %.2f peak %.2f frac %.4f" % (d["ms_per_step"], r["kernel_ms"], r["scatter_ms"], r["achieved"], r["peak"], r["frac"]))
except Exception as ex:
    print("$1 $2 rc=$rc: no line", ex); print(open("gpurun_out/r02e_sweep_$1_$2.err").read()[-1500:])
PY
done
echo "== C3 fast mode"; timeout -k 10 600 python bench.py --tangent analytic --no-cpu --no-res-only > gpurun_out/r02e_bench_c3_fast.json 2> gpurun_out/r02e_bench_c3_fast.err; echo "rc=$?"; tail -c 600 gpurun_out/r02e_bench_c3_fast.err
for gm in 0; do
  CMD="python bench.py --n 16 --steps 2 --warmup 3 --no-e2e --no-cpu --no-res-only --tangent analytic --fast-gemm $gm"
  timeout -k 10 600 ncu --set full --clock-control none --import-source on -k regex:element_fast -s 3 -c 1 -f -o gpurun_out/r02e_fast_g$gm $CMD > gpurun_out/r02e_ncu_g$gm.log 2>&1; echo "ncu rc=$?"
  ncu -i gpurun_out/r02e_fast_g$gm.ncu-rep --page details > gpurun_out/r02e_fast_g${gm}_details.txt 2>&1
  ncu -i gpurun_out/r02e_fast_g$gm.ncu-rep --page raw --csv > gpurun_out/r02e_fast_g${gm}_raw.csv 2>&1
  ncu -i gpurun_out/r02e_fast_g$gm.ncu-rep --page source --csv --print-source cuda,sass > gpurun_out/r02e_fast_g${gm}_source.csv 2>&1
  rm -f gpurun_out/r02e_fast_g$gm.ncu-rep
done
