#!/bin/bash
# round 2, call U: ncu --set full of the headline kernel as it stands at the end of the round (traffic.json must speak for the
# current source), and of csr_gather_kernel
set -u
mkdir -p gpurun_out
export MB2_BENCH_ALLOW_SHORT=1
O=gpurun_out/r02u
cap() {  # name, kernel regex, skip, command...
  local name=$1 k=$2 skip=$3; shift 3
  timeout -k 10 600 ncu --set full --clock-control none --import-source on -k regex:$k -s $skip -c 1 -f -o ${O}_$name "$@" > ${O}_ncu_$name.log 2>&1; echo "ncu $name rc=$?"
  ncu -i ${O}_$name.ncu-rep --page details > ${O}_${name}_details.txt 2>&1
  ncu -i ${O}_$name.ncu-rep --page raw --csv > ${O}_${name}_raw.csv 2>&1
  ncu -i ${O}_$name.ncu-rep --page source --csv --print-source cuda,sass > ${O}_${name}_source.csv 2>&1
  rm -f ${O}_$name.ncu-rep
}
CMD="python bench.py --n 16 --steps 2 --warmup 3 --no-cpu --no-e2e --no-res-only --no-fast --scaling weak"
cap elemfd 'element_fd_kernel' 3 $CMD
cap gather 'csr_gather_kernel' 3 $CMD
du -sh gpurun_out
