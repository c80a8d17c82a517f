#!/bin/bash
# round 2, call N (4 GPUs): config 4 with the NCCL exchange launched early / late (A/B), GMRES adapter time loop
set -u
mkdir -p gpurun_out
export MB2_BENCH_ALLOW_SHORT=1
N=${1:-4}
O=gpurun_out/r02n
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
echo "(GMRES test: see call N first attempt)"
for late in 0 1; do
  echo "== c4 n=64 --gpus $N mg_late=$late"; timeout -k 10 300 $TR --master-port 2951$late bench.py --workload c4 --size 64 --gpus $N --steps 5 --warmup 3 --no-cpu --no-e2e --no-res-only --mg-late $late > ${O}_c4_n64_${N}gpu_late$late.json 2> ${O}_c4_n64_${N}gpu_late$late.err; echo "rc=$?"; tail -c 300 ${O}_c4_n64_${N}gpu_late$late.err
  python -c "
import json;d=json.loads([l for l in open('${O}_c4_n64_${N}gpu_late$late.json').read().strip().splitlines() if l.startswith('{')][-1]);r=d['roofline'];print(d['ms_per_step'],r['kernel_ms'],r.get('frac'),r.get('scatter_ms'),d.get('parity_check',{}).get('ok'))"
done
