#!/bin/bash
# round 2, call V (2 GPUs): the driver's launch line for N = 2 on the final tree (fast_mode key over the NCCL exchange), reference arm
set -u
mkdir -p gpurun_out
N=${1:-2}
O=gpurun_out/r02v
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
echo "== bench --gpus $N"; S=$(date +%s); timeout -k 10 600 $TR --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 > ${O}_c3_${N}gpu.json 2> ${O}_c3_${N}gpu.err; echo "rc=$? wall=$(( $(date +%s) - S ))s"; tail -c 400 ${O}_c3_${N}gpu.err; python -c "
import json;d=json.loads([l for l in open('${O}_c3_${N}gpu.json').read().strip().splitlines() if l.startswith('{')][-1]);print(d['ms_per_step'],d['value'],d['e2e']['ms_per_step'],d['strong_scaling']['ms_per_step'],d['parity_check']['ok'],d['fast_mode']['ms_per_step'],d['fast_mode']['speedup_vs_parity_mode'],d['cpu_baseline']['value'],d['cpu_baseline']['cores'])"
echo "== reference arm under torchrun"; timeout -k 10 600 $TR --master-port 29512 bench.py --impl reference --gpus $N --steps 2 --warmup 1 > ${O}_ref_${N}gpu.json 2> ${O}_ref_${N}gpu.err; echo "rc=$?"; grep '^{' ${O}_ref_${N}gpu.json | cut -c1-300
