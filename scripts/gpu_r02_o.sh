#!/bin/bash
# round 2, call O (8 GPUs): C3 with weak + strong scaling (uneven slabs) and config 4 with the late exchange
set -u
mkdir -p gpurun_out
N=${1:-8}
O=gpurun_out/r02o
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
echo "== C3 --gpus $N"; timeout -k 10 360 $TR --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 > ${O}_c3_${N}gpu.json 2> ${O}_c3_${N}gpu.err; echo "rc=$?"; tail -c 500 ${O}_c3_${N}gpu.err; python -c "
import json;d=json.loads([l for l in open('${O}_c3_${N}gpu.json').read().strip().splitlines() if l.startswith('{')][-1]);print(d['ms_per_step'],d['value'],d['e2e']['ms_per_step'],d['strong_scaling'],d['parity_check']['ok'],d['cpu_baseline']['value'])"
echo "== c4 --gpus $N"; timeout -k 10 300 $TR --master-port 29512 bench.py --workload c4 --gpus $N --steps 5 --warmup 3 --no-cpu > ${O}_c4_${N}gpu.json 2> ${O}_c4_${N}gpu.err; echo "rc=$?"; tail -c 500 ${O}_c4_${N}gpu.err; python -c "
import json;d=json.loads([l for l in open('${O}_c4_${N}gpu.json').read().strip().splitlines() if l.startswith('{')][-1]);r=d['roofline'];print(d['ms_per_step'],d['value'],r['kernel_ms'],r['frac'],d['e2e']['ms_per_step'],d['parity_check']['ok'])"
