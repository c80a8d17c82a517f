#!/bin/bash
# round 2, call K (8 GPUs): the driver's torchrun line for C3 (weak + strong + e2e + parity_check), config 4 at 8 x 5e5 elements,
# config-5-style at 8 x 1e6, the NCCL check script, the one-process multi-GPU group on real devices.
set -u
mkdir -p gpurun_out
N=${1:-8}
O=gpurun_out/r02k
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
echo "== C3 --gpus $N"; timeout -k 10 300 $TR --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 > ${O}_c3_${N}gpu.json 2> ${O}_c3_${N}gpu.err; echo "rc=$?"; tail -c 800 ${O}_c3_${N}gpu.err; head -c 400 ${O}_c3_${N}gpu.json; echo
echo "== c4 --gpus $N"; timeout -k 10 240 $TR --master-port 29512 bench.py --workload c4 --gpus $N --steps 5 --warmup 3 --no-cpu > ${O}_c4_${N}gpu.json 2> ${O}_c4_${N}gpu.err; echo "rc=$?"; tail -c 800 ${O}_c4_${N}gpu.err; head -c 400 ${O}_c4_${N}gpu.json; echo
echo "== c5 --gpus $N"; timeout -k 10 240 $TR --master-port 29513 bench.py --workload c5 --gpus $N --steps 3 --warmup 3 --no-cpu > ${O}_c5_${N}gpu.json 2> ${O}_c5_${N}gpu.err; echo "rc=$?"; tail -c 800 ${O}_c5_${N}gpu.err; head -c 400 ${O}_c5_${N}gpu.json; echo
echo "== mg_nccl_check"; timeout -k 10 120 $TR --master-port 29514 tests/mg_nccl_check.py > ${O}_mg_nccl_${N}gpu.log 2>&1; echo "rc=$?"; grep -E "OK|MISMATCH|Error" ${O}_mg_nccl_${N}gpu.log | tail -8
echo "== one-process group on $N devices"; timeout -k 10 200 python -m pytest tests/test_gpu_adapter.py tests/test_gpu_partition.py -m gpu -q -k "group or multi_gpu" > ${O}_group_tests.log 2>&1; echo "rc=$?"; tail -3 ${O}_group_tests.log
