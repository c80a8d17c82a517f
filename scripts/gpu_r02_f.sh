#!/bin/bash
# 2-GPU validation: the driver's torchrun line for bench.py (parity_check, weak + strong, three e2e flavours), reference arm under torchrun,
# the NCCL check script.
set -u
mkdir -p gpurun_out
N=${1:-2}
echo "== bench --gpus $N"
timeout -k 10 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/r02f_bench_${N}gpu.json 2> gpurun_out/r02f_bench_${N}gpu.err; echo "rc=$?"; tail -c 1500 gpurun_out/r02f_bench_${N}gpu.err; head -c 600 gpurun_out/r02f_bench_${N}gpu.json
echo "== reference arm under torchrun"
timeout -k 10 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus $N --steps 2 --warmup 1 > gpurun_out/r02f_ref_${N}gpu.json 2> gpurun_out/r02f_ref_${N}gpu.err; echo "rc=$?"; cat gpurun_out/r02f_ref_${N}gpu.json | cut -c1-400
echo "== mg_nccl_check"
timeout -k 10 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 tests/mg_nccl_check.py > gpurun_out/r02f_mg_nccl_${N}gpu.log 2>&1; echo "rc=$?"; grep -E "OK|MISMATCH" gpurun_out/r02f_mg_nccl_${N}gpu.log | tail -8
