#!/bin/bash
set -u
mkdir -p gpurun_out
export MB2_BENCH_ALLOW_SHORT=1
echo "== multipatch tests"; timeout -k 10 900 python -m pytest tests/test_gpu_multipatch.py tests/test_gpu_partition.py tests/test_gpu_adapter.py tests/test_gpu_parity.py -m gpu -q -k "multipatch or beam or patch_by_patch or config4 or emulated" > gpurun_out/r02g_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r02g_tests.log; grep -E "passed|failed|FAILED|rc=|Error" gpurun_out/r02g_tests.log | cut -c1-220 | tail -12
echo "== c4 n=16"; timeout -k 10 300 python bench.py --workload c4 --n 16 --steps 3 --warmup 3 --no-cpu > gpurun_out/r02g_c4_n16.json 2> gpurun_out/r02g_c4_n16.err; echo "rc=$?"; tail -c 800 gpurun_out/r02g_c4_n16.err; head -c 500 gpurun_out/r02g_c4_n16.json; echo
echo "== c4 n=100 (5e5 elements)"; timeout -k 10 900 python bench.py --workload c4 --steps 5 --warmup 3 > gpurun_out/r02g_c4_n100.json 2> gpurun_out/r02g_c4_n100.err; echo "rc=$?"; tail -c 800 gpurun_out/r02g_c4_n100.err; head -c 500 gpurun_out/r02g_c4_n100.json; echo
nvidia-smi --query-gpu=memory.used,memory.total --format=csv
