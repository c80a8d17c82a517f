#!/bin/bash
set -u
mkdir -p gpurun_out
export MB2_BENCH_ALLOW_SHORT=1
echo "== multipatch tests"; timeout -k 10 900 python -m pytest tests/test_gpu_multipatch.py tests/test_gpu_partition.py tests/test_gpu_adapter.py tests/test_gpu_parity.py -m gpu -q -k "multipatch or beam or patch_by_patch or config4 or emulated" > gpurun_out/r02g_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r02g_tests.log; grep -E "passed|failed|FAILED|rc=|Error" gpurun_out/r02g_tests.log | cut -c1-220 | tail -12
echo "== c4 n=16"; timeout -k 10 300 python bench.py --workload c4 --n 16 --steps 3 --warmup 3 --no-cpu > gpurun_out/r02g_c4_n16.json 2> gpurun_out/r02g_c4_n16.err; echo "rc=$?"; tail -c 800 gpurun_out/r02g_c4_n16.err; head -c 500 gpurun_out/r02g_c4_n16.json; echo
echo "== c4 n=100 (5e5 elements)"; timeout -k 10 900 python bench.py --workload c4 --steps 5 --warmup 3 > gpurun_out/r02g_c4_n100.json 2> gpurun_out/r02g_c4_n100.err; echo "rc=$?"; tail -c 800 gpurun_out/r02g_c4_n100.err; head -c 500 gpurun_out/r02g_c4_n100.json; echo
nvidia-smi --query-gpu=memory.used,memory.total --format=csv
echo "== c5 n=16"; timeout -k 10 600 python bench.py --workload c5 --n 16 --steps 3 --warmup 3 --no-cpu > gpurun_out/r02g_c5_n16.json 2> gpurun_out/r02g_c5_n16.err; echo "rc=$?"; tail -c 800 gpurun_out/r02g_c5_n16.err; python -c "
import json;d=json.loads(open('gpurun_out/r02g_c5_n16.json').read().strip().splitlines()[-1]);print(d['ms_per_step'],d['roofline'],d['extra'])"
echo "== ncu c5 n=16 J2 kernel"
CMD="python bench.py --workload c5 --n 16 --steps 2 --warmup 3 --no-cpu --no-e2e"
timeout -k 10 900 ncu --set full --clock-control none --import-source on -k regex:element_fd_kernel -s 3 -c 1 -f -o gpurun_out/r02g_j2 $CMD > gpurun_out/r02g_ncu_j2.log 2>&1; echo "ncu rc=$?"
ncu -i gpurun_out/r02g_j2.ncu-rep --page details > gpurun_out/r02g_j2_details.txt 2>&1
ncu -i gpurun_out/r02g_j2.ncu-rep --page raw --csv > gpurun_out/r02g_j2_raw.csv 2>&1
ncu -i gpurun_out/r02g_j2.ncu-rep --page source --csv --print-source cuda,sass > gpurun_out/r02g_j2_source.csv 2>&1
rm -f gpurun_out/r02g_j2.ncu-rep
echo "== c5 n=100"; timeout -k 10 900 python bench.py --workload c5 --steps 3 --warmup 3 > gpurun_out/r02g_c5_n100.json 2> gpurun_out/r02g_c5_n100.err; echo "rc=$?"; tail -c 600 gpurun_out/r02g_c5_n100.err; python -c "
import json;d=json.loads(open('gpurun_out/r02g_c5_n100.json').read().strip().splitlines()[-1]);print(d['ms_per_step'],d['roofline'],d['extra'])"
du -sh gpurun_out
