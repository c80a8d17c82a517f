#!/bin/bash
# round 2, call L: GMRES adapter inside the reference's time loop, c4 / c5 at full single-GPU size with and without TMEM parking
set -u
mkdir -p gpurun_out
export MB2_BENCH_ALLOW_SHORT=1
O=gpurun_out/r02l
echo "== tests"; timeout -k 10 900 python -m pytest tests/test_gpu_adapter.py tests/test_gpu_partition.py -m gpu -q -s -k "config1 or group or emulated" > ${O}_tests.log 2>&1; echo "rc=$?" >> ${O}_tests.log; grep -E "passed|failed|FAILED|rc=|Error|GMRES|linear|reference  |max \|x" ${O}_tests.log | cut -c1-420 | tail -14
for park in 1 0; do
  echo "== c4 n=100 park=$park"; timeout -k 10 600 python bench.py --workload c4 --steps 3 --warmup 3 --no-cpu --no-e2e --no-res-only --park $park > ${O}_c4_n100_park$park.json 2> ${O}_c4_n100_park$park.err; echo "rc=$?"; tail -c 600 ${O}_c4_n100_park$park.err
  python -c "
import json;d=json.loads(open('${O}_c4_n100_park$park.json').read().strip().splitlines()[-1]);r=d['roofline'];print(d['ms_per_step'],r['kernel_ms'],r.get('frac'),r.get('scatter_ms'),d['clocks'])"
done
echo "== c4 n=64 park 1/0"
for park in 1 0; do timeout -k 10 600 python bench.py --workload c4 --n 64 --steps 3 --warmup 3 --no-cpu --no-e2e --no-res-only --park $park > ${O}_c4_n64_park$park.json 2> ${O}_c4_n64_park$park.err; python -c "
import json;d=json.loads(open('${O}_c4_n64_park$park.json').read().strip().splitlines()[-1]);r=d['roofline'];print($park,d['ms_per_step'],r['kernel_ms'],r.get('frac'),r.get('scatter_ms'),r.get('launches_per_step'))"; done
echo "== c5 n=100"; timeout -k 10 900 python bench.py --workload c5 --steps 3 --warmup 3 > ${O}_c5_n100.json 2> ${O}_c5_n100.err; echo "rc=$?"; tail -c 600 ${O}_c5_n100.err; python -c "
import json;d=json.loads(open('${O}_c5_n100.json').read().strip().splitlines()[-1]);print(d['ms_per_step'],d['roofline'],d['extra'],d['e2e'],d['cpu_baseline'])"
du -sh gpurun_out
