#!/bin/bash
# GPU call C: new contact / adapter tests, full suite, default bench (new keys), reference arm.
set -u
mkdir -p gpurun_out
echo "== new contact tests"; timeout -k 10 600 python -m pytest tests/test_gpu_contact.py tests/test_gpu_adapter.py -m gpu -x -q -s -k "spline or nl_contact" > gpurun_out/r02c_contact.log 2>&1; echo "rc=$?" >> gpurun_out/r02c_contact.log; tail -15 gpurun_out/r02c_contact.log
echo "== full gpu suite"; timeout -k 10 900 python -m pytest tests -m gpu -q > gpurun_out/r02c_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r02c_tests.log; tail -8 gpurun_out/r02c_tests.log
echo "== default bench"; timeout -k 10 900 python bench.py > gpurun_out/r02c_bench_default.json 2> gpurun_out/r02c_bench_default.err; echo "rc=$?"; tail -c 1500 gpurun_out/r02c_bench_default.err; head -c 3000 gpurun_out/r02c_bench_default.json
echo "== reference arm"; timeout -k 10 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02c_bench_reference.json 2> gpurun_out/r02c_bench_reference.err; echo "rc=$?"; cat gpurun_out/r02c_bench_reference.json
