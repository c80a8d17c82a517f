#!/bin/bash
# round 2, call W: the direct option of the linear solve (cuSOLVER sparse QR on the device CSR), alone and inside the reference's time loop
set -u
mkdir -p gpurun_out
O=gpurun_out/r02w
timeout -k 10 600 python -m pytest tests/test_gpu_newton.py tests/test_gpu_adapter.py -m gpu -q -s -k "direct or gmres or config1" > ${O}_tests.log 2>&1; echo "rc=$?" >> ${O}_tests.log; grep -E "passed|failed|FAILED|rc=|\[direct|\[config 1, direct|Error|error" ${O}_tests.log | cut -c1-400 | tail -14
