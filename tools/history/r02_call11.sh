#!/bin/bash
# round 2, 1-GPU call: pageable sweep after the way-back fix; host tests
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
O=gpurun_out
DGEMM_B200_HOST_TRACE=1 timeout 200 python tools/e2e_sweep.py 3000 --pageable 2>&1 | tail -n 2 | cut -c1-700
timeout 600 python tools/e2e_sweep.py 1024 2048 3000 4096 8192 16384 --pageable > $O/c11_e2e_pageable.jsonl 2>&1; cut -c1-260 $O/c11_e2e_pageable.jsonl
( time timeout 900 python -m pytest tests/test_gpu_out_of_core.py tests/test_gpu_host_api.py -m gpu -x -q ) 2>&1 | tail -n 6
