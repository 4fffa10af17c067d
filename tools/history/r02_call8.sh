#!/bin/bash
# round 2, 1-GPU call: streamed staging of pageable operands (tests + sweep), out-of-core tests again
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
O=gpurun_out
( time timeout 900 python -m pytest tests/test_gpu_out_of_core.py tests/test_gpu_host_api.py -m gpu -x -q ) 2>&1 | tail -n 25
timeout 600 python tools/e2e_sweep.py 1024 2048 3000 4096 8192 16384 --pageable > $O/c8_e2e_pageable.jsonl 2>&1; cut -c1-260 $O/c8_e2e_pageable.jsonl
for t in 4 6 12; do echo "threads $t"; DGEMM_B200_STAGE_THREADS=$t timeout 300 python tools/e2e_sweep.py 3000 8192 --pageable 2>&1 | cut -c1-200; done | tee $O/c8_e2e_pageable_threads.txt
timeout 300 python tools/e2e_sweep.py 3000 8192 16384 2>&1 | cut -c1-260 | tee $O/c8_e2e_pinned.jsonl
for b in 256 1024; do timeout 300 python tools/e2e_sweep.py 8192 --budget-mb=$b; done > $O/c8_e2e_out_of_core.jsonl 2>&1
timeout 300 python tools/e2e_sweep.py 16384 --budget-mb=2048 >> $O/c8_e2e_out_of_core.jsonl 2>&1
timeout 300 python tools/e2e_sweep.py 8192 --budget-mb=256 --pageable >> $O/c8_e2e_out_of_core.jsonl 2>&1
cut -c1-330 $O/c8_e2e_out_of_core.jsonl
