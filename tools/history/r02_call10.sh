#!/bin/bash
# round 2, 1-GPU call: launch floor by launch shape; host-side trace of one pageable / one pinned 3000^3 call
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
O=gpurun_out
timeout 300 python tools/launch_overhead_probe.py > $O/c10_launch_overhead.jsonl 2> $O/c10_launch_overhead.err; echo "probe rc=$?"; tail -n 9 $O/c10_launch_overhead.jsonl | cut -c1-200
DGEMM_B200_HOST_TRACE=1 timeout 200 python tools/e2e_sweep.py 3000 --pageable 2>&1 | tail -n 3 | cut -c1-900
DGEMM_B200_HOST_TRACE=1 timeout 200 python tools/e2e_sweep.py 3000 2>&1 | tail -n 3 | cut -c1-900
