#!/bin/bash
# round 2, 1-GPU call: staging chunk size / thread count / planner rate matrix for pageable operands
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
O=gpurun_out
for kb in 256 512 1024 2048 4096; do for t in 6 8; do for g in 30 45; do
  echo "chunk_kb $kb threads $t planner_gbps $g"
  DGEMM_B200_STAGE_CHUNK_KB=$kb DGEMM_B200_STAGE_THREADS=$t DGEMM_B200_PAGEABLE_GBPS=$g timeout 300 python tools/e2e_sweep.py 3000 8192 --pageable 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    try: r = json.loads(l); print('   ', r['n'], r['e2e_ms'], 'ms', r['e2e_tflops'], 'TF')
    except Exception: print(l.strip()[:200])
"
done; done; done | tee $O/c9_pageable_matrix.txt
