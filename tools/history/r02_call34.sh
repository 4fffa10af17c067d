#!/bin/bash
# round 2, 1-GPU call: the consumers' k-tile loop, one ingredient at a time (tools/loop_bubble_probe.cu)
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
timeout 120 optimizing-dgemm-on-intel-cpus-with-avx512f_b200/bin/loop_bubble_probe > gpurun_out/c34_loop_bubble.jsonl 2>&1
cat gpurun_out/c34_loop_bubble.jsonl
