#!/bin/bash
# round 2, 1-GPU call: the consumers' k-tile loop with the kernel's own fragment code, one ingredient at a time (tools/loop_bubble_probe2.cu)
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
for t in big mid small tiny; do timeout 120 optimizing-dgemm-on-intel-cpus-with-avx512f_b200/bin/lbp2_$t; done > gpurun_out/c35_loop_bubble2.jsonl 2>&1
cat gpurun_out/c35_loop_bubble2.jsonl
