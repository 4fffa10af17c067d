#!/bin/bash
# round 2, 1-GPU call: cluster rung v2 (default-semantics barriers) A/B + ncu, launch-overhead decomposition
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
O=gpurun_out
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "cluster" 2>&1 | tail -n 3
for n in 2048 4096 16384; do for v in 1 6; do timeout 300 python tools/one_shape.py $n $n $n $v 1 0; done; done > $O/c6_cluster_ab.jsonl 2>&1; cut -c1-120 $O/c6_cluster_ab.jsonl
timeout 300 python tools/launch_overhead_probe.py > $O/c6_launch_overhead.jsonl 2> $O/c6_launch_overhead.err; echo "probe rc=$?"; cut -c1-230 $O/c6_launch_overhead.jsonl
timeout 900 ncu --set full --clock-control none --import-source on -k regex:dgemm_tma_cluster_kernel -s 2 -c 1 -o $O/c6_ncu_cluster_16384 -f python tools/one_shape.py 16384 16384 16384 6 1 0 > $O/c6_ncu_cluster.log 2>&1; echo "ncu cluster rc=$?"
