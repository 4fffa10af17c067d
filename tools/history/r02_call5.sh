#!/bin/bash
# round 2, 1-GPU call: full -m gpu suite (new: cluster rung, busy-device, patched reference harness), small-tile A/B
# (one vs two CTAs per SM), timelines at 512^3 / 1024^3, cluster rung A/B, then ncu captures (after the plain runs).
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
O=gpurun_out
( time timeout 1500 python -m pytest tests -m gpu -x -q ) > $O/c5_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 6 $O/c5_pytest.log
timeout 600 python tools/small_tile_probe.py > $O/c5_small_tile_probe.jsonl 2> $O/c5_small_tile_probe.err; echo "probe rc=$?"
python - <<'PY'
import json
for l in open("gpurun_out/c5_small_tile_probe.jsonl"):
    r = json.loads(l)
    print(r["M"], r["N"], r["K"], "big", r["big_tflops"], "small", r["small_tflops"], "x2", r["small_x2_tflops"], "auto", r["auto_tflops"], r["auto_plan"])
PY
for n in 512 1024; do for c in 1 2; do echo "== timeline $n^3 small_ctas=$c"; DGEMM_B200_SMALL_CTAS=$c TL_VARIANT=0 timeout 120 python tools/timeline.py $n $n $n; done; done > $O/c5_timelines.txt 2>&1
cat $O/c5_timelines.txt | cut -c1-150
for n in 2048 4096 8192 16384; do for v in 1 6; do timeout 300 python tools/one_shape.py $n $n $n $v 1 0; done; done > $O/c5_cluster_ab.jsonl 2>&1; cat $O/c5_cluster_ab.jsonl | cut -c1-200
timeout 300 python tools/ladder.py 2048 4096 > $O/c5_ladder.jsonl 2>&1; tail -n 12 $O/c5_ladder.jsonl | cut -c1-200
# ---- ncu (each command above exited 0 plain)
timeout 900 ncu --set full --clock-control none --import-source on -k regex:dgemm_tma_cluster_kernel -s 2 -c 1 -o $O/c5_ncu_cluster_16384 -f python tools/one_shape.py 16384 16384 16384 6 1 0 > $O/c5_ncu_cluster.log 2>&1; echo "ncu cluster rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:dgemm_tma_ws_kernel -s 2 -c 1 -o $O/c5_ncu_tma_16384 -f python tools/one_shape.py 16384 16384 16384 1 1 0 > $O/c5_ncu_tma.log 2>&1; echo "ncu tma rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:dgemm_tma_ws_kernel -s 3 -c 1 -o $O/c5_ncu_dg64x2_1024 -f python tools/one_shape.py 1024 1024 1024 0 0 256 > $O/c5_ncu_dg64x2.log 2>&1; echo "ncu dg64x2 rc=$?"
ls -la $O | tail -n 15
