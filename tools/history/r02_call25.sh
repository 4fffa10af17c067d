#!/bin/bash
# round 2, 1-GPU call: per-CTA timelines of 512^3 on the 64 x 32 tile (one data-parallel wave) against the 64 x 64 tile, L2 flushed and warm
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
O=gpurun_out
for fl in 256 0; do for tile in tiny small; do for shape in "512 512 512" "256 256 256"; do
  echo "== tile=$tile flush_mb=$fl $shape"
  TL_FLUSH_MB=$fl DGEMM_B200_TILE=$tile TL_VARIANT=0 timeout 120 python tools/timeline.py $shape
done; done; done > $O/c25_timelines_tiny.txt 2>&1
cat $O/c25_timelines_tiny.txt
