#!/bin/bash
# round 2, 1-GPU call: raw per-CTA stamps of the mid sizes (stream-K on the 128 x 64 tile), L2 flushed, for offline analysis
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
for shape in "1024 1024 1024" "1000 1003 997" "2048 2048 2048" "1536 1536 1536"; do
  tag=$(echo $shape | tr ' ' 'x')
  TL_DUMP=gpurun_out/c32_stamps_$tag.json TL_FLUSH_MB=256 TL_VARIANT=0 timeout 120 python tools/timeline.py $shape
done > gpurun_out/c32_timelines.txt 2>&1
tail -30 gpurun_out/c32_timelines.txt
