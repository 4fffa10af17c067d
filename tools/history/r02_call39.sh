#!/bin/bash
# round 2, 1-GPU call: unroll factor of the k-tile loop (compiler's choice / 1 / 2 / 4), whole-tile loop durations in SM cycles, every tile build
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
PKG=optimizing-dgemm-on-intel-cpus-with-avx512f_b200
for lib in libdgemm_b200_clk.so libdgemm_b200_clk_u1.so libdgemm_b200_clk_u2.so libdgemm_b200_clk_u4.so; do
  for cfg in "big 2048 2048 2048" "mid 2048 2048 2048" "small 2048 2048 2048" "tiny 512 512 2048"; do
    set -- $cfg
    echo "== $lib tile=$1 $2x$3x$4 (128 k-tiles per tile)"
    DGEMM_B200_LIB=$PWD/$PKG/$lib DGEMM_B200_TILE=$1 TL_VARIANT=0 timeout 120 python tools/timeline.py $2 $3 $4 | grep -E "seg1 loop duration|seg0 loop duration" | tail -1
  done
done > gpurun_out/c39_unroll_clk.txt 2>&1
cat gpurun_out/c39_unroll_clk.txt
