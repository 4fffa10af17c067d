#!/bin/bash
# round 2, 1-GPU call: whole-tile loop durations of every tile build in SM CYCLES (make libclk: clock64 stamps), with and
# without TMA, next to the nanosecond stamps of the shipped library: what is the SM clock while a timeline is taken?
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
PKG=optimizing-dgemm-on-intel-cpus-with-avx512f_b200
for lib in libdgemm_b200.so libdgemm_b200_clk.so libdgemm_b200_clk_notma.so; do
  for cfg in "big 2048 2048 2048" "mid 2048 2048 2048" "small 2048 2048 2048" "tiny 512 512 2048"; do
    set -- $cfg
    echo "== $lib tile=$1 $2x$3x$4 (128 k-tiles per tile)"
    DGEMM_B200_LIB=$PWD/$PKG/$lib DGEMM_B200_TILE=$1 TL_VARIANT=0 timeout 120 python tools/timeline.py $2 $3 $4 | grep -E "seg0 loop duration|seg1 loop duration"
  done
done > gpurun_out/c36_loop_clk.txt 2>&1
cat gpurun_out/c36_loop_clk.txt
