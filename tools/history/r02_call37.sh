#!/bin/bash
# round 2, 1-GPU call: whole-tile loop durations in SM cycles with beta = 0 (the C prefetch warp exits at once) against beta = 1
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
PKG=optimizing-dgemm-on-intel-cpus-with-avx512f_b200
for lib in libdgemm_b200_clk.so libdgemm_b200_clk_notma.so; do
  for beta in 1.0 0.0; do
  for cfg in "mid 2048 2048 2048" "small 2048 2048 2048" "tiny 512 512 2048"; do
    set -- $cfg
    echo "== $lib beta=$beta tile=$1 $2x$3x$4 (128 k-tiles per tile)"
    TL_BETA=$beta DGEMM_B200_LIB=$PWD/$PKG/$lib DGEMM_B200_TILE=$1 TL_VARIANT=0 timeout 120 python tools/timeline.py $2 $3 $4 | grep -E "seg0 loop duration"
  done; done
done > gpurun_out/c37_loop_clk_beta.txt 2>&1
cat gpurun_out/c37_loop_clk_beta.txt
