#!/bin/bash
# round 2, 1-GPU call: does unrolling the k-tile loop explain the gap between the probe and the kernel?  Probe with the
# compiler's own choice (u0) and forced 2 (u2); the kernel with `#pragma unroll 1` on its k-tile loop (clock64 stamps)
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
PKG=optimizing-dgemm-on-intel-cpus-with-avx512f_b200
for t in mid small tiny; do for u in u0 u2; do echo "== probe $t $u"; timeout 120 $PKG/bin/lbp2_${t}_$u | grep -E '"mode": (3|7)'; done; done > gpurun_out/c38_unroll.txt 2>&1
for lib in libdgemm_b200_clk.so libdgemm_b200_clk_u1.so; do
  for cfg in "mid 2048 2048 2048" "small 2048 2048 2048" "tiny 512 512 2048"; do
    set -- $cfg
    echo "== $lib tile=$1 $2x$3x$4 (128 k-tiles per tile)"
    DGEMM_B200_LIB=$PWD/$PKG/$lib DGEMM_B200_TILE=$1 TL_VARIANT=0 timeout 120 python tools/timeline.py $2 $3 $4 | grep -E "seg0 loop duration"
  done
done >> gpurun_out/c38_unroll.txt 2>&1
cat gpurun_out/c38_unroll.txt
