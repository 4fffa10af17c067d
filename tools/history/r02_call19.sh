#!/bin/bash
# round 2, 1-GPU call: where does the main loop of the three tile builds lose time?  The shipped library against the
# measurement build whose producer hands stages over without loading them (make libdbg), timelines at 2048^3 (every
# segment 0 of the mid build, most segments of the others, is one whole tile of 128 k-tiles)
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
O=gpurun_out
PKG=optimizing-dgemm-on-intel-cpus-with-avx512f_b200
for lib in libdgemm_b200.so libdgemm_b200_dbg_notma.so; do
  for tile in big mid small; do
    for n in 1024 2048; do
      echo "== $lib tile=$tile $n^3"
      DGEMM_B200_LIB=$PWD/$PKG/$lib DGEMM_B200_TILE=$tile TL_VARIANT=0 timeout 120 python tools/timeline.py $n $n $n | grep -E "plan|loop duration|kernel span"
    done
  done
done > $O/c19_loop_with_and_without_tma.txt 2>&1
cat $O/c19_loop_with_and_without_tma.txt
