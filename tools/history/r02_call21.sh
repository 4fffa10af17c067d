#!/bin/bash
# round 2, 1-GPU call: consumer-warp skew, finer steps (n/64 of a stage): whole-tile loop durations and kernel times
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
O=gpurun_out
PKG=optimizing-dgemm-on-intel-cpus-with-avx512f_b200
for lib in libdgemm_b200.so libdgemm_b200_skew2.so libdgemm_b200_skew4.so libdgemm_b200_skew8.so libdgemm_b200_skew12.so libdgemm_b200_skew16.so libdgemm_b200.so; do
  for tile in big mid small; do
    echo "== $lib tile=$tile 2048^3"
    DGEMM_B200_LIB=$PWD/$PKG/$lib DGEMM_B200_TILE=$tile TL_VARIANT=0 timeout 120 python tools/timeline.py 2048 2048 2048 | grep -E "seg1 loop duration"
  done
  for n in 1024 2048 4096 8192 16384; do
    DGEMM_B200_LIB=$PWD/$PKG/$lib timeout 120 python tools/one_shape.py $n $n $n | python -c "import json,sys; r=json.loads(sys.stdin.read()); print('$lib', r['M'], round(r['ms']*1e3,2), 'us', round(r['tflops'],3), 'TF')"
  done
done > $O/c21_skew.txt 2>&1
cat $O/c21_skew.txt
