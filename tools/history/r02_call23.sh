#!/bin/bash
# round 2, 1-GPU call: unroll factor of the consumers' k-tile loop (make libunroll): whole-tile loop durations, kernel times
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
O=gpurun_out
PKG=optimizing-dgemm-on-intel-cpus-with-avx512f_b200
for lib in libdgemm_b200.so libdgemm_b200_unroll2.so libdgemm_b200_unroll4.so libdgemm_b200.so; do
  for tile in big mid small; do
    echo "== $lib tile=$tile 2048^3"
    DGEMM_B200_LIB=$PWD/$PKG/$lib DGEMM_B200_TILE=$tile TL_VARIANT=0 timeout 120 python tools/timeline.py 2048 2048 2048 | grep -E "seg1 loop duration"
  done
  for n in 512 1024 2048 4096 8192 16384; do
    DGEMM_B200_LIB=$PWD/$PKG/$lib timeout 120 python tools/one_shape.py $n $n $n | python -c "import json,sys; r=json.loads(sys.stdin.read()); print('$lib', r['M'], round(r['ms']*1e3,2), 'us', round(r['tflops'],3), 'TF')"
  done
done > $O/c23_unroll.txt 2>&1
cat $O/c23_unroll.txt
