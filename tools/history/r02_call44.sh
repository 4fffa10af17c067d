#!/bin/bash
# round 2, 1-GPU call: skew between the two warps of a sub-partition on the 64 x 64 build now that its k-tile loop is not unrolled
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
PKG=optimizing-dgemm-on-intel-cpus-with-avx512f_b200
for v in 0 4 8 16 0; do
  L=$PWD/$PKG/libdgemm_b200_smallskew$v.so
  echo "== skew $v/64: kclk per 128 k-tiles: $(DGEMM_B200_LIB=$L DGEMM_B200_TILE=small TL_VARIANT=0 timeout 120 python tools/timeline.py 2048 2048 2048 | grep -E 'seg1 loop duration')"
  for n in 768 1024; do DGEMM_B200_TILE=small DGEMM_B200_LIB=$L timeout 120 python tools/one_shape.py $n $n $n | python -c "import json,sys; r=json.loads(sys.stdin.read()); print('   skew$v', r['M'], round(r['ms']*1e3,2), 'us', round(r['tflops'],3), 'TF')"; done
done > gpurun_out/c44_small_skew.txt 2>&1
cat gpurun_out/c44_small_skew.txt
