#!/bin/bash
# round 2, 1-GPU call: the 64 x 32 tile with 4 consumer warps (one per sub-partition, 32 x 16 warp tiles) against the shipped 8
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
O=gpurun_out
PKG=optimizing-dgemm-on-intel-cpus-with-avx512f_b200
( DGEMM_B200_LIB=$PWD/$PKG/libdgemm_b200_tiny_w4.so timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "tiny" ) > $O/c29_tests_w4.log 2>&1
tail -2 $O/c29_tests_w4.log
for lib in libdgemm_b200.so libdgemm_b200_tiny_w4.so libdgemm_b200.so libdgemm_b200_tiny_w4.so; do
  for shape in "512 512 512" "256 256 256" "448 448 448" "500 501 499"; do
    echo "== $lib $shape"
    DGEMM_B200_LIB=$PWD/$PKG/$lib TL_FLUSH_MB=256 DGEMM_B200_TILE=tiny TL_VARIANT=0 timeout 120 python tools/timeline.py $shape | grep -E "landed|seg0 loop duration|span"
    DGEMM_B200_LIB=$PWD/$PKG/$lib DGEMM_B200_TILE=tiny timeout 120 python tools/one_shape.py $shape | python -c "import json,sys; r=json.loads(sys.stdin.read()); print('   events:', round(r['ms']*1e3,2), 'us', round(r['tflops'],3), 'TF')"
  done
done > $O/c29_tiny_w4.txt 2>&1
cat $O/c29_tiny_w4.txt
