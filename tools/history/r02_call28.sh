#!/bin/bash
# round 2, 1-GPU call: the 64 x 32 tile with 16 consumer warps (four per sub-partition, 16 x 8 warp tiles) against the shipped 8
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
O=gpurun_out
PKG=optimizing-dgemm-on-intel-cpus-with-avx512f_b200
( DGEMM_B200_LIB=$PWD/$PKG/libdgemm_b200_tiny_w16.so timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "tiny" ) > $O/c28_tests_w16.log 2>&1
tail -2 $O/c28_tests_w16.log
for lib in libdgemm_b200.so libdgemm_b200_tiny_w16.so libdgemm_b200.so libdgemm_b200_tiny_w16.so; do
  for shape in "512 512 512" "256 256 256" "448 448 448" "500 501 499"; do
    echo "== $lib $shape"
    DGEMM_B200_LIB=$PWD/$PKG/$lib TL_FLUSH_MB=256 DGEMM_B200_TILE=tiny TL_VARIANT=0 timeout 120 python tools/timeline.py $shape | grep -E "landed|seg0 loop duration|span"
    DGEMM_B200_LIB=$PWD/$PKG/$lib DGEMM_B200_TILE=tiny timeout 120 python tools/one_shape.py $shape | python -c "import json,sys; r=json.loads(sys.stdin.read()); print('   events:', round(r['ms']*1e3,2), 'us', round(r['tflops'],3), 'TF')"
  done
done > $O/c28_tiny_w16.txt 2>&1
cat $O/c28_tiny_w16.txt
