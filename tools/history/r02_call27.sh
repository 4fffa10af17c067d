#!/bin/bash
# round 2, 1-GPU call: variants of the 64 x 32 tile build (make libtiny): dual accumulators, early descriptor prefetch, k-tiles per slot
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
O=gpurun_out
PKG=optimizing-dgemm-on-intel-cpus-with-avx512f_b200
( timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "tiny" ) > $O/c27_tests.log 2>&1
tail -2 $O/c27_tests.log
for lib in libdgemm_b200.so libdgemm_b200_tiny_plain.so libdgemm_b200_tiny_nodual.so libdgemm_b200_tiny_latedesc.so libdgemm_b200_tiny_kps2.so libdgemm_b200.so; do
  for shape in "512 512 512" "256 256 256" "448 448 448"; do
    echo "== $lib $shape"
    DGEMM_B200_LIB=$PWD/$PKG/$lib TL_FLUSH_MB=256 DGEMM_B200_TILE=tiny TL_VARIANT=0 timeout 120 python tools/timeline.py $shape | grep -E "landed|seg0 loop duration|span"
    DGEMM_B200_LIB=$PWD/$PKG/$lib DGEMM_B200_TILE=tiny timeout 120 python tools/one_shape.py $shape | python -c "import json,sys; r=json.loads(sys.stdin.read()); print('   events:', round(r['ms']*1e3,2), 'us', round(r['tflops'],3), 'TF')"
  done
done > $O/c27_tiny_variants.txt 2>&1
cat $O/c27_tiny_variants.txt
