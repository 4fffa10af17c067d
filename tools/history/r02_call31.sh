#!/bin/bash
# round 2, 1-GPU call: is the 64 x 32 tile's loop held back by its producer?  The same loop with a producer that hands every
# ring slot over without loading it (results are garbage; timing is the consumers' alone), 8 and 4 consumer warps
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
O=gpurun_out
PKG=optimizing-dgemm-on-intel-cpus-with-avx512f_b200
for lib in libdgemm_b200.so libdgemm_b200_tiny_notma.so libdgemm_b200_tiny_w4.so libdgemm_b200_tiny_w4notma.so; do
  for shape in "512 512 512" "512 512 2048"; do
    echo "== $lib $shape"
    DGEMM_B200_LIB=$PWD/$PKG/$lib TL_FLUSH_MB=256 DGEMM_B200_TILE=tiny TL_VARIANT=0 timeout 120 python tools/timeline.py $shape | grep -E "landed|seg0 loop duration|span"
  done
done > $O/c31_tiny_notma.txt 2>&1
cat $O/c31_tiny_notma.txt
