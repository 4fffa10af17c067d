#!/bin/bash
# round 2, 1-GPU call: 64 x 32 tile variants (make libtiny): skew between the two warps of a sub-partition, 96-register
# consumers, 8 ring slots; whole-tile loop in SM cycles (512x512x2048: 128 k-tiles) and the event-timed 512^3 call
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
PKG=optimizing-dgemm-on-intel-cpus-with-avx512f_b200
for v in clk skew8 skew16 skew32 regs96 st8 clk; do
  L=$PWD/$PKG/libdgemm_b200_tiny_$v.so
  echo "== $v: kclk per 128 k-tiles: $(DGEMM_B200_LIB=$L DGEMM_B200_TILE=tiny TL_VARIANT=0 timeout 120 python tools/timeline.py 512 512 2048 | grep -E 'seg0 loop duration')"
  for i in 1 2; do DGEMM_B200_LIB=$L timeout 120 python tools/one_shape.py 512 512 512 | python -c "import json,sys; r=json.loads(sys.stdin.read()); print('   $v 512^3', round(r['ms']*1e3,2), 'us', round(r['tflops'],3), 'TF')"; done
done > gpurun_out/c43_tiny_variants2.txt 2>&1
cat gpurun_out/c43_tiny_variants2.txt
