#!/bin/bash
# round 2, 1-GPU call: warp-uniform barrier probe (vote) and no __syncwarp in front of the elected arrive, in the kernel:
# whole-tile loop durations in SM cycles for every tile build, event-timed calls at 4096^3 .. 16384^3
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
PKG=optimizing-dgemm-on-intel-cpus-with-avx512f_b200
for lib in base vote nosw votenosw base vote; do
  L=$PWD/$PKG/libdgemm_b200_$lib.so
  for cfg in "big 2048 2048 2048" "mid 2048 2048 2048" "small 2048 2048 2048" "tiny 512 512 2048"; do
    set -- $cfg
    echo "== $lib tile=$1 kclk per 128 k-tiles: $(DGEMM_B200_LIB=$L DGEMM_B200_TILE=$1 TL_VARIANT=0 timeout 120 python tools/timeline.py $2 $3 $4 | grep -E 'seg1 loop duration|seg0 loop duration' | tail -1)"
  done
  for n in 1024 2048 4096 8192 16384; do
    DGEMM_B200_LIB=$L timeout 120 python tools/one_shape.py $n $n $n | python -c "import json,sys; r=json.loads(sys.stdin.read()); print('$lib', r['M'], round(r['ms']*1e3,2), 'us', round(r['tflops'],3), 'TF')"
  done
done > gpurun_out/c40_vote.txt 2>&1
cat gpurun_out/c40_vote.txt
