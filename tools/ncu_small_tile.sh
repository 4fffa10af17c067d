#!/bin/bash
# ncu --set full of ONE launch of the 64x64-tile kernel (dg64) at 1024^3 and of the small-tile split kernel at
# 1000x1003x997 (LD = dim, LDB odd); each only after the same command exited 0 without ncu.
mkdir -p gpurun_out
for shape in "1024 1024 1024" "1000 1003 997"; do
  tag=$(echo $shape | tr ' ' 'x')
  timeout 60 python tools/one_shape.py $shape > gpurun_out/small_$tag.json || exit 1
  cat gpurun_out/small_$tag.json
  timeout 200 ncu --set full --clock-control none --import-source on -k regex:dgemm_tma -s 2 -c 1 \
      -o gpurun_out/small_$tag -f python tools/one_shape.py $shape > gpurun_out/small_ncu_$tag.log 2>&1
  ncu -i gpurun_out/small_$tag.ncu-rep --page raw --csv > gpurun_out/small_$tag.raw.csv 2>/dev/null
done
ls -la gpurun_out | tail -8
