#!/bin/bash
# round 2, 1-GPU call: ncu --set full with source-level sampling of the 64 x 32 tile kernel at 512^3 (one data-parallel wave)
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
timeout 60 python tools/one_shape.py 512 512 512 > gpurun_out/c26_tiny_512.json || exit 1
cat gpurun_out/c26_tiny_512.json
timeout 300 ncu --set full --clock-control none --import-source on -k regex:dgemm_tma -s 2 -c 1 \
    -o gpurun_out/c26_tiny_512 -f python tools/one_shape.py 512 512 512 > gpurun_out/c26_ncu.log 2>&1
tail -3 gpurun_out/c26_ncu.log
ncu -i gpurun_out/c26_tiny_512.ncu-rep --page raw --csv > gpurun_out/c26_tiny_512.raw.csv 2>/dev/null
ls -la gpurun_out | tail -5
