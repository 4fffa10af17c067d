#!/bin/bash
# DRAM traffic and duration of one dgemm_tma_ws_kernel launch for several rasterisation band heights
#   tools/traffic_probe.sh SIZE  ->  gpurun_out/traffic_probe_SIZE.txt
n=${1:-8192}
out=gpurun_out/traffic_probe_$n.txt
: > $out
for g in ${GROUPS_M:-4 8 12 16 24 32 64}; do
  echo "== group_m=$g" >> $out
  DGEMM_B200_GROUP_M=$g python tools/one_shape.py $n $n $n >> $out 2>&1
  DGEMM_B200_GROUP_M=$g ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,gpu__time_duration.sum --clock-control none -k regex:dgemm_tma -s 2 -c 1 python tools/one_shape.py $n $n $n 2>&1 | grep -E "dram__|lts__|gpu__time" >> $out
done
cat $out
