#!/bin/bash
# k-lockstep on/off: kernel time (plain run) and DRAM traffic / L2 hit rate (ncu) of one dgemm_tma_ws_kernel launch
#   tools/lockstep_probe.sh SIZE -> gpurun_out/lockstep_probe_SIZE.txt
n=${1:-16384}
out=gpurun_out/lockstep_probe_$n.txt
: > $out
run() {
  echo "== $1" >> $out
  env $1 python tools/one_shape.py $n $n $n >> $out 2>&1
  env $1 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,gpu__time_duration.sum --clock-control none -k regex:dgemm_tma -s 2 -c 1 python tools/one_shape.py $n $n $n 2>&1 | grep -E "dram__|lts__|gpu__time" >> $out
}
run DGEMM_B200_NO_LOCKSTEP=1
for kt in ${SYNC_KTS:-16 32 64 128}; do run DGEMM_B200_SYNC_KT=$kt; done
cat $out
