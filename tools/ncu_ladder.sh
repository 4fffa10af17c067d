#!/bin/bash
# ncu --set full of ONE launch of every rung of the optimisation ladder at N^3 (default 4096):
# gpurun_out/ladder_rungR.{ncu-rep,raw.csv,json}.  Each capture only after the same command exited 0 without ncu.
cd "$(dirname "$0")/.." || exit 1
N=${1:-4096}
mkdir -p gpurun_out
#        rung variant splits
RUNGS=("1 4 1" "2 5 1" "3 3 1" "4 1 1" "5 1 0" "6 6 1")
for r in "${RUNGS[@]}"; do
  set -- $r; rung=$1; variant=$2; splits=$3
  timeout 120 python tools/one_shape.py $N $N $N $variant $splits > gpurun_out/ladder_rung$rung.run.json || { echo "rung $rung failed"; continue; }
  cat gpurun_out/ladder_rung$rung.run.json
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:dgemm_ -s 2 -c 1 \
      -o gpurun_out/ladder_rung$rung -f python tools/one_shape.py $N $N $N $variant $splits > gpurun_out/ladder_rung$rung.ncu.log 2>&1
  ncu -i gpurun_out/ladder_rung$rung.ncu-rep --page raw --csv > gpurun_out/ladder_rung$rung.raw.csv 2>/dev/null
  python tools/ncu_summary.py gpurun_out/ladder_rung$rung.raw.csv gpurun_out/ladder_rung$rung.json | head -n 3
done
