#!/bin/bash
# round 2, 1-GPU call: parity of the 128 x 64 build, tile probe (big / small / mid), timelines, ncu of every ladder rung
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
O=gpurun_out
( time timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "mid_tile or small_tile_build" ) > $O/c17_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 6 $O/c17_pytest.log
PROBE_NO_X2=1 timeout 900 python tools/small_tile_probe.py 256 256 256 384 384 384 512 512 512 640 640 640 768 768 768 896 896 896 1000 1003 997 1024 1024 1024 1280 1280 1280 1536 1536 1536 1792 1792 1792 2048 2048 2048 2560 2560 2560 3000 3000 3000 4096 4096 4096 512 512 4096 512 512 16384 2048 2048 128 128 4096 512 4096 128 512 > $O/c17_tile_probe.jsonl 2> $O/c17_tile_probe.err; echo "probe rc=$?"
python - <<'PY'
import json
for l in open("gpurun_out/c17_tile_probe.jsonl"):
    r = json.loads(l); print(r["M"], r["N"], r["K"], "big", r["big_tflops"], "small", r["small_tflops"], "mid", r["mid_tflops"], "auto", r["auto_tflops"], "mid/best", r["mid_over_best"])
PY
for cfg in "mid 1024" "mid 2048" "big 2048" "mid 512" "mid 1536"; do
  set -- $cfg
  echo "== timeline $2^3 tile=$1"; DGEMM_B200_TILE=$1 TL_VARIANT=0 timeout 120 python tools/timeline.py $2 $2 $2
done > $O/c17_timelines.txt 2>&1
cat $O/c17_timelines.txt | head -n 80
bash tools/ncu_ladder.sh 4096 > $O/c17_ncu_ladder.log 2>&1; tail -n 30 $O/c17_ncu_ladder.log
rm -f $O/ladder_rung*.raw.csv
ls -la $O | tail -n 20
