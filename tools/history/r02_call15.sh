#!/bin/bash
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
O=gpurun_out
timeout 900 python bench.py --no-scale-base > $O/c15_bench_n1.json 2> $O/c15_bench_n1.err; echo "bench rc=$?"; tail -n 3 $O/c15_bench_n1.err
python - <<'PY'
import json
l=json.loads(open("gpurun_out/c15_bench_n1.json").read().strip().splitlines()[-1])
print("value", l["value"], "e2e", l["e2e"]["value"], l["e2e"]["ms_per_step"], l["e2e"].get("ms_per_rep_this_rank"))
PY
timeout 900 python bench.py --impl reference --steps 1 --warmup 0 > $O/c15_bench_ref.json 2> $O/c15_bench_ref.err; echo "ref rc=$?"; cut -c1-600 $O/c15_bench_ref.json
