#!/bin/bash
# round 2, 8-GPU call: the e2e experiment (default schedule vs plain ring, timelines, aggregate PCIe probe), then bench.py
cd "$(dirname "$0")/.." || exit 1
N=${1:-8}
mkdir -p gpurun_out
timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 tools/e2e_multi.py --size 65536 --variants ${VARIANTS:-default,ring} --passes 2 --out gpurun_out/c4_e2e_multi_n$N.jsonl --timeline-dir gpurun_out > gpurun_out/c4_e2e_n$N.log 2>&1; echo "e2e tool rc=$?"
grep -v "^\*\|OMP_NUM\|^$" gpurun_out/c4_e2e_n$N.log | cut -c 1-2500 | tail -n 8
BENCH_TIMELINE_FILE=gpurun_out/c4_bench_timeline_n$N.json timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29514 bench.py --gpus $N --steps 2 --warmup 1 > gpurun_out/c4_bench_n$N.json 2> gpurun_out/c4_bench_n$N.err; echo "bench rc=$?"
tail -n 5 gpurun_out/c4_bench_n$N.err
python - <<PY
import json
l=json.loads(open("gpurun_out/c4_bench_n$N.json").read().strip().splitlines()[-1])
print("value", l["value"], "e2e", l["e2e"]["value"], l["e2e"]["ms_per_step"], "v21", l.get("e2e_v21"), "parity", l.get("parity_panel_max_ratio"))
print(json.dumps(l["e2e"]["timeline"]["per_rank"]))
PY
