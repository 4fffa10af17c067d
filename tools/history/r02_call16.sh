#!/bin/bash
# round 2, N-GPU call: the multi-GPU tests that need N >= 2, then bench.py under torchrun
cd "$(dirname "$0")/.." || exit 1
N=${1:-2}
mkdir -p gpurun_out
O=gpurun_out
if [ "$N" = "2" ]; then ( time timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -x -q ) 2>&1 | tail -n 5; fi
BENCH_TIMELINE_FILE=$O/c16_bench_timeline_n$N.json timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29514 bench.py --gpus $N > $O/c16_bench_n$N.json 2> $O/c16_bench_n$N.err; echo "bench rc=$?"
tail -n 4 $O/c16_bench_n$N.err | cut -c1-300
python - <<PY
import json
l=json.loads(open("gpurun_out/c16_bench_n$N.json").read().strip().splitlines()[-1])
print("value", l["value"], "e2e", l["e2e"]["value"], l["e2e"]["ms_per_step"], "v21", l.get("e2e_v21"), "parity", l.get("parity_panel_max_ratio"))
print(json.dumps(l["e2e"]["timeline"]["per_rank"])[:1500])
PY
