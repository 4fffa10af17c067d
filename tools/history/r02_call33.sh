#!/bin/bash
# round 2, 2-GPU call on the final build: multi-GPU tests (torchrun fused / NCCL, mydgemm_gpu_v21, rank pipeline) and bench.py at N = 2
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
O=gpurun_out
( time timeout 900 python -m pytest tests/test_gpu_multi.py tests/test_gpu_rank_pipeline.py -x -q -m gpu ) > $O/c33_multi_tests.log 2>&1; echo "multi tests rc=$?"
tail -n 6 $O/c33_multi_tests.log
( time timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 2 --warmup 3 ) > $O/c33_bench_n2.json 2> $O/c33_bench_n2.err; echo "bench rc=$?"
tail -n 6 $O/c33_bench_n2.err
python - <<'P'
import json
l=[json.loads(x) for x in open("gpurun_out/c33_bench_n2.json") if x.startswith("{")][-1]
print("value", l["value"], "e2e", l["e2e"]["value"], "v21", (l.get("e2e_v21") or {}).get("value"), "parity", l.get("parity_panel_max_ratio"), l["clocks"])
P
