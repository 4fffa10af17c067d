#!/bin/bash
# round 2, 1-GPU call: out-of-core tests + timing, pageable staging probe
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
O=gpurun_out
( time timeout 900 python -m pytest tests/test_gpu_out_of_core.py tests/test_gpu_host_api.py -m gpu -x -q ) 2>&1 | tail -n 8
for b in 256 1024; do timeout 300 python tools/e2e_sweep.py 8192 --budget-mb=$b; done > $O/c7_e2e_out_of_core.jsonl 2>&1
timeout 300 python tools/e2e_sweep.py 16384 --budget-mb=2048 >> $O/c7_e2e_out_of_core.jsonl 2>&1
timeout 300 python tools/e2e_sweep.py 8192 16384 >> $O/c7_e2e_out_of_core.jsonl 2>&1
cut -c1-400 $O/c7_e2e_out_of_core.jsonl
timeout 600 optimizing-dgemm-on-intel-cpus-with-avx512f_b200/bin/stage_probe 2 > $O/c7_stage_probe.jsonl 2>&1; echo "stage probe rc=$?"
python - <<'PY'
import json
rows=[json.loads(l) for l in open("gpurun_out/c7_stage_probe.jsonl") if l.startswith("{")]
print(rows[0])
for r in rows[1:]:
    if r.get("what") in ("DMA from pinned memory","cudaMemcpy from pageable memory (driver staging)"): print(r)
def best(what, key):
    sel=[r for r in rows if r.get("what")==what]
    for m in sorted({r["method"] for r in sel}):
        top=sorted([r for r in sel if r["method"]==m], key=lambda r:-r["gb_per_s"])[:4]
        print(what, "|", m, [(r["threads"], r["slot_mb"], r.get("slots_per_thread"), r["gb_per_s"]) for r in top])
best("host copy only",0); best("staged pipeline",0); best("staged way back (d2h)",0)
PY
timeout 300 python tools/e2e_sweep.py 3000 8192 --pageable > $O/c7_e2e_pageable_before.jsonl 2>&1; cut -c1-300 $O/c7_e2e_pageable_before.jsonl
nproc; lscpu | grep -E "Model name|Socket|Thread|L3|NUMA node\(s\)"; free -g | head -2
