#!/bin/bash
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
DGEMM_B200_HOST_TRACE=1 timeout 200 python tools/e2e_sweep.py 3000 --pageable > gpurun_out/c13_trace.txt 2>&1
python - <<'PY'
import json
lines=[l for l in open("gpurun_out/c13_trace.txt") if l.startswith("{")]
# last call's trace: find last host_trace line index
idx=[i for i,l in enumerate(lines) if "host_trace_ms" in l]
last=idx[-1]; prev=idx[-2]
print(lines[last][:600])
for l in lines[last+1:]:
    if "staging_block" in l:
        r=json.loads(l); print(r["staging_block"], r["mb"], "sub", r["submitted"], "ready", r["ready"], "first", r["first_chunk_taken"], "retired", r["retired"])
print(lines[-1][:200])
PY
