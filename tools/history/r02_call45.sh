#!/bin/bash
# round 2, 1-GPU call: sweep of the final build (tile retune, 64 x 64 loop fix), cuBLAS beside it on the same box
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
( time timeout 600 python tools/sweep.py ) > gpurun_out/c45_sweep.jsonl 2> gpurun_out/c45_sweep.err
python - <<'P'
import json
for l in open("gpurun_out/c45_sweep.jsonl"):
    r = json.loads(l)
    print(r["M"], r["N"], r["K"], r["lda"], "kernel", r["kernel_tflops"], "TF", r["pct_fp64_peak"], "% cublas", r.get("cublas_tflops"), "e2e", r.get("e2e_tflops"), r["plan"]["ctas"], r["plan"]["stream_k_tiles"])
P
