#!/bin/bash
# round 2, 1-GPU call: concurrent-caller test, the sweep of the build with the 64 x 32 tile (cuBLAS beside it, same box)
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
O=gpurun_out
( timeout 300 python -m pytest tests/test_gpu_host_api.py -x -q -m gpu -k "concurrent or tiny" ) > $O/c30_tests.log 2>&1
tail -3 $O/c30_tests.log
( time timeout 600 python tools/sweep.py ) > $O/c30_sweep.jsonl 2> $O/c30_sweep.err
tail -4 $O/c30_sweep.err
python - <<'P'
import json
for l in open("gpurun_out/c30_sweep.jsonl"):
    r = json.loads(l)
    print(r["M"], r["N"], r["K"], r["lda"], "kernel", r["kernel_tflops"], "TF", r["pct_fp64_peak"], "% cublas", r.get("cublas_tflops"), "e2e", r.get("e2e_tflops"), r["plan"]["ctas"], r["plan"]["stream_k_tiles"])
P
