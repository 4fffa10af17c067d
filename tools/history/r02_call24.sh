#!/bin/bash
# round 2, 1-GPU call: the 64 x 32 tile build (dg64x32) and two k-tiles per ring slot (make libkps): parity of the new build,
# parity of the kps2 measurement library on every tile, tile probe of the three libraries
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
O=gpurun_out
PKG=optimizing-dgemm-on-intel-cpus-with-avx512f_b200
( time timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "tiny or mid_tile or small_tile or stream_k or split" ) > $O/c24_tests_default.log 2>&1
tail -3 $O/c24_tests_default.log
( time DGEMM_B200_LIB=$PWD/$PKG/libdgemm_b200_kps2.so timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "tiny_tile or mid_tile or small_tile or stream_k or split or lockstep or alpha_beta" ) > $O/c24_tests_kps2.log 2>&1
tail -3 $O/c24_tests_kps2.log
for lib in libdgemm_b200.so libdgemm_b200_kps2.so libdgemm_b200_tinykps1.so; do
  DGEMM_B200_LIB=$PWD/$PKG/$lib timeout 300 python tools/tile_probe2.py small
done > $O/c24_probe_small.jsonl 2> $O/c24_probe_small.err
for lib in libdgemm_b200.so libdgemm_b200_kps2.so libdgemm_b200.so libdgemm_b200_kps2.so; do
  DGEMM_B200_LIB=$PWD/$PKG/$lib timeout 300 python tools/tile_probe2.py mid
  DGEMM_B200_LIB=$PWD/$PKG/$lib timeout 300 python tools/tile_probe2.py large
done > $O/c24_probe_mid_large.jsonl 2> $O/c24_probe_mid_large.err

python - <<'P'
import json
for f in ("gpurun_out/c24_probe_small.jsonl", "gpurun_out/c24_probe_mid_large.jsonl"):
    for l in open(f):
        r = json.loads(l)
        print(r["lib"][13:-3] or "default", r["M"], r["N"], r["K"], " ".join(f"{k[:-3]}={r[k]}" for k in r if k.endswith("_us")))
P
