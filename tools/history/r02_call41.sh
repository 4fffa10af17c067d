#!/bin/bash
# round 2, 1-GPU call: tile probe after the 64 x 64 build's loop change (where do small / mid / big cross now?)
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
python - > gpurun_out/c41_probe.jsonl 2> gpurun_out/c41_probe.err <<'P'
import json, os, sys
sys.path.insert(0, os.getcwd())
import dgemm_b200 as dg
SHAPES = [(n, n, n) for n in (576, 640, 704, 768, 832, 896, 960, 1024, 1152, 1280, 1408, 1536)] + [(1000, 1003, 997), (768, 768, 2048), (1024, 1024, 512), (512, 512, 2048), (512, 512, 4096)]
ENVS = {"big": {"DGEMM_B200_TILE": "big"}, "small": {"DGEMM_B200_TILE": "small"}, "mid": {"DGEMM_B200_TILE": "mid"}, "auto": {}}
for rnd in range(2):
  for (M, N, K) in SHAPES:
    dA, dB, dC = dg.malloc(8 * M * K), dg.malloc(8 * K * N), dg.malloc(8 * M * N)
    dg.fill_signed(dA, M, K, M, 1); dg.fill_signed(dB, K, N, K, 2); dg.fill_signed(dC, M, N, M, 3)
    dg.lib.dgemm_b200_sync(None)
    row = {"M": M, "N": N, "K": K}
    for name in ENVS:
        os.environ.pop("DGEMM_B200_TILE", None)
        os.environ.update(ENVS[name]); dg.reload_env()
        ms = dg.time_device(M, N, K, 1.0, dA, M, dB, K, 1.0, dC, M, warmup=3, reps=20, flush_l2_bytes=256 << 20)
        row[name + "_us"] = round(ms * 1e3, 2)
    print(json.dumps(row), flush=True)
    for d in (dA, dB, dC): dg.free(d)
P
cat gpurun_out/c41_probe.jsonl; tail -3 gpurun_out/c41_probe.err
