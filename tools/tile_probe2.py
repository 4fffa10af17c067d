"""Every CTA-tile build of the TMA kernel on a list of shapes, kernel-only (CUDA events, L2 flushed between reps), one JSON
line per shape; the library under test is the one DGEMM_B200_LIB names (measurement builds: make libkps):
    python tools/tile_probe2.py [small|mid|large] """
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dgemm_b200 as dg
which = sys.argv[1] if len(sys.argv) > 1 else "small"
SHAPES = {
    "small": [(n, n, n) for n in (128, 192, 256, 320, 384, 448, 512, 576, 640, 768)] +
             [(512, 512, 128), (512, 512, 2048), (512, 512, 4096), (512, 384, 512), (500, 501, 499), (256, 1024, 512), (1024, 256, 512)],
    "mid": [(n, n, n) for n in (896, 1024, 1280, 1536, 2048, 3000)] + [(1000, 1003, 997), (512, 512, 16384)],
    "large": [(n, n, n) for n in (4096, 8192, 16384)],
}[which]
TILES = {"small": ("big", "small", "mid", "tiny", "tiny_sk", "auto"), "mid": ("big", "small", "mid", "auto"),
         "large": ("big", "auto")}[which]
ENVS = {"big": {"DGEMM_B200_TILE": "big"}, "small": {"DGEMM_B200_TILE": "small"}, "mid": {"DGEMM_B200_TILE": "mid"},
        "tiny": {"DGEMM_B200_TILE": "tiny"}, "tiny_sk": {"DGEMM_B200_TILE": "tiny", "DGEMM_B200_TINY_SK": "1"}, "auto": {}}
for (M, N, K) in SHAPES:
    dA, dB, dC = dg.malloc(8 * M * K), dg.malloc(8 * K * N), dg.malloc(8 * M * N)
    dg.fill_signed(dA, M, K, M, 1); dg.fill_signed(dB, K, N, K, 2); dg.fill_signed(dC, M, N, M, 3)
    dg.lib.dgemm_b200_sync(None)
    row = {"lib": os.path.basename(dg.LIB_PATH), "M": M, "N": N, "K": K}
    for name in TILES:
        for k in ("DGEMM_B200_TILE", "DGEMM_B200_TINY_SK"):
            os.environ.pop(k, None)
        os.environ.update(ENVS[name])
        dg.reload_env()
        reps = 20 if 2.0 * M * N * K < 1e12 else 5
        ms = dg.time_device(M, N, K, 1.0, dA, M, dB, K, 1.0, dC, M, warmup=3, reps=reps, flush_l2_bytes=256 << 20)
        row[name + "_us"] = round(ms * 1e3, 2); row[name + "_tf"] = round(2.0 * M * N * K / ms / 1e9, 3)
        p = dg.plan(M, N, K, dA, M, dB, K)
        row[name + "_plan"] = f"{p['ctas']}c/{p['stream_k_tiles']}sk"
    print(json.dumps(row), flush=True)
    for d in (dA, dB, dC): dg.free(d)
