"""The 64 x 64-tile builds of the TMA kernel (one / two CTAs per SM) against the 128 x 128 one, kernel-only (CUDA events, L2
flushed between reps; LD = dimension, so 1000 x 1003 x 997 runs the parity-split kernels):
    python tools/small_tile_probe.py [M N K ...]"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dgemm_b200 as dg
args = [int(x) for x in sys.argv[1:]]
shapes = [tuple(args[i:i + 3]) for i in range(0, len(args), 3)] or \
    [(n, n, n) for n in (100, 200, 256, 300, 384, 500, 512, 640, 768, 900, 1000, 1024, 1280, 1536, 2048, 3000)] + \
    [(1000, 1003, 997), (512, 512, 4096), (512, 512, 16384), (256, 256, 8192), (1024, 1024, 256), (2048, 2048, 128), (128, 4096, 512), (4096, 128, 512)]
for (M, N, K) in shapes:
    dA, dB, dC = dg.malloc(8 * M * K), dg.malloc(8 * K * N), dg.malloc(8 * M * N)
    dg.fill_signed(dA, M, K, M, 1); dg.fill_signed(dB, K, N, K, 2); dg.fill_signed(dC, M, N, M, 3)
    dg.lib.dgemm_b200_sync(None)
    row = {"M": M, "N": N, "K": K}
    for name, env in (("big", {"DGEMM_B200_NO_SMALL_TILE": "1"}),
                      ("small", {"DGEMM_B200_SMALL_TILE_MAX_TILES": "100000", "DGEMM_B200_SMALL_TILE_MAX_FLOPS": "1e30",
                                 "DGEMM_B200_SMALL_CTAS": "1"}),
                      ("small_x2", {"DGEMM_B200_SMALL_TILE_MAX_TILES": "100000", "DGEMM_B200_SMALL_TILE_MAX_FLOPS": "1e30",
                                    "DGEMM_B200_SMALL_CTAS": "2"}),
                      ("mid", {"DGEMM_B200_TILE": "mid"}),
                      ("auto", {})):
        if name == "small_x2" and os.environ.get("PROBE_NO_X2"):
            continue
        for k in ("DGEMM_B200_NO_SMALL_TILE", "DGEMM_B200_SMALL_TILE_MAX_TILES", "DGEMM_B200_SMALL_TILE_MAX_FLOPS",
                  "DGEMM_B200_SMALL_CTAS", "DGEMM_B200_TILE"):
            os.environ.pop(k, None)
        os.environ.update(env)
        dg.reload_env()   # the library reads its switches once
        ms = dg.time_device(M, N, K, 1.0, dA, M, dB, K, 1.0, dC, M, warmup=3, reps=20, flush_l2_bytes=256 << 20)
        row[name + "_us"] = round(ms * 1e3, 2); row[name + "_tflops"] = round(2.0 * M * N * K / ms / 1e9, 3)
        row[name + "_plan"] = dg.plan(M, N, K, dA, M, dB, K)
    row["small_over_big"] = round(row["big_us"] / row["small_us"], 3)
    if "small_x2_us" in row:
        row["x2_over_big"] = round(row["big_us"] / row["small_x2_us"], 3)
        row["x2_over_small"] = round(row["small_us"] / row["small_x2_us"], 3)
    row["mid_over_best"] = round(min(row["big_us"], row["small_us"]) / row["mid_us"], 3)
    print(json.dumps(row), flush=True)
    for d in (dA, dB, dC): dg.free(d)
