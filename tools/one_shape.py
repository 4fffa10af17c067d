"""Runs one shape a few times through the device API (for ncu / timing experiments):
    python tools/one_shape.py M N K [variant] [splits] [flush_mb]"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dgemm_b200 as dg
M, N, K = (int(x) for x in sys.argv[1:4])
variant = int(sys.argv[4]) if len(sys.argv) > 4 else 0
splits = int(sys.argv[5]) if len(sys.argv) > 5 else 0
flush = (int(sys.argv[6]) if len(sys.argv) > 6 else 256) << 20
dA, dB, dC = dg.malloc(8 * M * K), dg.malloc(8 * K * N), dg.malloc(8 * M * N)
dg.fill_signed(dA, M, K, M, 1); dg.fill_signed(dB, K, N, K, 2); dg.fill_signed(dC, M, N, M, 3)
dg.lib.dgemm_b200_sync(None)
ms = dg.time_device(M, N, K, 1.0, dA, M, dB, K, 1.0, dC, M, variant=variant, splits=splits, warmup=2, reps=5,
                    flush_l2_bytes=flush)
print(json.dumps({"M": M, "N": N, "K": K, "variant": variant, "splits": splits, "flush_mb": flush >> 20, "ms": ms,
                  "tflops": 2.0 * M * N * K / ms / 1e9, "plan": dg.plan(M, N, K, dA, M, dB, K)}))
