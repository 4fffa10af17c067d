"""Odd-leading-dimension shapes (no tensor map possible): cp.async kernel vs realign pre-pass + TMA kernel.
    DGEMM_B200_REALIGN_MIN_FLOPS=... python tools/odd_ld_probe.py"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dgemm_b200 as dg
for (M, N, K) in [(500, 501, 499), (1000, 1003, 997), (1500, 1501, 1499), (2001, 2003, 1999), (3001, 3003, 2999), (4097, 4096, 4095)]:
    lda, ldb, ldc = M, K, M
    dA, dB, dC = dg.malloc(8 * lda * K), dg.malloc(8 * ldb * N), dg.malloc(8 * ldc * N)
    dg.fill_signed(dA, M, K, lda, 1); dg.fill_signed(dB, K, N, ldb, 2); dg.fill_signed(dC, M, N, ldc, 3)
    dg.lib.dgemm_b200_sync(None)
    ms = dg.time_device(M, N, K, 2.5, dA, lda, dB, ldb, 0.5, dC, ldc, warmup=2, reps=10, flush_l2_bytes=256 << 20)
    print(json.dumps({"M": M, "N": N, "K": K, "ms": round(ms, 4), "tflops": round(2.0 * M * N * K / ms / 1e9, 2),
                      "realign_min_flops": os.environ.get("DGEMM_B200_REALIGN_MIN_FLOPS", "default")}), flush=True)
    for d in (dA, dB, dC): dg.free(d)
