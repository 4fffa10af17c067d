"""Odd-leading-dimension shapes (no single tensor map possible): parity-split tensor maps (default) vs realign
pre-pass + TMA kernel vs cp.async kernel.
    DGEMM_B200_ODD_LD=split|realign|cpasync python tools/odd_ld_probe.py"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dgemm_b200 as dg
for (M, N, K, pa) in [(500, 501, 499, 0), (1000, 1003, 997, 0), (1000, 1003, 997, 7), (1500, 1501, 1499, 0),
                      (2001, 2003, 1999, 0), (3001, 3003, 2999, 0), (4097, 4096, 4095, 0), (8192, 8192, 8192, 1)]:
    lda, ldb, ldc = M + pa, K + pa, M + pa
    dA, dB, dC = dg.malloc(8 * lda * K), dg.malloc(8 * ldb * N), dg.malloc(8 * ldc * N)
    dg.fill_signed(dA, M, K, lda, 1); dg.fill_signed(dB, K, N, ldb, 2); dg.fill_signed(dC, M, N, ldc, 3)
    dg.lib.dgemm_b200_sync(None)
    ms = dg.time_device(M, N, K, 2.5, dA, lda, dB, ldb, 0.5, dC, ldc, warmup=2, reps=10, flush_l2_bytes=256 << 20)
    print(json.dumps({"M": M, "N": N, "K": K, "ms": round(ms, 4), "tflops": round(2.0 * M * N * K / ms / 1e9, 2),
                      "lda": lda, "ldb": ldb, "mode": os.environ.get("DGEMM_B200_ODD_LD", "split")}), flush=True)
    for d in (dA, dB, dC): dg.free(d)
