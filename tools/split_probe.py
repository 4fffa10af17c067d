"""Kernel-only timing of one shape with chosen paddings of A and B (parity-split / cp.async halves experiments):
    [DGEMM_B200_FORCE_SPLIT=bits] [DGEMM_B200_ODD_LD=mode] python tools/split_probe.py M N K pad_a pad_b [reps]"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dgemm_b200 as dg
M, N, K, pa, pb = (int(x) for x in sys.argv[1:6])
reps = int(sys.argv[6]) if len(sys.argv) > 6 else 5
lda, ldb, ldc = M + pa, K + pb, M
dA, dB, dC = dg.malloc(8 * lda * K), dg.malloc(8 * ldb * N), dg.malloc(8 * ldc * N)
dg.fill_signed(dA, M, K, lda, 1); dg.fill_signed(dB, K, N, ldb, 2); dg.fill_signed(dC, M, N, ldc, 3)
dg.lib.dgemm_b200_sync(None)
ms = dg.time_device(M, N, K, 1.0, dA, lda, dB, ldb, 1.0, dC, ldc, warmup=2, reps=reps, flush_l2_bytes=256 << 20)
print(json.dumps({"M": M, "N": N, "K": K, "lda": lda, "ldb": ldb, "ms": round(ms, 4),
                  "tflops": round(2.0 * M * N * K / ms / 1e9, 2), "force_split": os.environ.get("DGEMM_B200_FORCE_SPLIT"),
                  "mode": os.environ.get("DGEMM_B200_ODD_LD", "split")}), flush=True)
