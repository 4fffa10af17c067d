"""Kernel-only timing sweep through the C ABI (no torch): python tools/quick_bench.py [sizes...]"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dgemm_b200 as dg

sizes = [int(s) for s in sys.argv[1:]] or [1024, 2048, 4096, 8192]
L2_FLUSH = 256 << 20
for n in sizes:
    nbytes = 8 * n * n
    dA, dB, dC = (dg.lib.dgemm_b200_malloc(nbytes) for _ in range(3))
    dg.fill_signed(dA, n, n, n, 1); dg.fill_signed(dB, n, n, n, 2); dg.fill_signed(dC, n, n, n, 3)
    dg.lib.dgemm_b200_sync(None)
    for variant in (1, 3, 2):
        reps = 3 if n >= 8192 else 10
        ms = dg.time_device(n, n, n, 1.0, dA, n, dB, n, 1.0, dC, n, variant=variant, splits=0, warmup=2, reps=reps,
                            flush_l2_bytes=L2_FLUSH if n < 8192 else 0)
        print(json.dumps({"n": n, "variant": dg.VARIANT_NAMES[variant], "ms": round(ms, 4),
                          "tflops": round(2.0 * n**3 / ms / 1e9, 3), "plan": dg.plan(n, n, n, dA, n, dB, n),
                          "sm_mhz": round(dg.sm_clock_mhz(), 1)}), flush=True)
    for p in (dA, dB, dC):
        dg.lib.dgemm_b200_free(p)
