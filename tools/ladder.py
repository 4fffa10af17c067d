"""The Blackwell optimisation ladder: every rung timed on the same shapes.
    python tools/ladder.py [sizes...] > profiles/rNN_ladder.jsonl"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dgemm_b200 as dg
RUNGS = [("1 naive (one thread per entry)", 4, 1), ("2 smem tiles + 8x8 register tile, DFMA", 5, 1),
         ("3 DMMA + cp.async 4-stage", 3, 1), ("4 DMMA + TMA/mbarrier warp-specialised persistent", 1, 1),
         ("5 + stream-K tail (product path)", 1, 0),
         ("6 rung 4 on 2-CTA clusters, A by TMA multicast", 6, 1)]
for n in [int(x) for x in sys.argv[1:]] or [2048, 4096, 8192]:
    dA, dB, dC = dg.malloc(8 * n * n), dg.malloc(8 * n * n), dg.malloc(8 * n * n)
    dg.fill_signed(dA, n, n, n, 1); dg.fill_signed(dB, n, n, n, 2); dg.fill_signed(dC, n, n, n, 3)
    for name, variant, splits in RUNGS:
        if variant == 4 and n > 4096:
            continue
        ms = dg.time_device(n, n, n, 1.0, dA, n, dB, n, 1.0, dC, n, variant=variant, splits=splits, warmup=1,
                            reps=2 if variant == 4 else 5, flush_l2_bytes=256 << 20 if n < 8192 else 0)
        print(json.dumps({"n": n, "rung": name, "ms": round(ms, 4), "tflops": round(2.0 * n**3 / ms / 1e9, 3),
                          "pct_fp64_peak": round(100 * 2.0 * n**3 / ms / 1e9 / 37.225, 2)}), flush=True)
    for p in (dA, dB, dC):
        dg.free(p)
