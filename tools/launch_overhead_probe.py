"""Where do the microseconds of a small call go?  Kernel-only event times (dgemm_b200_time_device, 30 reps, L2 flushed
between reps unless noted) of the same tiny / small problems under different launch configurations, next to the span the
CTAs themselves see (tools/timeline.py).  One JSON line per experiment.
    python tools/launch_overhead_probe.py"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dgemm_b200 as dg

KEYS = ("DGEMM_B200_NO_PDL", "DGEMM_B200_SMALL_CTAS", "DGEMM_B200_NO_SMALL_TILE", "DGEMM_B200_NO_TMAP_CACHE", "DGEMM_B200_NO_WAIT")


def run(label, M, N, K, env=None, variant=0, splits=0, flush=256 << 20, beta=1.0):
    for k in KEYS:
        os.environ.pop(k, None)
    os.environ.update(env or {})
    dg.reload_env()
    dA, dB, dC = dg.malloc(8 * M * K), dg.malloc(8 * K * N), dg.malloc(8 * M * N)
    dg.fill_signed(dA, M, K, M, 1); dg.fill_signed(dB, K, N, K, 2); dg.fill_signed(dC, M, N, M, 3)
    dg.lib.dgemm_b200_sync(None)
    ms = dg.time_device(M, N, K, 1.0, dA, M, dB, K, beta, dC, M, variant=variant, splits=splits, warmup=3, reps=30,
                        flush_l2_bytes=flush)
    print(json.dumps({"what": label, "M": M, "N": N, "K": K, "env": env or {}, "variant": variant, "splits": splits,
                      "flush_mb": flush >> 20, "beta": beta, "us": round(ms * 1e3, 2),
                      "tflops": round(2.0 * M * N * K / ms / 1e9, 3), "plan": dg.plan(M, N, K, dA, M, dB, K)}), flush=True)
    for d in (dA, dB, dC):
        dg.free(d)


run("floor: naive ladder kernel, 64^3", 64, 64, 64, variant=4, splits=1)
run("floor: naive ladder kernel, 64^3, no flush", 64, 64, 64, variant=4, splits=1, flush=0)
run("one small tile, 64^3", 64, 64, 64)
run("one small tile, 64^3, no flush", 64, 64, 64, flush=0)
run("one small tile, 64^3, no PDL attribute", 64, 64, 64, {"DGEMM_B200_NO_PDL": "1"})
run("one big tile, 128^2 x 64, explicit TMA variant", 128, 128, 64, variant=1, splits=1)
run("one big tile, 128^2 x 64, cp.async variant", 128, 128, 64, variant=3, splits=1)
for n in (256, 512):
    run(f"{n}^3 default", n, n, n)
    run(f"{n}^3 no flush", n, n, n, flush=0)
    run(f"{n}^3 no PDL attribute", n, n, n, {"DGEMM_B200_NO_PDL": "1"})
    run(f"{n}^3 waits dropped (no cooperative launch)", n, n, n, {"DGEMM_B200_NO_WAIT": "1"})
    run(f"{n}^3 one CTA per SM build", n, n, n, {"DGEMM_B200_SMALL_CTAS": "1"})
    run(f"{n}^3 one CTA per SM build, waits dropped", n, n, n, {"DGEMM_B200_SMALL_CTAS": "1", "DGEMM_B200_NO_WAIT": "1"})
    run(f"{n}^3 beta = 0", n, n, n, beta=0.0)
    run(f"{n}^3 big tile data-parallel (splits=1)", n, n, n, variant=1, splits=1)
    run(f"{n}^3 cp.async kernel auto split", n, n, n, variant=3, splits=0)

# ---- what the timing method costs for an EMPTY kernel, by launch shape (dgemm_b200_debug_launch_floor)
for label, ctas, threads, smem, flags in [
        ("empty kernel <<<1, 32>>>", 1, 32, 0, 0),
        ("empty kernel <<<148, 384>>>", 148, 384, 0, 0),
        ("empty kernel <<<148, 384, 199 KB>>> (the 128x128 kernel's shape)", 148, 384, 199 * 1024, 0),
        ("empty kernel <<<148, 384, 132 KB>>> (dg64)", 148, 384, 132 * 1024, 0),
        ("empty kernel <<<296, 384, 66 KB>>> (dg64x2)", 296, 384, 66 * 1024, 0),
        ("empty kernel <<<148, 384, 199 KB>>> + 512 B of parameters", 148, 384, 199 * 1024, 4),
        ("empty kernel <<<148, 384, 199 KB>>> + PDL attribute", 148, 384, 199 * 1024, 2),
        ("empty kernel <<<148, 384, 199 KB>>> cooperative", 148, 384, 199 * 1024, 1),
        ("empty kernel <<<128, 128, 48 KB>>> (a light data-parallel kernel)", 128, 128, 48 * 1024, 0)]:
    us = dg.lib.dgemm_b200_debug_launch_floor(ctas, threads, smem, flags, 50)
    print(json.dumps({"what": label, "ctas": ctas, "threads": threads, "smem": smem, "flags": flags, "us": round(us, 2)}), flush=True)
