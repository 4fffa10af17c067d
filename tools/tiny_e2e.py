"""Latency of mydgemm_gpu_v20 on tiny problems (pageable numpy operands, like the harness):
    python tools/tiny_e2e.py"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import dgemm_b200 as dg
for n in (64, 100, 128, 200, 256, 300):
    A = np.asfortranarray(np.random.rand(n, n)); B = np.asfortranarray(np.random.rand(n, n)); C = np.asfortranarray(np.random.rand(n, n))
    for _ in range(5): dg.mydgemm_gpu_v20(n, n, n, 2.0, A, n, B, n, 0.0, C, n)
    reps = 200
    t0 = time.perf_counter()
    for _ in range(reps): dg.mydgemm_gpu_v20(n, n, n, 2.0, A, n, B, n, 0.0, C, n)
    dt = (time.perf_counter() - t0) / reps
    print(json.dumps({"n": n, "us": round(dt * 1e6, 1), "gflops": round(2.0 * n**3 / dt / 1e9, 1),
                      "tiny_bytes": os.environ.get("DGEMM_B200_TINY_BYTES", "default")}), flush=True)
