"""End-to-end timing of mydgemm_gpu_v21 (one process, all visible GPUs) from pinned host buffers:
    python tools/v21_e2e.py [sizes...]"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import dgemm_b200 as dg
for n in [int(x) for x in sys.argv[1:]] or [8192, 16384, 32768]:
    h = [torch.rand(n * n, dtype=torch.float64).pin_memory() for _ in range(3)]
    dg.mydgemm_gpu_v21(n, n, n, 1.0, h[0].data_ptr(), n, h[1].data_ptr(), n, 1.0, h[2].data_ptr(), n)
    best = 1e9
    for _ in range(3):
        t0 = time.perf_counter()
        dg.mydgemm_gpu_v21(n, n, n, 1.0, h[0].data_ptr(), n, h[1].data_ptr(), n, 1.0, h[2].data_ptr(), n)
        best = min(best, time.perf_counter() - t0)
    print(json.dumps({"n": n, "gpus": torch.cuda.device_count(), "e2e_ms": round(best * 1e3, 2),
                      "e2e_tflops": round(2.0 * n ** 3 / best / 1e12, 2)}), flush=True)
    del h
