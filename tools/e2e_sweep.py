"""End-to-end (host pointers) timing of mydgemm_gpu_v20 over square sizes.
    python tools/e2e_sweep.py [sizes...] [--pageable] [--budget-mb=N]
--pageable: malloc'ed operands, as the reference harness passes them (test.c:27-30); default pinned.
--budget-mb: forces the out-of-core path (DGEMM_B200_HBM_BUDGET), the row then names the blocking."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import dgemm_b200 as dg
PAGEABLE = "--pageable" in sys.argv
BUDGET = next((int(x.split("=")[1]) << 20 for x in sys.argv if x.startswith("--budget-mb=")), 0)
if BUDGET:
    os.environ["DGEMM_B200_HBM_BUDGET"] = str(BUDGET)
    dg.reload_env()
for n in [int(x) for x in sys.argv[1:] if x.isdigit()] or [1024, 2048, 3072, 4096, 6144, 8192, 16384]:
    h = [torch.rand(n * n, dtype=torch.float64) for _ in range(3)]
    if not PAGEABLE:
        h = [x.pin_memory() for x in h]
    dg.mydgemm_gpu_v20(n, n, n, 1.0, h[0].data_ptr(), n, h[1].data_ptr(), n, 1.0, h[2].data_ptr(), n)
    best = 1e9
    for _ in range(4):
        t0 = time.perf_counter()
        dg.mydgemm_gpu_v20(n, n, n, 1.0, h[0].data_ptr(), n, h[1].data_ptr(), n, 1.0, h[2].data_ptr(), n)
        best = min(best, time.perf_counter() - t0)
    nb, kb, od, pred = dg.host_plan(n, n, n)
    extra = {"hbm_budget_mb": BUDGET >> 20, "blocking": dg.host_blocking(n, n, n, BUDGET), "device_bytes_held": dg.device_bytes()} if BUDGET else {}
    print(json.dumps({"n": n, "host_memory": "pageable" if PAGEABLE else "pinned", **extra, "plan": {"panels": len(nb) - 1, "slabs": len(kb) - 1, "calls": len(od), "model_ms": round(pred, 3)}, "e2e_ms": round(best * 1e3, 3), "e2e_tflops": round(2.0 * n**3 / best / 1e12, 3),
                      "pcie_floor_ms": round(24.0 * n * n / 55e9 * 1e3, 3)}), flush=True)
