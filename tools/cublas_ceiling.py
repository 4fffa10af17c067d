"""Ceiling reference only (never on the product path): cuBLAS cublasDgemm via torch.matmul(float64).
Prints TFLOPS for a few square sizes and the kernel names cuBLAS picked."""
import json, sys, torch
sizes = [int(s) for s in sys.argv[1:]] or [2048, 4096, 8192, 16384]
dev = torch.device("cuda:0")
for n in sizes:
    a = torch.rand(n, n, dtype=torch.float64, device=dev) - 0.5
    b = torch.rand(n, n, dtype=torch.float64, device=dev) - 0.5
    for _ in range(2): c = a @ b
    torch.cuda.synchronize()
    best = 1e30
    for _ in range(3):
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record(); c = a @ b; e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    print(json.dumps({"cublas_dgemm": n, "ms": best, "tflops": 2.0 * n**3 / best / 1e9}), flush=True)
    del a, b, c
from torch.profiler import profile, ProfilerActivity
a = torch.rand(4096, 4096, dtype=torch.float64, device=dev); b = torch.rand(4096, 4096, dtype=torch.float64, device=dev)
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    c = a @ b; torch.cuda.synchronize()
for ev in prof.key_averages():
    print(json.dumps({"cublas_kernel": ev.key, "cuda_us": ev.device_time_total}))
