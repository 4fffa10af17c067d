"""BASELINE.json configs[1] and [2]: square size sweep and edge/irregular shapes on one B200.
Kernel-only (CUDA events, operands resident, L2 flushed between reps for operands smaller than L2) and end-to-end
(mydgemm_gpu_v20 with pinned host buffers), next to cuBLAS cublasDgemm as a ceiling reference.
    python tools/sweep.py > profiles/rNN_sweep.jsonl"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import dgemm_b200 as dg

PEAK = 148 * 128 * 1.965e9 / 1e12
dev = torch.device("cuda:0")
L2_FLUSH = 256 << 20


def cublas_ms(M, N, K, a, b, reps):
    A = a[:M * K].view(K, M).t(); B = b[:K * N].view(N, K).t()
    for _ in range(2): c = A @ B
    torch.cuda.synchronize()
    tot = 0.0
    flush = torch.empty(L2_FLUSH, dtype=torch.uint8, device=dev)
    for i in range(reps):
        flush.fill_(i)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); c = A @ B; e1.record(); torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / reps


def run(M, N, K, alpha=1.0, beta=1.0, ld_pad=(0, 0, 0), e2e=True, label=None):
    lda, ldb, ldc = M + ld_pad[0], K + ld_pad[1], M + ld_pad[2]
    a = torch.empty(lda * K, dtype=torch.float64, device=dev)
    b = torch.empty(ldb * N, dtype=torch.float64, device=dev)
    c = torch.empty(ldc * N, dtype=torch.float64, device=dev)
    dg.fill_signed(a.data_ptr(), lda, K, lda, 1); dg.fill_signed(b.data_ptr(), ldb, N, ldb, 2)
    dg.fill_signed(c.data_ptr(), ldc, N, ldc, 3)
    torch.cuda.synchronize()
    flops = 2.0 * M * N * K
    reps = 3 if flops > 5e11 else 10
    flush = L2_FLUSH if 8 * (M * K + K * N + M * N) < (512 << 20) else 0
    ms = dg.time_device(M, N, K, alpha, a.data_ptr(), lda, b.data_ptr(), ldb, beta, c.data_ptr(), ldc, warmup=2,
                        reps=reps, flush_l2_bytes=flush)
    row = {"M": M, "N": N, "K": K, "alpha": alpha, "beta": beta, "lda": lda, "ldb": ldb, "ldc": ldc,
           "plan": dg.plan(M, N, K, a.data_ptr(), lda, b.data_ptr(), ldb), "kernel_ms": round(ms, 4),
           "kernel_tflops": round(flops / ms / 1e9, 3), "pct_fp64_peak": round(100 * flops / ms / 1e9 / PEAK, 2),
           "algorithmic_GBps": round(8.0 * (M * K + K * N + (2 if beta else 1) * M * N) / ms / 1e6, 1)}
    if label: row["label"] = label
    if ld_pad == (0, 0, 0):
        cms = cublas_ms(M, N, K, a, b, reps)
        row["cublas_ms"] = round(cms, 4); row["cublas_tflops"] = round(flops / cms / 1e9, 3)
    if e2e:
        hA = torch.empty(lda * K, dtype=torch.float64).pin_memory(); hB = torch.empty(ldb * N, dtype=torch.float64).pin_memory()
        hC = torch.empty(ldc * N, dtype=torch.float64).pin_memory()
        hA.copy_(a); hB.copy_(b); hC.copy_(c); torch.cuda.synchronize()
        dg.mydgemm_gpu_v20(M, N, K, alpha, hA.data_ptr(), lda, hB.data_ptr(), ldb, beta, hC.data_ptr(), ldc)
        t0 = time.perf_counter()
        for _ in range(3):
            dg.mydgemm_gpu_v20(M, N, K, alpha, hA.data_ptr(), lda, hB.data_ptr(), ldb, beta, hC.data_ptr(), ldc)
        dt = (time.perf_counter() - t0) / 3
        row["e2e_ms"] = round(dt * 1e3, 4); row["e2e_tflops"] = round(flops / dt / 1e12, 3)
    print(json.dumps(row), flush=True)


for n in [256, 512, 1024, 1536, 2048, 3072, 4096, 6144, 8192, 12288, 16384]:
    run(n, n, n, label="square")
run(2048, 2048, 2048, alpha=2.0, beta=0.0, label="harness pair (test.c:25)")
run(512, 512, 16384, label="tall-skinny")
run(1000, 1003, 997, alpha=2.5, beta=0.5, label="edge, LD=dim (LDB odd -> B by column parity: TMA + cp.async halves)")
run(1000, 1003, 997, alpha=2.5, beta=0.5, ld_pad=(7, 7, 7), label="edge, LD=dim+7 (LDA odd -> A by k parity)")
run(1000, 1003, 997, alpha=2.5, beta=0.5, ld_pad=(24, 27, 24), label="edge, LD=1024 (TMA)")
run(1001, 1003, 997, alpha=2.5, beta=0.5, ld_pad=(23, 27, 23), label="edge odd M, LD=1024 (TMA)")
run(3000, 3000, 3000, alpha=2.0, beta=0.0, label="reference's largest published size")
