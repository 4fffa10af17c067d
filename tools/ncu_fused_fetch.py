"""Evidence for the fused compute + NVLink kernel (needs 2 GPUs, ONE process): the same DGEMM call on GPU 0 twice, once
pulling a K-slab of A from GPU 1's HBM with its spare warps (dgemm_b200_device_fetch), once without.

    python tools/ncu_fused_fetch.py [M N K]          -> CUDA-event times of both calls (JSON line)
    ncu --metrics gpu__time_duration.sum,nvlrx__bytes.sum,nvltx__bytes.sum,dram__bytes_read.sum,dram__bytes_write.sum,\
lts__t_bytes.sum,smsp__pipe_tensor_subpipe_dmma_cycles_active.avg,sm__cycles_elapsed.avg -k regex:dgemm_tma_ws_kernel ...

Default shape = one rank's call of the 8-GPU 65536^3 pass scaled to fit quickly: M = 16384, N = 8192, K = 8192, pulling a
slab of the same size (M x K doubles = 1.07 GB): the algorithmic NVLink volume per call."""
import json
import sys
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import ctypes as C

import numpy as np

import dgemm_b200 as dg

M, N, K = (int(x) for x in sys.argv[1:4]) if len(sys.argv) > 3 else (16384, 8192, 8192)
cudart = C.CDLL("libcudart.so.12")
assert dg.device_count() >= 2, "needs two GPUs"
# peer access both ways: a small v21 call sets it up (and proves the path works on this box)
dg.lib.dgemm_b200_set_max_gpus(2)
a = np.asfortranarray(np.ones((512, 512))); b = np.asfortranarray(np.ones((512, 512))); c = np.asfortranarray(np.zeros((512, 512)))
dg.mydgemm_gpu_v21(512, 512, 512, 1.0, a, 512, b, 512, 0.0, c, 512)
assert abs(c[3, 400] - 512.0) < 1e-9
cudart.cudaSetDevice(1)
src = dg.malloc(8 * M * K)
dg.fill_signed(src, M, K, M, 7)
dg.lib.dgemm_b200_sync(None)
cudart.cudaSetDevice(0)
dA, dB, dC, dst = dg.malloc(8 * M * K), dg.malloc(8 * K * N), dg.malloc(8 * M * N), dg.malloc(8 * M * K)
dg.fill_signed(dA, M, K, M, 1); dg.fill_signed(dB, K, N, K, 2); dg.fill_signed(dC, M, N, M, 3)
dg.lib.dgemm_b200_sync(None)
e = [C.c_void_p() for _ in range(2)]
for x in e:
    cudart.cudaEventCreate(C.byref(x))


def timed(fetch):
    ms = C.c_float()
    for rep in range(3):
        cudart.cudaEventRecord(e[0], None)
        if fetch:
            dg.dgemm_device_fetch(M, N, K, 1.0, dA, M, dB, K, 1.0, dC, M, 0, src, dst, 8 * M * K)
        else:
            dg.dgemm_device(M, N, K, 1.0, dA, M, dB, K, 1.0, dC, M)
        cudart.cudaEventRecord(e[1], None)
        cudart.cudaEventSynchronize(e[1])
        cudart.cudaEventElapsedTime(C.byref(ms), e[0], e[1])
    return ms.value


t_plain, t_fetch = timed(False), timed(True)
# the pulled slab must be bit-identical to its source
h0, h1 = np.zeros(1024), np.zeros(1024)
off = 8 * (M * (K - 1) + M - 1024)
dg.lib.dgemm_b200_d2h(h0.ctypes.data, 1024, dst + off, 1024, 1024, 1, None)
dg.lib.dgemm_b200_sync(None)
cudart.cudaSetDevice(1)
dg.lib.dgemm_b200_d2h(h1.ctypes.data, 1024, src + off, 1024, 1024, 1, None)
dg.lib.dgemm_b200_sync(None)
cudart.cudaSetDevice(0)
assert np.array_equal(h0, h1), "pulled slab differs from its source"
print(json.dumps({"M": M, "N": N, "K": K, "ms_plain": round(t_plain, 4), "ms_with_fetch": round(t_fetch, 4),
                  "tflops_plain": round(2.0 * M * N * K / t_plain / 1e9, 2), "tflops_with_fetch": round(2.0 * M * N * K / t_fetch / 1e9, 2),
                  "fetch_bytes": 8 * M * K, "nvlink_gb_per_s_needed": round(8 * M * K / t_fetch / 1e6, 1)}))
