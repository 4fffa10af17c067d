"""Small shapes through every kernel path, for compute-sanitizer (memcheck / racecheck / synccheck):
TMA data-parallel, TMA stream-K (owner + contributors), forced split-K + reduce, cp.async 8/16, edges, scale-only,
fused fetch (local copy), host-pointer 2-D pipeline."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import dgemm_b200 as dg
import gpu_util as gu

cases = [  # M, N, K, alpha, beta, pad, variant, splits
    (256, 256, 64, 1.0, 1.0, (0, 0, 0), 1, 1),
    (384, 256, 512, 1.5, 0.5, (0, 0, 0), 1, 0),      # stream-K, several ranges per tile
    (1920, 1664, 128, 1.0, 1.0, (0, 0, 0), 1, 0),    # data-parallel wave + stream-K tail
    (200, 190, 100, 2.5, 0.5, (0, 2, 0), 1, 1),      # TMA edges (zero fill), predicated epilogue
    (256, 128, 256, 1.0, 0.0, (0, 0, 0), 1, 4),      # split-K + reduce
    (130, 70, 97, 1.0, 1.0, (1, 0, 1), 2, 0),        # cp.async 8 B, odd LD
    (256, 256, 128, 1.0, 1.0, (0, 0, 0), 3, 2),      # cp.async 16 B + split-K
    (64, 64, 0, 1.0, 0.5, (0, 0, 0), 0, 0),          # scale only
]
for (M, N, K, alpha, beta, pad, variant, splits) in cases:
    r = gu.run_device(dg, M, N, K, alpha, beta, pad=pad, variant=variant, splits=splits)
    print("ok", M, N, K, variant, splits, "ratio %.3g" % r, flush=True)
# fused fetch with a local source
A, B, C0 = gu.make_inputs(256, 256, 256)
dA, dB, dC = gu.DevMat(dg, A), gu.DevMat(dg, B), gu.DevMat(dg, C0)
src, dst = dg.malloc(1 << 20), dg.malloc(1 << 20)
dg.fill_signed(src, 1 << 17, 1, 1 << 17, 9)
dg.dgemm_device_fetch(256, 256, 256, 1.0, dA.ptr, 256, dB.ptr, 256, 1.0, dC.ptr, 256, 0, src, dst, 1 << 20)
dg.lib.dgemm_b200_sync(None)
h1, h2 = np.zeros(1 << 17), np.zeros(1 << 17)
dg.lib.dgemm_b200_d2h(h1.ctypes.data, 1 << 17, src, 1 << 17, 1 << 17, 1, None)
dg.lib.dgemm_b200_d2h(h2.ctypes.data, 1 << 17, dst, 1 << 17, 1 << 17, 1, None)
dg.lib.dgemm_b200_sync(None)
assert np.array_equal(h1, h2)
gu.check(256, 256, 256, 1.0, A, B, 1.0, C0, dC.to_host())
print("ok fused fetch", flush=True)
# host-pointer entry
A, B, C0 = gu.make_inputs(300, 520, 200, (1, 1, 1)); Cg = C0.copy(order="F")
dg.mydgemm_gpu_v20(300, 520, 200, 1.5, A, 301, B, 201, 0.5, Cg, 301)
gu.check(300, 520, 200, 1.5, A, B, 0.5, C0, Cg)
print("ok v20", flush=True)
dg.shutdown()
