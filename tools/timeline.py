"""Per-CTA phase timeline of one dgemm_tma_ws_kernel launch (dgemm_b200_debug_timeline):
    python tools/timeline.py M N K [splits]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import dgemm_b200 as dg
M, N, K = (int(x) for x in sys.argv[1:4])
splits = int(sys.argv[4]) if len(sys.argv) > 4 else 0
ROWS = 1024   # one row of 16 stamps per CTA (the grid is at most 2 CTAs x 148 SMs)
dA, dB, dC = dg.malloc(8 * M * K), dg.malloc(8 * K * N), dg.malloc(8 * M * N)
dg.fill_signed(dA, M, K, M, 1); dg.fill_signed(dB, K, N, K, 2); dg.fill_signed(dC, M, N, M, 3)
tl = dg.malloc(8 * ROWS * 16)
host = np.zeros((ROWS, 16), dtype=np.uint64)
FLUSH = int(os.environ.get("TL_FLUSH_MB", "0")) << 20   # > 0: evict the operands from L2 before every rep (as the sweep does)
fl = dg.malloc(FLUSH) if FLUSH else 0
for rep in range(3):
    if FLUSH: dg.fill_signed(fl, FLUSH // 8, 1, FLUSH // 8, rep)
    dg.lib.dgemm_b200_h2d(tl, ROWS * 16, host.ctypes.data, ROWS * 16, ROWS * 16, 1, None)
    dg.lib.dgemm_b200_sync(None)
    dg.lib.dgemm_b200_debug_timeline(tl)
    dg.dgemm_device(M, N, K, 1.0, dA, M, dB, K, float(os.environ.get("TL_BETA", "1.0")), dC, M, 0, int(os.environ.get("TL_VARIANT", "1")), splits)
    dg.lib.dgemm_b200_sync(None)
    dg.lib.dgemm_b200_debug_timeline(None)
out = np.zeros((ROWS, 16), dtype=np.uint64)
dg.lib.dgemm_b200_d2h(out.ctypes.data, ROWS * 16, tl, ROWS * 16, ROWS * 16, 1, None)
dg.lib.dgemm_b200_sync(None)
act = out[out[:, 0] > 0]
t0 = act[:, 0].min()
rel = (act.astype(np.int64) - np.int64(t0)) / 1e3
rel[act == 0] = np.nan
print("plan", dg.plan(M, N, K, dA, M, dB, K), "ctas", len(act))
names = ["entry", "cons.start"] + [f"seg{i//3}.{['landed','loop','epi'][i%3]}" for i in range(14)]
for j in range(16):
    col = rel[:, j]
    if np.all(np.isnan(col)): continue
    print(f"{names[j]:12s} min {np.nanmin(col):8.2f}  median {np.nanmedian(col):8.2f}  max {np.nanmax(col):8.2f} us  (n={np.sum(~np.isnan(col))})")
for i in range(4):   # per-CTA main-loop duration of segment i (landed -> loop end): a whole tile where the segment is one
    d = rel[:, 2 + 3 * i + 1] - rel[:, 2 + 3 * i]
    if not np.all(np.isnan(d)):
        print(f"seg{i} loop duration  min {np.nanmin(d):8.2f}  median {np.nanmedian(d):8.2f}  max {np.nanmax(d):8.2f} us")
print("kernel span (first entry -> last stamp): %.2f us" % np.nanmax(rel))
if os.environ.get("TL_DUMP"):   # raw per-CTA stamps (us relative to the first entry; NaN = not stamped), CTA = row
    import json
    json.dump({"M": M, "N": N, "K": K, "plan": dg.plan(M, N, K, dA, M, dB, K), "names": names,
               "stamps_us": [[None if np.isnan(v) else round(float(v), 3) for v in row] for row in rel]},
              open(os.environ["TL_DUMP"], "w"))
