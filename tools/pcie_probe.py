"""Pinned-host <-> HBM copy rates through the library's own 2-D copy helpers (calibrates csrc/dgemm_hostplan.cpp):
    python tools/pcie_probe.py > profiles/rNN_pcie_probe.jsonl"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dgemm_b200 as dg

rows = 16384
for cols in [16, 128, 1024, 4096, 16384]:
    nbytes = 8 * rows * cols
    h = dg.lib.dgemm_b200_host_alloc(nbytes)
    d = dg.malloc(nbytes)
    out = {"rows": rows, "cols": cols, "MB": nbytes / 1e6}
    for name, fn, args in [("h2d", dg.lib.dgemm_b200_h2d, (d, rows, h, rows)), ("d2h", dg.lib.dgemm_b200_d2h, (h, rows, d, rows))]:
        fn(*args, rows, cols, None); dg.lib.dgemm_b200_sync(None)
        reps = 5
        t0 = time.perf_counter()
        for _ in range(reps):
            fn(*args, rows, cols, None)
        dg.lib.dgemm_b200_sync(None)
        dt = (time.perf_counter() - t0) / reps
        out[name + "_GBps"] = round(nbytes / dt / 1e9, 2)
        out[name + "_us"] = round(dt * 1e6, 1)
    # both directions at once (two buffers, two streams are inside the library only for the GEMM entry; here: serial)
    print(json.dumps(out), flush=True)
    dg.lib.dgemm_b200_host_free(h); dg.free(d)
