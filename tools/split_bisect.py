"""Runs single parity-split cases in separate processes (a device fault kills only that process):
    python tools/split_bisect.py"""
import os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CASES = [  # M N K pad_a pad_b off_a off_b splits
    (128, 128, 16, 0, 0, 0, 0, 1), (128, 128, 64, 0, 0, 0, 0, 1), (1000, 1004, 1000, 0, 0, 0, 0, 0), (128, 128, 16, 1, 0, 0, 0, 1), (128, 128, 16, 0, 1, 0, 0, 1),
    (128, 128, 16, 0, 0, 1, 0, 1), (128, 128, 16, 0, 0, 0, 1, 1), (128, 128, 64, 1, 1, 0, 0, 1),
    (1000, 1003, 997, 0, 0, 0, 0, 1), (1000, 1003, 997, 0, 0, 0, 0, 0), (1000, 1003, 997, 7, 7, 0, 0, 0),
]
if len(sys.argv) > 1:
    sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
    import numpy as np
    import dgemm_b200 as dg, gpu_util as gu
    M, N, K, pa, pb, oa, ob, splits = (int(x) for x in sys.argv[1:9])
    r = gu.run_device(dg, M, N, K, 1.5, 0.5, pad=(pa, pb, 0), offset=(oa, ob, 0), variant=1, splits=splits, nan_pad=True)
    print("ok ratio %.3g" % r)
else:
    for c in CASES:
        p = subprocess.run([sys.executable, __file__] + [str(x) for x in c], capture_output=True, text=True, timeout=120)
        tail = (p.stdout.strip().splitlines() or [""])[-1] if p.returncode == 0 else (p.stderr.strip().splitlines() or [""])[-1]
        print(c, "rc", p.returncode, tail[:160], flush=True)
