"""Generates tests/golden/ref_vectors.npz by running the UNMODIFIED reference kernels (oracle/_ref/libref.so,
compiled from /root/reference by oracle/Makefile) on small seeded inputs.  Run in the build container:

    make -C oracle all && python tests/golden/make_golden.py

The reference keeps no golden vectors of its own for this path, so these are "outputs of the
reference itself run here".  Inputs are regenerated from seeds by the oracle's counter-based generator; only the
reference OUTPUTS are stored (float64, exact bits).  Shapes stay inside the reference's valid domain (even M, N;
every kernel19 thread band even; N <= 9216) except the kernel1 cases, which take any shape.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import oracle_lib as ol  # noqa: E402

# (kernel, M, N, K, pad, alpha, beta, seed, threads)
CASES = [
    ("v18", 64, 64, 64, 0, 1.0, 1.0, 101, 1),
    ("v18", 100, 100, 100, 0, 2.0, 0.0, 102, 1),        # the harness' own (alpha, beta), test.c:25
    ("v18", 26, 10, 385, 3, 1.0, 0.25, 103, 1),         # every edge micro-kernel: m 24/2, n 8/2, K crossing 384
    ("v18", 250, 38, 800, 1, 3.0, 2.0, 104, 1),         # M > M_BLOCKING=192, three K blocks
    ("v18", 48, 30, 0, 0, 1.0, 0.5, 105, 1),            # K == 0 -> C = beta*C
    ("v18", 48, 30, 17, 0, 0.0, 0.5, 106, 1),           # alpha == 0 -> C = beta*C
    ("v19", 96, 40, 500, 2, -1.5, 1.0, 107, 2),
    ("v19", 192, 24, 384, 0, 1.0, 0.0, 108, 4),
    ("v19", 48, 30, 17, 0, 0.0, 0.5, 109, 2),           # F4: kernel19 returns BEFORE scaling by beta
    ("v1", 33, 17, 29, 5, 2.5, 0.5, 110, 1),            # odd everything: only kernel1 is valid here
    ("v1", 7, 9, 13, 0, 1.0, 1.0, 111, 1),
]


def run_case(kernel, M, N, K, pad, alpha, beta, seed, threads):
    A = ol.fill_signed(M, K, max(M, 1) + pad, seed, pad_value=0.0)
    B = ol.fill_signed(K, N, max(K, 1) + pad, seed + 1000, pad_value=0.0)
    C = ol.fill_signed(M, N, max(M, 1) + pad, seed + 2000, pad_value=-777.0)
    ref = ol.ref()
    ref.ref_set_num_threads(threads)
    ol.call_gemm(getattr(ref, "mydgemm_cpu_" + kernel), M, N, K, alpha, A, B, beta, C)
    return C


if __name__ == "__main__":
    assert ol.have_ref(), "oracle/_ref/libref.so missing (make -C oracle all) or no AVX-512 on this host"
    out = {}
    for i, case in enumerate(CASES):
        out[f"case{i}"] = run_case(*case)
    np.savez_compressed(os.path.join(HERE, "ref_vectors.npz"), **out)
    print("wrote", os.path.join(HERE, "ref_vectors.npz"), {k: v.shape for k, v in out.items()})
