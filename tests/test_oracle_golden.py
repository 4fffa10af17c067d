"""CPU: the oracle (oracle/dgemm_oracle.c) against the committed golden outputs of the reference kernels
(tests/golden/ref_vectors.npz, produced by tests/golden/make_golden.py from oracle/_ref/libref.so).
Bit-exact: the oracle restates the reference's per-entry operation order."""
import importlib.util
import os

import numpy as np
import pytest

import oracle_lib as ol

HERE = os.path.dirname(os.path.abspath(__file__))
spec = importlib.util.spec_from_file_location("make_golden", os.path.join(HERE, "golden", "make_golden.py"))
mg = importlib.util.module_from_spec(spec)
spec.loader.exec_module(mg)
GOLD = np.load(os.path.join(HERE, "golden", "ref_vectors.npz"))


@pytest.mark.parametrize("idx", range(len(mg.CASES)))
def test_oracle_reproduces_reference_bits(idx):
    kernel, M, N, K, pad, alpha, beta, seed, threads = mg.CASES[idx]
    A = ol.fill_signed(M, K, max(M, 1) + pad, seed, pad_value=0.0)
    B = ol.fill_signed(K, N, max(K, 1) + pad, seed + 1000, pad_value=0.0)
    C = ol.fill_signed(M, N, max(M, 1) + pad, seed + 2000, pad_value=-777.0)
    if kernel == "v18":
        ol.oracle_k18(M, N, K, alpha, A, B, beta, C)
    elif kernel == "v19":
        ol.oracle_k19(M, N, K, alpha, A, B, beta, C, threads)
    else:
        ol.oracle_k1(M, N, K, alpha, A, B, beta, C)
    want = GOLD[f"case{idx}"]
    assert C.shape == want.shape
    assert np.array_equal(C, want), f"case {idx} {mg.CASES[idx]}: max diff {np.max(np.abs(C - want))}"


def test_kernel19_quirk_is_kept_only_in_k19():
    """kernel19 returns before the beta scaling when alpha == 0; kernel18 scales first."""
    M, N, K = 16, 12, 8
    A = ol.fill_signed(M, K, M, 1); B = ol.fill_signed(K, N, K, 2); C0 = ol.fill_signed(M, N, M, 3)
    c18, c19 = C0.copy(order="F"), C0.copy(order="F")
    ol.oracle_k18(M, N, K, 0.0, A, B, 0.5, c18)
    ol.oracle_k19(M, N, K, 0.0, A, B, 0.5, c19, 2)
    assert np.array_equal(c18, 0.5 * C0) and np.array_equal(c19, C0)


def test_oracle_vs_long_double_dot_products():
    """Independent check of the oracle itself: within the north-star bound of exact (long double) dot products."""
    M, N, K = 37, 29, 1000
    A = ol.fill_signed(M, K, M, 5); B = ol.fill_signed(K, N, K, 6); C0 = ol.fill_signed(M, N, M, 7)
    C = C0.copy(order="F")
    ol.oracle_k18(M, N, K, 2.5, A, B, 0.5, C)
    exact = (2.5 * (A.astype(np.longdouble) @ B.astype(np.longdouble)) + 0.5 * C0.astype(np.longdouble))
    ratio, _ = ol.bound_ratio(M, N, K, 2.5, A, B, 0.5, C0, C, np.asfortranarray(exact.astype(np.float64)))
    assert ratio <= 1.0


def test_signed_generator_matches_formula():
    x = ol.fill_signed(5, 4, 7, 42, row0=3, col0=9, pad_value=0.0)
    lib = ol.oracle()
    for i in range(5):
        for j in range(4):
            assert x[i, j] == lib.oracle_signed_value(42, 3 + i, 9 + j)
    assert np.all(x[5:] == 0.0) and np.all(np.abs(x[:5]) < 1.0)
