"""CPU: the oracle against the UNMODIFIED reference compiled from /root/reference (oracle/_ref/libref.so).
Skipped where the compiled reference is absent or the host lacks AVX-512.  Bit-exact."""
import ctypes as C

import numpy as np
import pytest

import oracle_lib as ol

pytestmark = pytest.mark.skipif(not ol.have_ref(), reason="oracle/_ref/libref.so not built or no AVX-512")

# even M and N only: the reference drops odd rows/columns or crashes
SHAPES = [(2, 2, 1), (8, 8, 8), (24, 8, 4), (26, 10, 385), (64, 64, 64), (100, 100, 100), (192, 16, 384),
          (194, 18, 383), (300, 300, 300), (418, 30, 800), (512, 512, 1000), (1000, 1002, 997)]


@pytest.mark.parametrize("shape", SHAPES)
@pytest.mark.parametrize("alpha,beta,pad", [(1.0, 1.0, 0), (2.0, 0.0, 0), (2.5, 0.5, 7)])
def test_k18_bit_identical(shape, alpha, beta, pad):
    M, N, K = shape
    A = ol.fill_signed(M, K, M + pad, 1, pad_value=0.0)
    B = ol.fill_signed(K, N, K + pad, 2, pad_value=0.0)
    C0 = ol.fill_signed(M, N, M + pad, 3, pad_value=-777.0)
    mine, ref = C0.copy(order="F"), C0.copy(order="F")
    ol.oracle_k18(M, N, K, alpha, A, B, beta, mine)
    ol.call_gemm(ol.ref().mydgemm_cpu_v18, M, N, K, alpha, A, B, beta, ref)
    assert np.array_equal(mine, ref)
    assert np.array_equal(ref[M:], C0[M:])  # the reference leaves padding rows alone; so must every implementation


@pytest.mark.parametrize("nthr", [1, 2, 4])
@pytest.mark.parametrize("shape", [(64, 64, 64), (96, 40, 500), (512, 512, 1000), (1024, 1024, 1024)])
def test_k19_bit_identical(shape, nthr):
    M, N, K = shape  # every band M/nthr is even for these
    A = ol.fill_signed(M, K, M, 4); B = ol.fill_signed(K, N, K, 5); C0 = ol.fill_signed(M, N, M, 6)
    mine, ref = C0.copy(order="F"), C0.copy(order="F")
    ol.oracle_k19(M, N, K, -1.5, A, B, 0.5, mine, nthr)
    ol.ref().ref_set_num_threads(nthr)
    ol.call_gemm(ol.ref().mydgemm_cpu_v19, M, N, K, -1.5, A, B, 0.5, ref)
    assert np.array_equal(mine, ref)


@pytest.mark.parametrize("shape", [(7, 9, 13), (33, 17, 29), (64, 64, 64), (101, 53, 77)])
def test_k1_bit_identical_any_shape(shape):
    M, N, K = shape
    A = ol.fill_signed(M, K, M + 3, 7, pad_value=0.0)
    B = ol.fill_signed(K, N, K + 3, 8, pad_value=0.0)
    C0 = ol.fill_signed(M, N, M + 3, 9, pad_value=-777.0)
    mine, ref = C0.copy(order="F"), C0.copy(order="F")
    ol.oracle_k1(M, N, K, 2.5, A, B, 0.5, mine)
    ol.call_gemm(ol.ref().mydgemm_cpu_v1, M, N, K, 2.5, A, B, 0.5, ref)
    assert np.array_equal(mine, ref)


def test_reference_style_distribution():
    """D1 inputs (utils.c:58-59 with a fixed seed): two-decimal, non-negative; oracle == reference on them too."""
    n = 200
    A = ol.fill_reference_style(n, n, 1); B = ol.fill_reference_style(n, n, 2); C0 = ol.fill_reference_style(n, n, 3)
    assert A.min() >= 0 and A.max() < 100 and np.allclose(A * 100, np.round(A * 100))
    mine, ref = C0.copy(order="F"), C0.copy(order="F")
    ol.oracle_k18(n, n, n, 2.0, A, B, 0.0, mine)
    ol.call_gemm(ol.ref().mydgemm_cpu_v18, n, n, n, 2.0, A, B, 0.0, ref)
    assert np.array_equal(mine, ref)


@pytest.mark.parametrize("m", [1, 2, 7, 100, 1000, 2048, 4097])
@pytest.mark.parametrize("nthrs", [1, 2, 3, 7, 8, 16])
def test_partition_m_dim_matches_reference(m, nthrs):
    for ithr in range(nthrs):
        o1, b1, o2, b2 = C.c_int(), C.c_int(), C.c_int(), C.c_int()
        ol.oracle().oracle_partition_m_dim(ithr, nthrs, m, C.byref(o1), C.byref(b1))
        ol.ref().ref_partition_m_dim(ithr, nthrs, m, C.byref(o2), C.byref(b2))
        assert (o1.value, b1.value) == (o2.value, b2.value)
