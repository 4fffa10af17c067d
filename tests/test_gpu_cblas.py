"""-m gpu: the CBLAS-signature entry (dgemm_b200_cblas_dgemm) and its device-pointer twin (dgemm_b200_device_op).

The reference calls cblas_dgemm(CblasColMajor, CblasNoTrans, CblasNoTrans, ...) for kernel 0 and for its own
verification (test.c:38, utils.c:87-89).  The entry takes the full CBLAS surface: both layouts and op(A) / op(B).
Every case is checked against the oracle (kernel18 restatement) run on explicitly materialised column-major
op(A), op(B) and C, under the north-star bound."""
import itertools

import numpy as np
import pytest

import gpu_util as gu
import oracle_lib as ol

pytestmark = pytest.mark.gpu
ROW, COL, NT, TR = 101, 102, 111, 112


def stored(mat, layout, trans, pad, rng_fill=0.0):
    """Lay out the logical matrix `mat` (already op()'d: rows x cols) the way the caller would store it for the given
    layout / trans flag; returns (flat array holder, leading dimension)."""
    logical = mat.T if trans == TR else mat             # the matrix as the caller holds it (before op)
    r, c = logical.shape
    if layout == COL:
        ld = max(r, 1) + pad
        buf = np.full((ld, max(c, 1)), rng_fill, dtype=np.float64, order="F")
        buf[:r, :c] = logical
    else:                                               # row-major: rows contiguous, ld >= cols
        ld = max(c, 1) + pad
        buf = np.full((max(r, 1), ld), rng_fill, dtype=np.float64, order="C")
        buf[:r, :c] = logical
    return buf, ld


def expected(M, N, K, alpha, A, B, beta, C0):
    Af, Bf = np.asfortranarray(A), np.asfortranarray(B)
    C0f = np.asfortranarray(C0.copy())
    Cref = gu.reference_result(M, N, K, alpha, Af, Bf, beta, C0f)
    return Af, Bf, C0f, Cref


CASES = list(itertools.product([COL, ROW], [NT, TR], [NT, TR]))


@pytest.mark.parametrize("layout,ta,tb", CASES)
@pytest.mark.parametrize("shape,pad,alpha,beta", [
    ((300, 260, 200), 0, 1.0, 1.0),
    ((1000, 1003, 997), 0, 2.5, 0.5),
    ((257, 131, 65), 3, -1.0, 0.0),
    ((1500, 1100, 1300), 1, 1.0, 1.0),       # large enough for the pageable staging path
])
def test_cblas_host_entry(dg, layout, ta, tb, shape, pad, alpha, beta):
    M, N, K = shape
    A = np.array(ol.fill_signed(M, K, M, 11, pad_value=0.0)[:M, :K])
    B = np.array(ol.fill_signed(K, N, K, 12, pad_value=0.0)[:K, :N])
    C0 = np.array(ol.fill_signed(M, N, M, 13, pad_value=0.0)[:M, :N])
    Abuf, lda = stored(A, layout, ta, pad)
    Bbuf, ldb = stored(B, layout, tb, pad)
    Cbuf, ldc = stored(C0, layout, NT, pad, rng_fill=-777.0)
    sentinel = Cbuf.copy()
    dg.cblas_dgemm(layout, ta, tb, M, N, K, alpha, Abuf, lda, Bbuf, ldb, beta, Cbuf, ldc)
    got = Cbuf[:M, :N]
    Af, Bf, C0f, Cref = expected(M, N, K, alpha, A, B, beta, C0)
    C0b = C0f if beta != 0.0 else np.zeros_like(C0f)
    ratio, max_abs = ol.bound_ratio(M, N, K, alpha, Af, Bf, beta, C0b, np.asfortranarray(got), Cref)
    assert ratio <= gu.BOUND_C, f"bound ratio {ratio:.3g} (max abs err {max_abs:.3g})"
    # padding of C (rows M.. in column-major, columns N.. in row-major) must be untouched
    if layout == COL:
        assert np.array_equal(Cbuf[M:, :], sentinel[M:, :])
    else:
        assert np.array_equal(Cbuf[:, N:], sentinel[:, N:])


@pytest.mark.parametrize("layout,ta,tb", CASES)
def test_device_op(dg, layout, ta, tb):
    M, N, K, alpha, beta = 520, 390, 270, 1.5, 0.5
    A = np.array(ol.fill_signed(M, K, M, 21, pad_value=0.0)[:M, :K])
    B = np.array(ol.fill_signed(K, N, K, 22, pad_value=0.0)[:K, :N])
    C0 = np.array(ol.fill_signed(M, N, M, 23, pad_value=0.0)[:M, :N])
    Abuf, lda = stored(A, layout, ta, 2)
    Bbuf, ldb = stored(B, layout, tb, 2)
    Cbuf, ldc = stored(C0, layout, NT, 2)

    def dev(buf):  # raw bytes to the device, whatever the layout
        flat = np.frombuffer(buf.tobytes(order="A"), dtype=np.float64).reshape(-1, 1, order="F")
        return gu.DevMat(dg, np.asfortranarray(flat))

    dA, dB, dC = dev(Abuf), dev(Bbuf), dev(Cbuf)
    try:
        dg.dgemm_device_op(layout, ta, tb, M, N, K, alpha, dA.ptr, lda, dB.ptr, ldb, beta, dC.ptr, ldc)
        assert dg.lib.dgemm_b200_sync(None) == 0, dg.last_error()
        out = dC.to_host().reshape(Cbuf.shape, order="F" if layout == COL else "C")
    finally:
        dA.free(); dB.free(); dC.free()
    Af, Bf, C0f, Cref = expected(M, N, K, alpha, A, B, beta, C0)
    ratio, _ = ol.bound_ratio(M, N, K, alpha, Af, Bf, beta, C0f, np.asfortranarray(out[:M, :N]), Cref)
    assert ratio <= gu.BOUND_C


def test_notrans_colmajor_is_the_v20_path_bit_for_bit(dg):
    """cblas(ColMajor, N, N) and mydgemm_gpu_v20 are the same code path: identical bits."""
    M, N, K = 640, 512, 384
    A, B, C0 = gu.make_inputs(M, N, K)
    C1, C2 = C0.copy(order="F"), C0.copy(order="F")
    dg.mydgemm_gpu_v20(M, N, K, 2.0, A, M, B, K, 0.5, C1, M)
    dg.cblas_dgemm(COL, NT, NT, M, N, K, 2.0, A, M, B, K, 0.5, C2, M)
    assert np.array_equal(C1, C2)


def test_bad_codes_and_leading_dimensions(dg):
    A, B, C0 = gu.make_inputs(64, 48, 32)
    with pytest.raises(dg.DgemmError) as e:
        dg.cblas_dgemm(100, NT, NT, 64, 48, 32, 1.0, A, 64, B, 32, 0.0, C0, 64)
    assert e.value.status == dg.ERR_BAD_ARG
    with pytest.raises(dg.DgemmError):
        dg.cblas_dgemm(COL, 110, NT, 64, 48, 32, 1.0, A, 64, B, 32, 0.0, C0, 64)
    with pytest.raises(dg.DgemmError):   # A^T is stored K x M: LDA must be >= K
        dg.cblas_dgemm(COL, TR, NT, 32, 48, 64, 1.0, A, 32, B, 64, 0.0, C0, 64)
