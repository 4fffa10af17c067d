"""Helpers for the -m gpu parity tests: run the product DGEMM through the C ABI and compare with the oracle."""
import numpy as np

import oracle_lib as ol

BOUND_C = 2.0  # |C - C_ref| <= c*K*eps*(|alpha||A||B| + |beta*C0|), c = 2, eps = 2^-52


class DevMat:
    """A column-major matrix on the device, allocated through the C ABI (no torch)."""

    def __init__(self, dg, host, offset_elems=0):
        self.dg = dg
        self.ld, self.cols = host.shape
        self.nbytes = host.nbytes + 8 * offset_elems
        self.base = dg.lib.dgemm_b200_malloc(self.nbytes + 64)
        assert self.base, dg.last_error()
        self.ptr = self.base + 8 * offset_elems  # offset_elems=1 -> only 8-byte aligned (forces the non-TMA loader)
        st = dg.lib.dgemm_b200_h2d(self.ptr, self.ld, host.ctypes.data, self.ld, self.ld, self.cols, None)
        assert st == 0, dg.last_error()
        dg.lib.dgemm_b200_sync(None)

    def to_host(self):
        out = np.zeros((self.ld, self.cols), dtype=np.float64, order="F")
        st = self.dg.lib.dgemm_b200_d2h(out.ctypes.data, self.ld, self.ptr, self.ld, self.ld, self.cols, None)
        assert st == 0, self.dg.last_error()
        self.dg.lib.dgemm_b200_sync(None)
        return out

    def free(self):
        self.dg.lib.dgemm_b200_free(self.base)


def make_inputs(M, N, K, pad=(0, 0, 0), seed=1, dist="signed", c_fill=None):
    """A (ld=M+pad_a), B (ld=K+pad_b), C0 (ld=M+pad_c). Pad rows hold a sentinel that must survive."""
    lda, ldb, ldc = max(M, 1) + pad[0], max(K, 1) + pad[1], max(M, 1) + pad[2]
    if dist == "signed":
        A = ol.fill_signed(M, K, lda, seed, pad_value=0.0)
        B = ol.fill_signed(K, N, ldb, seed + 1, pad_value=0.0)
        C0 = ol.fill_signed(M, N, ldc, seed + 2, pad_value=-777.0)
    else:  # the reference harness' own distribution (utils.c:58-59), dense only
        assert pad == (0, 0, 0)
        A = ol.fill_reference_style(M, K, seed) if M * K else ol.colmajor(lda, K)
        B = ol.fill_reference_style(K, N, seed + 1) if K * N else ol.colmajor(ldb, N)
        C0 = ol.fill_reference_style(M, N, seed + 2) if M * N else ol.colmajor(ldc, N)
    if c_fill is not None:
        C0[:M, :] = c_fill
    return A, B, C0


def reference_result(M, N, K, alpha, A, B, beta, C0):
    """Oracle result (kernel18 restatement; BLAS semantics for beta == 0: C is not read)."""
    Cref = C0.copy(order="F")
    if beta == 0.0:
        Cref[:M, :] = 0.0  # the GPU path does not read C when beta == 0; neither may the check
    ol.oracle_k18(M, N, K, alpha, A, B, beta, Cref)
    return Cref


def check(M, N, K, alpha, A, B, beta, C0, Cgpu, Cref=None):
    if Cref is None:
        Cref = reference_result(M, N, K, alpha, A, B, beta, C0)
    if M > 0 and N > 0:
        C0b = C0 if beta != 0.0 else np.zeros_like(C0)
        ratio, max_abs = ol.bound_ratio(M, N, K, alpha, A, B, beta, C0b, Cgpu, Cref)
        assert ratio <= BOUND_C, f"bound ratio {ratio:.3g} > {BOUND_C} (max abs err {max_abs:.3g})"
    else:
        ratio = 0.0
    # caller-owned padding rows M..ld-1 must come back bit-identical
    assert np.array_equal(Cgpu[M:, :], C0[M:, :]), "padding rows of C were modified"
    return ratio


def run_device(dg, M, N, K, alpha, beta, pad=(0, 0, 0), variant=0, splits=0, seed=1, dist="signed", offset=(0, 0, 0),
               c_fill=None, nan_pad=False):
    A, B, C0 = make_inputs(M, N, K, pad, seed, dist, c_fill)
    if nan_pad:   # padding rows of A and B must never reach the result (TMA out-of-bounds fill, shifted parity maps)
        A[M:, :] = np.nan
        B[K:, :] = np.nan
    dA, dB, dC = DevMat(dg, A, offset[0]), DevMat(dg, B, offset[1]), DevMat(dg, C0, offset[2])
    try:
        dg.dgemm_device(M, N, K, alpha, dA.ptr, dA.ld, dB.ptr, dB.ld, beta, dC.ptr, dC.ld, 0, variant, splits)
        assert dg.lib.dgemm_b200_sync(None) == 0, dg.last_error()
        Cgpu = dC.to_host()
    finally:
        dA.free(); dB.free(); dC.free()
    return check(M, N, K, alpha, A, B, beta, C0, Cgpu)
