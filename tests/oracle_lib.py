"""ctypes loaders for the CHECKER libraries (test infrastructure only):
  oracle/liboracle.so      plain-C restatement of kernel18/kernel19/kernel1 arithmetic
  oracle/_ref/libref.so    the unmodified reference kernels, compiled by oracle/Makefile
Nothing in the product package imports this module."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
_dp = C.POINTER(C.c_double)
_GEMM_ARGS = [C.c_int, C.c_int, C.c_int, C.c_double, _dp, C.c_int, _dp, C.c_int, C.c_double, _dp, C.c_int]


def build_oracle():
    subprocess.run(["make", "-s", "-C", ORACLE_DIR, "all"], check=True, stdout=subprocess.DEVNULL)


def _ptr(a):
    assert a.dtype == np.float64
    return a.ctypes.data_as(_dp)


_oracle = None
_ref = None


def oracle():
    global _oracle
    if _oracle is None:
        path = os.path.join(ORACLE_DIR, "liboracle.so")
        if not os.path.exists(path):
            build_oracle()
        lib = C.CDLL(path)
        for name in ("oracle_dgemm_k18", "oracle_dgemm_k1"):
            getattr(lib, name).argtypes = _GEMM_ARGS
            getattr(lib, name).restype = None
        lib.oracle_dgemm_k19.argtypes = _GEMM_ARGS + [C.c_int]
        lib.oracle_dgemm_k19.restype = None
        lib.oracle_partition_m_dim.argtypes = [C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        lib.oracle_bound_term.argtypes = [C.c_int, C.c_int, C.c_int, C.c_double, _dp, C.c_int, _dp, C.c_int,
                                          C.c_double, _dp, C.c_int, _dp]
        lib.oracle_max_bound_ratio.argtypes = [C.c_int, C.c_int, C.c_int, _dp, C.c_int, _dp, C.c_int, _dp, _dp]
        lib.oracle_max_bound_ratio.restype = C.c_double
        lib.oracle_fill_reference_style.argtypes = [_dp, C.c_size_t, C.c_uint]
        lib.oracle_fill_signed.argtypes = [_dp, C.c_int, C.c_int, C.c_int, C.c_uint64, C.c_uint64, C.c_uint64]
        lib.oracle_signed_value.argtypes = [C.c_uint64, C.c_uint64, C.c_uint64]
        lib.oracle_signed_value.restype = C.c_double
        lib.oracle_max_threads.restype = C.c_int
        _oracle = lib
    return _oracle


def have_ref():
    return os.path.exists(os.path.join(ORACLE_DIR, "_ref", "libref.so")) and _cpu_has_avx512()


def _cpu_has_avx512():
    try:
        with open("/proc/cpuinfo") as f:
            return "avx512f" in f.read()
    except OSError:
        return False


def ref():
    """The unmodified reference kernels (needs AVX-512 on the host)."""
    global _ref
    if _ref is None:
        lib = C.CDLL(os.path.join(ORACLE_DIR, "_ref", "libref.so"))
        for name in ("mydgemm_cpu_v1", "mydgemm_cpu_v18", "mydgemm_cpu_v19"):
            getattr(lib, name).argtypes = _GEMM_ARGS
            getattr(lib, name).restype = None
        lib.ref_partition_m_dim.argtypes = [C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        lib.ref_set_num_threads.argtypes = [C.c_int]
        lib.ref_get_max_threads.restype = C.c_int
        _ref = lib
    return _ref


# ---- matrices: column-major numpy arrays of shape (ld, cols), element (i, j) = X[i, j] in Fortran order ----
def colmajor(ld, cols, fill=None):
    x = np.zeros((ld, cols), dtype=np.float64, order="F")
    if fill is not None:
        x[...] = fill
    return x


def fill_signed(rows, cols, ld, seed, row0=0, col0=0, pad_value=np.nan):
    """D2 distribution: U(-1,1), 53-bit, counter based. Pad rows rows..ld-1 get pad_value."""
    x = colmajor(ld, cols, pad_value)
    oracle().oracle_fill_signed(_ptr(x), rows, cols, ld, seed, row0, col0)
    return x


def fill_reference_style(rows, cols, seed):
    """D1 distribution: utils.c:58-59 with a fixed seed; dense (ld == rows)."""
    x = colmajor(rows, cols)
    oracle().oracle_fill_reference_style(_ptr(x), rows * cols, seed)
    return x


def call_gemm(fn, M, N, K, alpha, A, B, beta, C_, extra=()):
    fn(M, N, K, float(alpha), _ptr(A), A.shape[0], _ptr(B), B.shape[0], float(beta), _ptr(C_), C_.shape[0], *extra)


def oracle_k18(M, N, K, alpha, A, B, beta, C_):
    call_gemm(oracle().oracle_dgemm_k18, M, N, K, alpha, A, B, beta, C_)


def oracle_k19(M, N, K, alpha, A, B, beta, C_, nthr):
    call_gemm(oracle().oracle_dgemm_k19, M, N, K, alpha, A, B, beta, C_, (nthr,))


def oracle_k1(M, N, K, alpha, A, B, beta, C_):
    call_gemm(oracle().oracle_dgemm_k1, M, N, K, alpha, A, B, beta, C_)


def bound_ratio(M, N, K, alpha, A, B, beta, C0, Cx, Cy):
    """max |Cx-Cy| / (K*eps*(|alpha||A||B| + |beta*C0|)) over the M x N window, and max |Cx-Cy|."""
    lib = oracle()
    S = np.zeros((max(M, 1), max(N, 1)), dtype=np.float64, order="F")
    lib.oracle_bound_term(M, N, K, float(alpha), _ptr(A), A.shape[0], _ptr(B), B.shape[0], float(beta), _ptr(C0),
                          C0.shape[0], _ptr(S))
    mx = C.c_double(0.0)
    r = lib.oracle_max_bound_ratio(M, N, K, _ptr(Cx), Cx.shape[0], _ptr(Cy), Cy.shape[0], _ptr(S), C.byref(mx))
    return r, mx.value
