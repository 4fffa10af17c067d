"""Python host-side mirror of the reference's dispatch interface for the DGEMM hot path, over the C ABI.

The reference is C: `test_kernel(kernel_num, m, n, k, alpha, A, B, beta, C)` switches to
`test_mydgemm_vNN`, which fixes LDA=m, LDB=k, LDC=m and calls the 11-argument kernel
(reference utils.c:159-165, 167-191; include/kernel18.h:629-640).  This module binds
`libdgemm_b200.so` (include/dgemm_b200.h) with ctypes and exposes the same three levels:

    mydgemm_gpu_v20(M, N, K, alpha, A, LDA, B, LDB, beta, C, LDC)   # host numpy arrays, in place
    test_mydgemm_v20(m, n, k, alpha, A, B, beta, C)                 # LDA=m, LDB=k, LDC=m
    test_kernel(kernel_num, m, n, k, alpha, A, B, beta, C)          # kernel_num 20 / 21

plus the device-pointer entry used for kernel-only timing.  There is no CPU fallback: if the CUDA
library is missing, importing this module raises; if there is no sm_100 device, calls raise DgemmError.
(The package directory name contains hyphens; `import dgemm_b200` at the repo root loads it.)
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("DGEMM_B200_LIB") or os.path.join(_HERE, "libdgemm_b200.so")  # env: experimental builds

OK, ERR_NO_DEVICE, ERR_BAD_ARG, ERR_OOM, ERR_CUDA, ERR_NCCL = range(6)
VARIANT_AUTO, VARIANT_TMA, VARIANT_CPASYNC8, VARIANT_CPASYNC16, VARIANT_NAIVE, VARIANT_DFMA, VARIANT_CLUSTER = range(7)
VARIANT_NAMES = {VARIANT_AUTO: "auto", VARIANT_TMA: "tma_ws", VARIANT_CPASYNC8: "cpasync8",
                 VARIANT_CPASYNC16: "cpasync16", VARIANT_NAIVE: "naive", VARIANT_DFMA: "dfma_regtile",
                 VARIANT_CLUSTER: "tma_cluster_multicast"}

# every symbol include/dgemm_b200.h declares (tests check the library exports all of them)
EXPORTS = [
    "mydgemm_gpu_v20", "mydgemm_gpu_v21", "dgemm_b200_last_status", "dgemm_b200_last_error",
    "dgemm_b200_device_count", "dgemm_b200_init", "dgemm_b200_shutdown", "dgemm_b200_set_max_gpus",
    "dgemm_b200_device", "dgemm_b200_device_ex", "dgemm_b200_device_fetch", "dgemm_b200_ipc_export",
    "dgemm_b200_ipc_open", "dgemm_b200_ipc_close", "dgemm_b200_plan", "dgemm_b200_malloc", "dgemm_b200_free",
    "dgemm_b200_host_alloc", "dgemm_b200_host_free", "dgemm_b200_h2d", "dgemm_b200_d2h", "dgemm_b200_sync",
    "dgemm_b200_fill_signed", "dgemm_b200_time_device", "dgemm_b200_launch_count",
    "dgemm_b200_reset_launch_count", "dgemm_b200_sm_clock_mhz", "dgemm_b200_debug_timeline", "dgemm_b200_debug_occupy_sms", "dgemm_b200_debug_launch_floor", "dgemm_b200_host_plan", "dgemm_b200_host_blocking", "dgemm_b200_tile_plan",
    "dgemm_b200_device_bytes",
    "dgemm_b200_cblas_dgemm", "dgemm_b200_device_op", "dgemm_b200_add_scaled", "dgemm_b200_rank_pipeline",
    "dgemm_b200_exchange_bytes", "dgemm_b200_exchange_alloc", "dgemm_b200_rank_plan", "dgemm_b200_reload_env",
    "dgemm_b200_device_static_scratch",
]


class DgemmError(RuntimeError):
    def __init__(self, status, message):
        super().__init__(f"dgemm_b200 status {status}: {message}")
        self.status = status


if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} is missing: build it with `make lib` (or __graft_entry__.build()). "
        "The DGEMM path is CUDA-only; there is no CPU fallback.")

lib = C.CDLL(LIB_PATH)
_dp = C.POINTER(C.c_double)
_vp = C.c_void_p
_ll = C.c_longlong

lib.mydgemm_gpu_v20.argtypes = [C.c_int, C.c_int, C.c_int, C.c_double, _vp, C.c_int, _vp, C.c_int, C.c_double, _vp,
                                C.c_int]
lib.mydgemm_gpu_v20.restype = None
lib.mydgemm_gpu_v21.argtypes = lib.mydgemm_gpu_v20.argtypes
lib.mydgemm_gpu_v21.restype = None
lib.dgemm_b200_last_error.restype = C.c_char_p
lib.dgemm_b200_init.argtypes = [C.c_int]
lib.dgemm_b200_set_max_gpus.argtypes = [C.c_int]
_DEV_ARGS = [C.c_int, C.c_int, C.c_int, C.c_double, _vp, _ll, _vp, _ll, C.c_double, _vp, _ll]
lib.dgemm_b200_device.argtypes = _DEV_ARGS + [_vp]
lib.dgemm_b200_device_ex.argtypes = _DEV_ARGS + [_vp, C.c_int, C.c_int]
lib.dgemm_b200_device_static_scratch.argtypes = _DEV_ARGS + [_vp, C.c_int]
lib.dgemm_b200_device_fetch.argtypes = _DEV_ARGS + [_vp, _vp, _vp, C.c_size_t]
lib.dgemm_b200_ipc_export.argtypes = [_vp, C.c_char_p]
lib.dgemm_b200_ipc_open.argtypes = [C.c_char_p, C.POINTER(_vp)]
lib.dgemm_b200_ipc_close.argtypes = [_vp]
lib.dgemm_b200_plan.argtypes = [C.c_int, C.c_int, C.c_int, _vp, _ll, _vp, _ll, C.POINTER(C.c_int),
                                C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]
lib.dgemm_b200_cblas_dgemm.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_double, _vp, C.c_int,
                                       _vp, C.c_int, C.c_double, _vp, C.c_int]
lib.dgemm_b200_cblas_dgemm.restype = None
lib.dgemm_b200_device_op.argtypes = [C.c_int, C.c_int, C.c_int] + _DEV_ARGS + [_vp]
lib.dgemm_b200_add_scaled.argtypes = [C.c_int, C.c_int, C.c_double, _vp, _ll, _vp, _ll, _vp]
_ip = C.POINTER(C.c_int)
lib.dgemm_b200_host_plan.argtypes = [C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, _ip, _ip, _ip, _ip, _ip, _ip,
                                     C.c_int, C.POINTER(C.c_double), _ip]
lib.dgemm_b200_rank_pipeline.argtypes = [C.c_int, C.c_int, C.c_int, C.c_double, _vp, C.c_int, _vp, C.c_int, C.c_double, _vp,
                                         C.c_int, C.c_int, C.c_int, _ip, C.POINTER(_vp), C.c_uint, C.POINTER(C.c_double),
                                         C.c_int, _ip]
lib.dgemm_b200_exchange_bytes.argtypes = [C.c_int, C.c_int]
lib.dgemm_b200_exchange_bytes.restype = C.c_size_t
lib.dgemm_b200_exchange_alloc.argtypes = [C.c_int, C.c_int]
lib.dgemm_b200_exchange_alloc.restype = _vp
lib.dgemm_b200_rank_plan.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _ip, C.c_int, C.c_int, _ip, C.c_int, _ip,
                                     _ip, C.c_int, _ip, _ip]
lib.dgemm_b200_malloc.argtypes = [C.c_size_t]
lib.dgemm_b200_malloc.restype = _vp
lib.dgemm_b200_free.argtypes = [_vp]
lib.dgemm_b200_free.restype = None
lib.dgemm_b200_host_alloc.argtypes = [C.c_size_t]
lib.dgemm_b200_host_alloc.restype = _vp
lib.dgemm_b200_host_free.argtypes = [_vp]
lib.dgemm_b200_host_free.restype = None
lib.dgemm_b200_h2d.argtypes = [_vp, _ll, _vp, _ll, _ll, _ll, _vp]
lib.dgemm_b200_d2h.argtypes = [_vp, _ll, _vp, _ll, _ll, _ll, _vp]
lib.dgemm_b200_sync.argtypes = [_vp]
lib.dgemm_b200_fill_signed.argtypes = [_vp, C.c_int, C.c_int, _ll, C.c_ulonglong, C.c_ulonglong, C.c_ulonglong, _vp]
lib.dgemm_b200_time_device.argtypes = _DEV_ARGS + [C.c_int, C.c_int, C.c_int, C.c_int, C.c_size_t]
lib.dgemm_b200_time_device.restype = C.c_double
lib.dgemm_b200_debug_timeline.argtypes = [_vp]
lib.dgemm_b200_debug_launch_floor.restype = C.c_double
lib.dgemm_b200_debug_launch_floor.argtypes = [C.c_int, C.c_int, C.c_uint, C.c_int, C.c_int]
lib.dgemm_b200_host_blocking.argtypes = [C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, C.c_double, C.POINTER(C.c_int),
                                         C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_double)]
lib.dgemm_b200_tile_plan.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]
lib.dgemm_b200_device_bytes.restype = C.c_size_t
lib.dgemm_b200_device_bytes.argtypes = []
lib.dgemm_b200_debug_occupy_sms.argtypes = [C.c_int, C.c_uint, C.c_double]
lib.dgemm_b200_reset_launch_count.restype = None
lib.dgemm_b200_sm_clock_mhz.restype = C.c_double


def last_status():
    return lib.dgemm_b200_last_status()


def last_error():
    return (lib.dgemm_b200_last_error() or b"").decode()


def _check(status=None):
    st = lib.dgemm_b200_last_status() if status is None else status
    if st != OK:
        raise DgemmError(st, last_error())


ROW_MAJOR, COL_MAJOR, NO_TRANS, TRANS, CONJ_TRANS = 101, 102, 111, 112, 113


def cblas_dgemm(layout, transa, transb, M, N, K, alpha, A, lda, B, ldb, beta, Cm, ldc):
    """cblas_dgemm's own signature (reference test.c:38) on host arrays / raw host addresses, in place."""
    pa = A.ctypes.data if isinstance(A, np.ndarray) else int(A)
    pb = B.ctypes.data if isinstance(B, np.ndarray) else int(B)
    pc = Cm.ctypes.data if isinstance(Cm, np.ndarray) else int(Cm)
    lib.dgemm_b200_cblas_dgemm(layout, transa, transb, M, N, K, alpha, pa, lda, pb, ldb, beta, pc, ldc)
    _check()


def add_scaled(M, N, beta, dX, ldx, dC, ldc, stream=0):
    """C += beta * X on device pointers (the deferred-C step of a pipelined caller)."""
    _check(lib.dgemm_b200_add_scaled(M, N, float(beta), int(dX), ldx, int(dC), ldc, int(stream) or None))


def dgemm_device_op(layout, transa, transb, M, N, K, alpha, dA, lda, dB, ldb, beta, dC, ldc, stream=0):
    _check(lib.dgemm_b200_device_op(layout, transa, transb, M, N, K, alpha, dA, lda, dB, ldb, beta, dC, ldc, stream))


def host_plan(M, N, K, alpha=1.0, beta=1.0, cap=4096):
    """Schedule mydgemm_gpu_v20 uses for this shape: (panel bounds, slab bounds, [(slab, panel, end_slab), ...], predicted
    ms); a call covers K-slabs [slab, end_slab) of one column panel.
    Pure host logic — works without a GPU."""
    nb, kb, od = (C.c_int * cap)(), (C.c_int * cap)(), (C.c_int * cap)()
    n_p, n_s, n_i, ms, n_d = C.c_int(), C.c_int(), C.c_int(), C.c_double(), C.c_int()
    _check(lib.dgemm_b200_host_plan(M, N, K, alpha, beta, nb, C.byref(n_p), kb, C.byref(n_s), od, C.byref(n_i), cap,
                                    C.byref(ms), C.byref(n_d)))
    host_plan.deferred_panels = n_d.value
    return (list(nb[: n_p.value + 1]), list(kb[: n_s.value + 1]),
            [(od[3 * i], od[3 * i + 1], od[3 * i + 2]) for i in range(n_i.value)], ms.value)


def host_blocking(M, N, K, budget_bytes, alpha=1.0, beta=1.0):
    """Out-of-core blocking of mydgemm_gpu_v20 under an HBM budget: {"Mb", "Nb", "Kb", "pcie_bytes"} (whole dimensions when
    the call fits in core).  Pure host logic — works without a GPU."""
    mb, nb, kb, by = C.c_int(), C.c_int(), C.c_int(), C.c_double()
    _check(lib.dgemm_b200_host_blocking(M, N, K, alpha, beta, float(budget_bytes), C.byref(mb), C.byref(nb), C.byref(kb),
                                        C.byref(by)))
    return {"Mb": mb.value, "Nb": nb.value, "Kb": kb.value, "pcie_bytes": by.value}


TILE_NAMES = {0: "128x128", 1: "64x64", 2: "64x64x2", 3: "128x64", 4: "64x32"}


def tile_plan(M, N, K, num_sms=148):
    """Which tile build of the TMA kernel a device call of this shape runs, its CTAs and stream-K tiles (no GPU needed)."""
    t, c, k = C.c_int(), C.c_int(), C.c_int()
    _check(lib.dgemm_b200_tile_plan(M, N, K, num_sms, C.byref(t), C.byref(c), C.byref(k)))
    return {"tile": TILE_NAMES[t.value], "ctas": c.value, "stream_k_tiles": k.value}


def device_bytes():
    """Bytes of device memory the library's cached buffers hold on the current device."""
    return int(lib.dgemm_b200_device_bytes())


def device_count():
    return lib.dgemm_b200_device_count()


def reload_env():
    """The DGEMM_B200_* switches are read once; call this after changing one in os.environ."""
    lib.dgemm_b200_reload_env()


# ---- multi-GPU: one rank's pass (one process per GPU) ------------------------------------------------------------
RANK_CALL_FIELDS = ("k0", "kc", "c0", "nc", "first", "last", "f_owner", "f_k0", "f_kc", "f_need_piece", "wait_copy",
                    "wait_x_copy")
RANK_COPY_FIELDS = ("kind", "k0", "kc", "c0", "nc", "piece")


def rank_plan(M, ncols, K, world, rank, kb, need_c_in=True, p1=0, cap=1 << 16):
    """Schedule of dgemm_b200_rank_pipeline as ({p1, calls: [dict], copies: [dict]}); pure host logic, no GPU needed."""
    kbv = (C.c_int * (world + 1))(*kb)
    calls, copies = (C.c_int * cap)(), (C.c_int * cap)()
    nc, ncp, p1o = C.c_int(), C.c_int(), C.c_int()
    _check(lib.dgemm_b200_rank_plan(M, ncols, K, world, rank, kbv, int(need_c_in), p1, calls, cap, C.byref(nc), copies, cap,
                                    C.byref(ncp), C.byref(p1o)))
    return {"p1": p1o.value,
            "calls": [dict(zip(RANK_CALL_FIELDS, calls[12 * i: 12 * i + 12])) for i in range(nc.value)],
            "copies": [dict(zip(RANK_COPY_FIELDS, copies[6 * i: 6 * i + 6])) for i in range(ncp.value)]}


def exchange_alloc(M, K):
    p = lib.dgemm_b200_exchange_alloc(M, K)
    if not p:
        _check()
        raise DgemmError(ERR_OOM, "dgemm_b200_exchange_alloc failed")
    return p


def rank_pipeline(M, ncols, K, alpha, A_slab, LDA, B_panel, LDB, beta, C_panel, LDC, rank, world, kb, xbuf, pass_no,
                  timeline_cap=0):
    """One rank's pass of the column-panel sharded DGEMM from host operands (numpy arrays or raw host addresses).
    Returns the timeline as a list of (kind, index, ms) when timeline_cap > 0."""
    kbv = (C.c_int * (world + 1))(*kb)
    xb = (_vp * world)(*[int(x) for x in xbuf])
    tl = (C.c_double * max(timeline_cap, 1))()
    n = C.c_int()
    st = lib.dgemm_b200_rank_pipeline(M, ncols, K, float(alpha), _host_ptr(A_slab), LDA, _host_ptr(B_panel), LDB, float(beta),
                                      _host_ptr(C_panel), LDC, rank, world, kbv, xb, pass_no,
                                      tl if timeline_cap else None, timeline_cap, C.byref(n))
    _check(st)
    return [(int(tl[i]), int(tl[i + 1]), tl[i + 2]) for i in range(0, n.value, 3)]


def _host_ptr(a):
    if isinstance(a, np.ndarray):
        if a.dtype != np.float64:
            raise TypeError("matrices must be float64")
        if not a.flags.f_contiguous:
            raise ValueError("matrices must be column-major (Fortran-ordered) numpy arrays of shape (ld, cols)")
        return a.ctypes.data
    return int(a)  # raw host address (e.g. pinned memory)


# ---- the reference's three dispatch levels ------------------------------------------------------------------
def mydgemm_gpu_v20(M, N, K, alpha, A, LDA, B, LDB, beta, Cm, LDC):
    """C = alpha*A*B + beta*C on one B200; host operands; same 11 arguments as mydgemm_cpu_v18."""
    lib.mydgemm_gpu_v20(M, N, K, float(alpha), _host_ptr(A), LDA, _host_ptr(B), LDB, float(beta), _host_ptr(Cm), LDC)
    _check()


def mydgemm_gpu_v21(M, N, K, alpha, A, LDA, B, LDB, beta, Cm, LDC):
    """Same call sharded over all visible B200s (column panels of B and C, A replicated with NCCL)."""
    lib.mydgemm_gpu_v21(M, N, K, float(alpha), _host_ptr(A), LDA, _host_ptr(B), LDB, float(beta), _host_ptr(Cm), LDC)
    _check()


def test_mydgemm_v20(m, n, k, alpha, A, B, beta, Cm):
    """reference utils.c:159-165 wrapper shape: LDA=m, LDB=k, LDC=m."""
    mydgemm_gpu_v20(m, n, k, alpha, A, m, B, k, beta, Cm, m)


def test_mydgemm_v21(m, n, k, alpha, A, B, beta, Cm):
    mydgemm_gpu_v21(m, n, k, alpha, A, m, B, k, beta, Cm, m)


def test_kernel(kernel_num, m, n, k, alpha, A, B, beta, Cm):
    """reference utils.c:167-191 switch, for the two kernel numbers this package adds."""
    if kernel_num == 20:
        test_mydgemm_v20(m, n, k, alpha, A, B, beta, Cm)
    elif kernel_num == 21:
        test_mydgemm_v21(m, n, k, alpha, A, B, beta, Cm)
    else:
        raise ValueError("this package provides kernel_num 20 (one B200) and 21 (all B200s of the box)")


# ---- device-pointer entry -----------------------------------------------------------------------------------------
def dgemm_device(M, N, K, alpha, dA, lda, dB, ldb, beta, dC, ldc, stream=0, variant=VARIANT_AUTO, splits=0):
    """Asynchronous DGEMM on device pointers of the current CUDA device (ints / data_ptr())."""
    st = lib.dgemm_b200_device_ex(M, N, K, float(alpha), int(dA), lda, int(dB), ldb, float(beta), int(dC), ldc,
                                  int(stream) or None, variant, splits)
    _check(st)


STATIC_SCRATCH_BYTES = 16 << 20   # DGEMM_B200_STATIC_SCRATCH_BYTES: 128 partial tiles of 128 x 128 doubles


def dgemm_device_static_scratch(M, N, K, alpha, dA, lda, dB, ldb, beta, dC, ldc, stream=0, splits=0):
    """Split-K DGEMM on device pointers that allocates nothing (include/dgemm_b200.h: dgemm_b200_device_static_scratch).

    For callers that cannot let the library call cudaMalloc (their own pool holds the HBM, or they cannot afford the
    device-wide stall of cudaMalloc / cudaFree).  The partial tiles live in a fixed 16 MiB array inside the library, so at
    most 128 (tile, split) units fit: larger calls raise DgemmError(ERR_BAD_ARG).  Calls on different streams race."""
    st = lib.dgemm_b200_device_static_scratch(M, N, K, float(alpha), int(dA), lda, int(dB), ldb, float(beta), int(dC), ldc,
                                              int(stream) or None, splits)
    _check(st)


def dgemm_device_fetch(M, N, K, alpha, dA, lda, dB, ldb, beta, dC, ldc, stream=0, fetch_src=0, fetch_dst=0,
                       fetch_bytes=0):
    """dgemm_device fused with an NVLink pull of fetch_bytes from fetch_src (peer memory) to fetch_dst."""
    st = lib.dgemm_b200_device_fetch(M, N, K, float(alpha), int(dA), lda, int(dB), ldb, float(beta), int(dC), ldc,
                                     int(stream) or None, int(fetch_src) or None, int(fetch_dst) or None, fetch_bytes)
    _check(st)


def ipc_export(dptr):
    buf = C.create_string_buffer(64)
    _check(lib.dgemm_b200_ipc_export(int(dptr), buf))
    return buf.raw


def ipc_open(handle):
    out = _vp()
    _check(lib.dgemm_b200_ipc_open(handle, C.byref(out)))
    return out.value


def ipc_close(dptr):
    _check(lib.dgemm_b200_ipc_close(int(dptr)))


def malloc(nbytes):
    p = lib.dgemm_b200_malloc(nbytes)
    if not p:
        _check()
        raise DgemmError(ERR_OOM, "dgemm_b200_malloc failed")
    return p


def free(dptr):
    lib.dgemm_b200_free(int(dptr))


def plan(M, N, K, dA=0, lda=None, dB=0, ldb=None):
    v, s, c, k = C.c_int(), C.c_int(), C.c_int(), C.c_int()
    st = lib.dgemm_b200_plan(M, N, K, int(dA) or None, lda or max(M, 1), int(dB) or None, ldb or max(K, 1),
                             C.byref(v), C.byref(s), C.byref(c), C.byref(k))
    _check(st)
    return {"variant": VARIANT_NAMES[v.value], "splits": s.value, "ctas": c.value, "stream_k_tiles": k.value}


def time_device(M, N, K, alpha, dA, lda, dB, ldb, beta, dC, ldc, variant=VARIANT_AUTO, splits=0, warmup=3, reps=5,
                flush_l2_bytes=0):
    ms = lib.dgemm_b200_time_device(M, N, K, float(alpha), int(dA), lda, int(dB), ldb, float(beta), int(dC), ldc,
                                    variant, splits, warmup, reps, flush_l2_bytes)
    if ms < 0:
        _check()
        raise DgemmError(ERR_CUDA, "timed run failed")
    return ms


def fill_signed(dX, rows, cols, ld, seed, row0=0, col0=0, stream=0):
    _check(lib.dgemm_b200_fill_signed(int(dX), rows, cols, ld, seed, row0, col0, int(stream) or None))


def launch_count():
    return lib.dgemm_b200_launch_count()


def reset_launch_count():
    lib.dgemm_b200_reset_launch_count()


def sm_clock_mhz():
    return lib.dgemm_b200_sm_clock_mhz()


def shutdown():
    lib.dgemm_b200_shutdown()
