"""dgemm_b200_device_static_scratch — the DGEMM entry for callers that cannot let the library allocate.

Some programs must keep cudaMalloc out of their hot path: their own pool has already taken the HBM (a library-side
cudaMalloc would then fail, or steal from the pool), or they run many streams and cannot afford the device-wide lock and
the device-wide wait that cudaMalloc / cudaFree bring.  For them the library keeps the split-K partial tiles in a fixed
16 MiB array of its own image instead of a grown workspace (include/dgemm_b200.h).  These tests drive that entry through
the C ABI and hold it to the existing entry: same bits as dgemm_b200_device_ex with the same variant and split count, and
inside the oracle's bound.  Two calls are made back to back on one stream, the second smaller than the first, so the second
runs on an array still full of the first call's partial tiles.

This file sorts last on purpose (pytest -x stops at the first failure, so a failure here hides nothing behind it).  It has
to pass alone in a fresh process — the first call on the device does the entry's one-time set-up — and after every other
test has used the library.
"""
import numpy as np
import pytest

import gpu_util as gu

TMA, CPASYNC8 = 1, 2


def _run_pair(dg, shapes, pad, offset, variant, sync_between):
    """shapes: two (M, N, K, alpha, beta, splits); returns after checking both results against device_ex and the oracle."""
    mats = []
    try:
        for i, (M, N, K, alpha, beta, splits) in enumerate(shapes):
            A, B, C0 = gu.make_inputs(M, N, K, pad, seed=11 + 7 * i)
            dA, dB = gu.DevMat(dg, A, offset), gu.DevMat(dg, B, offset)
            dC_new, dC_old = gu.DevMat(dg, C0), gu.DevMat(dg, C0)
            mats.append((A, B, C0, dA, dB, dC_new, dC_old))
        # the two calls under test, one after the other on the default stream
        for (M, N, K, alpha, beta, splits), (A, B, C0, dA, dB, dC_new, dC_old) in zip(shapes, mats):
            dg.dgemm_device_static_scratch(M, N, K, alpha, dA.ptr, dA.ld, dB.ptr, dB.ld, beta, dC_new.ptr, dC_new.ld, 0,
                                           splits)
            if sync_between:
                assert dg.lib.dgemm_b200_sync(None) == 0, dg.last_error()
        assert dg.lib.dgemm_b200_sync(None) == 0, dg.last_error()
        for (M, N, K, alpha, beta, splits), (A, B, C0, dA, dB, dC_new, dC_old) in zip(shapes, mats):
            got = dC_new.to_host()
            gu.check(M, N, K, alpha, A, B, beta, C0, got)           # the oracle's bound, padding rows untouched
            if splits:                                             # same kernels, same split count: the same bits
                dg.dgemm_device(M, N, K, alpha, dA.ptr, dA.ld, dB.ptr, dB.ld, beta, dC_old.ptr, dC_old.ld, 0, variant, splits)
                assert dg.lib.dgemm_b200_sync(None) == 0, dg.last_error()
                assert np.array_equal(got, dC_old.to_host()), "differs from dgemm_b200_device_ex with the same splits"
    finally:
        for m in mats:
            for d in m[3:]:
                d.free()


@pytest.mark.gpu
@pytest.mark.parametrize("sync_between", [False, True])
def test_two_calls_in_a_row_match_the_existing_entry(dg, sync_between):
    # 512 x 512 is 16 tiles; 8 splits fill all 128 slots of the array.  The second call (6 tiles x 4 splits) then reads
    # and writes slots that still hold the first call's partials.
    shapes = [(512, 512, 4096, 1.0, 1.0, 8), (200, 300, 1000, -2.5, 0.5, 4)]
    _run_pair(dg, shapes, pad=(0, 0, 0), offset=0, variant=TMA, sync_between=sync_between)


@pytest.mark.gpu
def test_unaligned_operands_take_the_cp_async_kernels(dg):
    # odd leading dimensions and an 8-byte aligned base: no tensor map can describe them, the cp.async loader runs
    shapes = [(257, 130, 999, 2.0, 0.0, 3), (129, 131, 515, 1.0, 1.0, 2)]
    _run_pair(dg, shapes, pad=(0, 0, 7), offset=1, variant=CPASYNC8, sync_between=False)


@pytest.mark.gpu
def test_chosen_split_and_degenerate_calls(dg):
    # splits = 0: the entry picks (checked against the oracle only), also for a shape with a single k-tile
    shapes = [(384, 256, 2048, 1.0, 1.0, 0), (100, 60, 16, 0.5, 2.0, 0)]
    _run_pair(dg, shapes, pad=(3, 1, 5), offset=0, variant=TMA, sync_between=False)
    # alpha == 0 and K == 0 scale C; M == 0 / N == 0 do nothing
    M, N = 70, 33
    A, B, C0 = gu.make_inputs(M, N, 8, (0, 0, 2), seed=5)
    dA, dB, dC = gu.DevMat(dg, A), gu.DevMat(dg, B), gu.DevMat(dg, C0)
    try:
        dg.dgemm_device_static_scratch(M, N, 8, 0.0, dA.ptr, dA.ld, dB.ptr, dB.ld, 0.5, dC.ptr, dC.ld)
        dg.dgemm_device_static_scratch(M, N, 0, 3.0, dA.ptr, dA.ld, dB.ptr, dB.ld, 2.0, dC.ptr, dC.ld)   # together: x 1.0
        dg.dgemm_device_static_scratch(0, N, 8, 1.0, dA.ptr, dA.ld, dB.ptr, dB.ld, 0.0, dC.ptr, dC.ld)
        dg.dgemm_device_static_scratch(M, 0, 8, 1.0, dA.ptr, dA.ld, dB.ptr, dB.ld, 0.0, dC.ptr, dC.ld)
        assert dg.lib.dgemm_b200_sync(None) == 0, dg.last_error()
        assert np.array_equal(dC.to_host(), C0)
    finally:
        dA.free(); dB.free(); dC.free()


def _free_device_bytes():
    """Free device memory as the driver reports it, or None when no stand-alone libcudart can be loaded."""
    import ctypes
    for name in ("libcudart.so", "libcudart.so.12", "/usr/local/cuda/lib64/libcudart.so"):
        try:
            rt = ctypes.CDLL(name)
        except OSError:
            continue
        free, total = ctypes.c_size_t(0), ctypes.c_size_t(0)
        if rt.cudaMemGetInfo(ctypes.byref(free), ctypes.byref(total)) == 0:
            return free.value
    return None


@pytest.mark.gpu
def test_calls_beyond_the_scratch_are_refused_and_nothing_is_allocated(dg):
    M = N = 2048   # 256 output tiles: more than the 64 the array can serve at two splits
    K = 256
    A, B, C0 = gu.make_inputs(M, N, K, seed=3)
    dA, dB, dC = gu.DevMat(dg, A), gu.DevMat(dg, B), gu.DevMat(dg, C0)
    try:
        held = dg.lib.dgemm_b200_device_bytes()
        with pytest.raises(dg.DgemmError) as e:
            dg.dgemm_device_static_scratch(M, N, K, 1.0, dA.ptr, dA.ld, dB.ptr, dB.ld, 1.0, dC.ptr, dC.ld)
        assert e.value.status == dg.ERR_BAD_ARG and "static scratch" in str(e.value)
        with pytest.raises(dg.DgemmError) as e:   # 16 tiles x 9 splits = 144 > 128
            dg.dgemm_device_static_scratch(512, 512, K, 1.0, dA.ptr, dA.ld, dB.ptr, dB.ld, 1.0, dC.ptr, dC.ld, 0, 9)
        assert e.value.status == dg.ERR_BAD_ARG and "128" in str(e.value)
        with pytest.raises(dg.DgemmError):        # splits = 1 has no partial tiles: not this entry's job
            dg.dgemm_device_static_scratch(512, 512, K, 1.0, dA.ptr, dA.ld, dB.ptr, dB.ld, 1.0, dC.ptr, dC.ld, 0, 1)
        assert dg.lib.dgemm_b200_sync(None) == 0
        assert np.array_equal(dC.to_host(), C0), "a refused call must leave C alone"
        # a call that is served leaves the library's cached device buffers as they were
        dg.dgemm_device_static_scratch(512, 512, K, 1.0, dA.ptr, dA.ld, dB.ptr, dB.ld, 1.0, dC.ptr, dC.ld, 0, 2)
        assert dg.lib.dgemm_b200_sync(None) == 0, dg.last_error()
        assert dg.lib.dgemm_b200_device_bytes() == held
        # ... and the driver's free-memory figure does not move either (kernels are loaded by now; a second CUDA runtime
        # instance sees the same primary context, so its cudaMemGetInfo is the device's)
        free0 = _free_device_bytes()
        if free0 is not None:
            for splits in (2, 4, 0):
                dg.dgemm_device_static_scratch(512, 512, K, 1.0, dA.ptr, dA.ld, dB.ptr, dB.ld, 1.0, dC.ptr, dC.ld, 0, splits)
            assert dg.lib.dgemm_b200_sync(None) == 0, dg.last_error()
            assert _free_device_bytes() == free0
    finally:
        dA.free(); dB.free(); dC.free()


def test_the_entry_is_declared_and_exported(dg):
    """CPU-side: the symbol is in the library and in the header, with the documented size of the array."""
    import os
    assert hasattr(dg.lib, "dgemm_b200_device_static_scratch")
    hdr = open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include", "dgemm_b200.h")).read()
    assert "dgemm_b200_device_static_scratch(" in hdr and "#define DGEMM_B200_STATIC_SCRATCH_BYTES (16u << 20)" in hdr
    assert dg.STATIC_SCRATCH_BYTES == 16 << 20
