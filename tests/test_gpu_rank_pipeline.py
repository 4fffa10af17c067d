"""-m gpu, ONE B200 is enough: the multi-GPU rank pipeline (dgemm_b200_rank_pipeline, csrc/dgemm_rank.cu -- the code
mydgemm_gpu_v21 runs per GPU and bench.py times per rank) with its peers EMULATED on the same device.

Rank r's pass is run for every r of a world of G, one after the other.  The peers' exchange buffers are stand-ins that
already hold the whole of A and a sequence word that says "everything has landed", so every pull finds its data without
any other rank running (two kernels that wait for each other must never share a GPU).  Rank r's own exchange buffer
starts as NaN: every foreign column it multiplies must really have been pulled, every own column really uploaded.
Each rank's C panel is compared with the oracle entry by entry, and the panels are stitched into the full C."""
import ctypes

import numpy as np
import pytest

import gpu_util as gu
import oracle_lib as ol
from dgemm_b200.sharding import partition

pytestmark = pytest.mark.gpu


def _exchange(dg, M, K, fill_from=None, flag=0):
    """An exchange buffer for (M, K): A region NaN (or a copy of fill_from), sequence word = flag."""
    ldaD = (M + 1) & ~1
    nbytes = dg.lib.dgemm_b200_exchange_bytes(M, K)
    flag_off = nbytes - 256
    x = dg.exchange_alloc(M, K)
    host = np.full((ldaD, max(K, 1)), np.nan, order="F")
    if fill_from is not None:
        host[:M, :K] = fill_from[:M, :K]
    assert dg.lib.dgemm_b200_h2d(x, ldaD, host.ctypes.data, ldaD, ldaD, max(K, 1), None) == 0
    word = np.array([flag, 0], dtype=np.uint32)
    assert dg.lib.dgemm_b200_h2d(x + flag_off, 1, word.ctypes.data, 1, 1, 1, None) == 0   # 8 bytes: flag + padding
    assert dg.lib.dgemm_b200_sync(None) == 0
    return x


def run_world(dg, M, N, K, G, alpha, beta, pad=(0, 0, 0), timeline=False):
    A, B, C0 = gu.make_inputs(M, N, K, pad)
    nb, kb = partition(N, G, 128), partition(K, G, 16)
    Cg = C0.copy(order="F")
    full = _exchange(dg, M, K, fill_from=A, flag=1 << 20)    # what every peer's buffer looks like once it has uploaded
    worst, tls = 0.0, []
    try:
        for r in range(G):
            own = _exchange(dg, M, K)
            try:
                xbuf = [own if g == r else full for g in range(G)]
                n0, ncols = nb[r], nb[r + 1] - nb[r]
                A_slab = A[:, kb[r]:kb[r + 1]]          # column slices keep their leading dimension
                tl = dg.rank_pipeline(M, ncols, K, alpha, A_slab.ctypes.data if A_slab.size else 0, A.shape[0],
                                      B[:, n0:].ctypes.data if ncols else 0, B.shape[0], beta,
                                      Cg[:, n0:].ctypes.data if ncols else 0, Cg.shape[0], r, G, kb, xbuf, 0,
                                      30000 if timeline else 0)
                tls.append(tl)
            finally:
                dg.free(own)
    finally:
        dg.free(full)
    worst = gu.check(M, N, K, alpha, A, B, beta, C0, Cg)
    return worst, tls


@pytest.mark.parametrize("G", [1, 2, 3, 4, 8])
@pytest.mark.parametrize("shape,pad,alpha,beta", [((1024, 2048, 1024), (0, 0, 0), 1.0, 1.0),
                                                  ((1000, 1003, 997), (3, 1, 5), 2.5, 0.5),
                                                  ((512, 900, 333), (0, 0, 0), -1.0, 0.0),
                                                  ((2048, 8192, 2048), (0, 0, 0), 1.0, 1.0)])
def test_rank_pipeline_every_rank_of_a_world(dg, G, shape, pad, alpha, beta):
    M, N, K = shape
    if partition(K, G, 16)[G] - partition(K, G, 16)[G - 1] <= 0:
        pytest.skip("K too small for this many ranks (v21 falls back to its simple path)")
    run_world(dg, M, N, K, G, alpha, beta, pad)


def test_rank_pipeline_large_slab_lead_in_and_panel_major_phase(dg):
    """Big enough for several upload pieces per slab, several sub-panels and a panel-major second phase."""
    worst, tls = run_world(dg, 4096, 4096 * 3, 8192, 4, 1.0, 1.0, timeline=True)
    assert worst <= gu.BOUND_C
    for tl in tls:
        kinds = {k for k, _i, _ms in tl}
        assert kinds == {0, 1, 2, 3}
        starts = [ms for k, _i, ms in tl if k == 1]
        assert starts == sorted(starts)


def test_rank_pipeline_degenerate_calls(dg):
    # alpha == 0 and K == 0: C = beta * C on every rank's panel (kernel18 order, reference kernel18.h:644-645)
    for (M, N, K, alpha, beta) in [(300, 700, 64, 0.0, 0.5), (300, 700, 0, 1.0, 0.5), (256, 512, 64, 0.0, 0.0)]:
        A, B, C0 = gu.make_inputs(M, N, K)
        nb, kb = partition(N, 2, 128), partition(K, 2, 16)
        Cg = C0.copy(order="F")
        x = [_exchange(dg, M, K), _exchange(dg, M, K)]
        try:
            for r in range(2):
                n0, ncols = nb[r], nb[r + 1] - nb[r]
                dg.rank_pipeline(M, ncols, K, alpha, A.ctypes.data, A.shape[0], B[:, n0:].ctypes.data, B.shape[0], beta,
                                 Cg[:, n0:].ctypes.data, Cg.shape[0], r, 2, kb, x, 0)
        finally:
            for p in x:
                dg.free(p)
        gu.check(M, N, K, alpha, A, B, beta, C0, Cg)


def test_rank_pipeline_bad_arguments(dg):
    x = _exchange(dg, 64, 32)
    try:
        kb = partition(32, 4, 16)   # ranks 2 and 3 own no K-slab
        A, B, C0 = gu.make_inputs(64, 512, 32)
        with pytest.raises(dg.DgemmError) as e:
            dg.rank_pipeline(64, 128, 32, 1.0, A.ctypes.data, 64, B.ctypes.data, 32, 1.0, C0.ctypes.data, 64, 3, 4, kb,
                             [x, x, x, x], 0)
        assert e.value.status == dg.ERR_BAD_ARG
        with pytest.raises(dg.DgemmError):
            dg.rank_pipeline(64, 128, 32, 1.0, A.ctypes.data, 8, B.ctypes.data, 32, 1.0, C0.ctypes.data, 64, 0, 1, [0, 32],
                             [x], 0)   # LDA < M
    finally:
        dg.free(x)


def test_peer_slab_that_never_arrives_is_an_error_not_a_hang(dg, monkeypatch):
    """The wait for an owner's sequence word is bounded (DGEMM_B200_FLAG_TIMEOUT_MS shortens the 20 s for this test)."""
    monkeypatch.setenv("DGEMM_B200_FLAG_TIMEOUT_MS", "300")
    dg.reload_env()
    M, N, K = 256, 512, 256
    A, B, C0 = gu.make_inputs(M, N, K)
    own, dead = _exchange(dg, M, K), _exchange(dg, M, K)   # the "peer" never uploads: its word stays 0
    try:
        kb = partition(K, 2, 16)
        Cg = C0.copy(order="F")   # (kept alive: the library writes the panel back before it reports the failure)
        with pytest.raises(dg.DgemmError) as e:
            dg.rank_pipeline(M, 256, K, 1.0, A.ctypes.data, M, B.ctypes.data, K, 1.0, Cg.ctypes.data, M, 0, 2, kb, [own, dead], 0)
        assert "never arrived" in str(e.value)
    finally:
        dg.free(own); dg.free(dead)
        monkeypatch.delenv("DGEMM_B200_FLAG_TIMEOUT_MS")
        dg.reload_env()
