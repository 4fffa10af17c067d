"""-m gpu: mydgemm_gpu_v20 with operands that do not fit the HBM budget (DGEMM_B200_HBM_BUDGET): C is cut into blocks (and K
into chunks when one block's operands still do not fit), every block runs through the in-core pipeline
(csrc/dgemm_api.cu: run_host_single / choose_host_blocking).  Small shapes against the full oracle, 8192^3 under a 256 MB
budget against sampled long-double entries."""
import numpy as np
import pytest

import bench
import gpu_util as gu

pytestmark = pytest.mark.gpu
EPS = 2.220446049250313e-16
WS = 148 * 2 * 128 * 128 * 8   # the workspace term of the footprint


@pytest.fixture
def budget(dg, monkeypatch):
    def set_budget(nbytes):
        dg.shutdown()   # drop buffers an earlier test grew beyond the budget
        monkeypatch.setenv("DGEMM_B200_HBM_BUDGET", str(int(nbytes)))
        dg.reload_env()
    yield set_budget
    monkeypatch.delenv("DGEMM_B200_HBM_BUDGET", raising=False)
    dg.reload_env()


@pytest.mark.parametrize("shape,pad,alpha,beta,extra", [
    ((1000, 1003, 997), (0, 0, 0), 2.5, 0.5, 12 << 20),     # C in 2 x 2 blocks, K whole
    ((1000, 1003, 997), (7, 3, 5), -1.0, 1.0, 12 << 20),    # padded leading dimensions: the caller's padding rows survive
    ((700, 2000, 600), (0, 0, 0), 1.0, 0.0, 8 << 20),       # beta == 0: C is not read, NaN in it must not leak
    ((300, 300, 4000), (0, 0, 0), 1.5, 0.5, 4_200_000),     # K cut into chunks: C goes in and out once per chunk
    ((640, 512, 64), (0, 0, 0), 0.0, 0.5, 3 << 20),         # alpha == 0: C = beta * C block by block
])
def test_v20_out_of_core_against_the_oracle(dg, budget, shape, pad, alpha, beta, extra):
    M, N, K = shape
    budget(WS + extra)
    blk = dg.host_blocking(M, N, K, WS + extra, alpha, beta)
    assert (blk["Mb"], blk["Nb"], blk["Kb"]) != (M, N, K), "the budget was meant to force blocking"
    A, B, C0 = gu.make_inputs(M, N, K, pad, c_fill=np.nan if beta == 0.0 else None)
    A_before, B_before = A.copy(order="F"), B.copy(order="F")
    Cg = C0.copy(order="F")
    dg.mydgemm_gpu_v20(M, N, K, alpha, A, A.shape[0], B, B.shape[0], beta, Cg, Cg.shape[0])
    if beta == 0.0:
        C0[:M, :] = 0.0
    gu.check(M, N, K, alpha, A, B, beta, C0, Cg)
    assert np.array_equal(A, A_before) and np.array_equal(B, B_before)
    assert dg.device_bytes() <= WS + extra, "the library holds more device memory than the budget"


def test_v20_out_of_core_8192_under_256_mb(dg, budget):
    M = N = K = 8192
    # operands generated on the device in 512-column pieces (same counter-based generator as the oracle), kept on the host
    host = [np.empty((M, N), order="F") for _ in range(3)]
    piece = dg.malloc(8 * M * 512)
    for h, seed in zip(host, (1, 2, 3)):
        for j0 in range(0, N, 512):
            dg.fill_signed(piece, M, 512, M, seed, 0, j0)
            assert dg.lib.dgemm_b200_d2h(h[:, j0:].ctypes.data, M, piece, M, M, 512, None) == 0
            assert dg.lib.dgemm_b200_sync(None) == 0
    dg.free(piece)
    A, B, Cm = host
    rng = np.random.default_rng(7)
    idx = [(0, 0), (M - 1, N - 1), (2815, 1663), (2816, 1664), (4095, 4096)] + \
          [(int(rng.integers(M)), int(rng.integers(N))) for _ in range(19)]
    before = [float(Cm[i, j]) for i, j in idx]
    budget(256 << 20)
    blk = dg.host_blocking(M, N, K, 256 << 20)
    assert blk["Mb"] < M and blk["Nb"] < N
    dg.mydgemm_gpu_v20(M, N, K, 1.0, A, M, B, K, 1.0, Cm, M)
    assert dg.device_bytes() <= 256 << 20
    after = [float(Cm[i, j]) for i, j in idx]
    assert bench.spot_check(before, after, idx, M, K, 0, 1.0, 1.0) <= 1.0


def test_buffers_of_an_earlier_shape_do_not_sit_beside_the_blocks(dg, monkeypatch):
    """An in-core call of another shape leaves ~100 MB of staging buffers cached in the context.  The blocked run that follows
    sizes its blocks to the budget, so those buffers have to be given back first (run_host_single: release_cached) — without
    a dgemm_b200_shutdown() in between, which is what a real caller would not think of."""
    M0 = 2048
    A, B, C0 = gu.make_inputs(M0, M0, M0)
    dg.mydgemm_gpu_v20(M0, M0, M0, 1.0, A, M0, B, M0, 1.0, C0.copy(order="F"), M0)
    assert dg.device_bytes() > 90 << 20
    M, N, K = 1000, 1003, 997
    limit = WS + (12 << 20)
    monkeypatch.setenv("DGEMM_B200_HBM_BUDGET", str(limit))
    dg.reload_env()
    A, B, C0 = gu.make_inputs(M, N, K, (7, 3, 5))
    Cg = C0.copy(order="F")
    dg.mydgemm_gpu_v20(M, N, K, 2.5, A, A.shape[0], B, B.shape[0], 0.5, Cg, Cg.shape[0])
    gu.check(M, N, K, 2.5, A, B, 0.5, C0, Cg)
    assert dg.device_bytes() <= limit, "buffers of the earlier call are still allocated next to the blocks"
