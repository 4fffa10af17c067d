"""-m gpu: the CUDA path (through the C ABI) against the oracle on identical seeded inputs.

Tolerance (BASELINE.json north_star): |C - C_ref| <= c*K*eps*(|alpha||A||B| + |beta*C0|) elementwise with
c = 2, eps = 2^-52.  Shapes follow BASELINE.json configs 1-3 / SURVEY.md 8(d): squares, the 1000x1003x997
edge shape with dense / padded / rounded leading dimensions, odd M and N (which the reference itself drops,
SURVEY.md F1), alpha in {0, 2.5, -1}, beta in {0, 0.5, 1}, K = 0, tiny M / N, tall-skinny split-K.
"""
import numpy as np
import pytest

import gpu_util as gu

pytestmark = pytest.mark.gpu

TMA, CPA8, CPA16 = 1, 2, 3


@pytest.mark.parametrize("variant", [TMA, CPA8, CPA16])
@pytest.mark.parametrize("size", [128, 256, 384, 512, 1024])
def test_square_all_variants(dg, variant, size):
    gu.run_device(dg, size, size, size, 1.0, 1.0, variant=variant)


@pytest.mark.parametrize("size", [100, 200, 300, 700, 1000, 1500, 2048])
def test_square_reference_harness_sizes(dg, size):
    """test.c:12-14 style sizes with the harness' own (alpha, beta) = (2, 0) and its value distribution."""
    gu.run_device(dg, size, size, size, 2.0, 0.0, dist="reference")
    gu.run_device(dg, size, size, size, 1.0, 1.0, dist="reference")


@pytest.mark.parametrize("pad", [(0, 0, 0), (7, 7, 7), (24, 27, 24), (1, 0, 3)])
@pytest.mark.parametrize("shape", [(1000, 1003, 997), (1001, 1003, 997), (1000, 1002, 997)])
def test_edge_shapes_and_leading_dims(dg, shape, pad):
    M, N, K = shape
    gu.run_device(dg, M, N, K, 2.5, 0.5, pad=pad)


@pytest.mark.parametrize("variant", [TMA, CPA16])
def test_edge_shape_aligned_variants(dg, variant):
    # LD rounded to 1024: TMA-eligible although M, N, K are not tile multiples (TMA zero-fills the edges)
    gu.run_device(dg, 1000, 1003, 997, 2.5, 0.5, pad=(24, 27, 24), variant=variant)
    gu.run_device(dg, 1001, 1003, 999, -1.0, 1.0, pad=(23, 25, 23), variant=variant)


@pytest.mark.parametrize("alpha", [0.0, 2.5, -1.0, 1.0])
@pytest.mark.parametrize("beta", [0.0, 0.5, 1.0])
def test_alpha_beta_grid(dg, alpha, beta):
    gu.run_device(dg, 260, 250, 300, alpha, beta, pad=(4, 2, 6))


def test_beta_zero_ignores_nan_in_c(dg):
    for variant in (TMA, CPA8):
        gu.run_device(dg, 256, 192, 128, 1.5, 0.0, variant=variant, c_fill=np.nan)
    gu.run_device(dg, 256, 192, 0, 1.5, 0.0, c_fill=np.nan)    # K == 0 -> C = 0
    gu.run_device(dg, 256, 192, 64, 0.0, 0.0, c_fill=np.inf)   # alpha == 0 -> C = 0


@pytest.mark.parametrize("K", [0, 1, 3, 15, 16, 17, 383, 385])
def test_k_edges(dg, K):
    gu.run_device(dg, 130, 70, K, 1.25, 0.75)


@pytest.mark.parametrize("mn", [(1, 1), (1, 300), (300, 1), (7, 9), (8, 8), (9, 7), (127, 129), (129, 127)])
def test_tiny_m_n(dg, mn):
    M, N = mn
    gu.run_device(dg, M, N, 200, -0.5, 2.0, pad=(3, 1, 5))


def test_empty_is_noop(dg):
    gu.run_device(dg, 0, 5, 5, 1.0, 0.0)
    gu.run_device(dg, 5, 0, 5, 1.0, 0.0)


@pytest.mark.parametrize("offset", [(1, 0, 0), (0, 1, 0), (0, 0, 1), (1, 1, 1)])
def test_unaligned_base_pointers(dg, offset):
    """8-byte (not 16-byte) aligned operands: no tensor map possible -> device-side cp.async loader."""
    gu.run_device(dg, 300, 260, 200, 1.0, 1.0, offset=offset)


@pytest.mark.parametrize("variant", [TMA, CPA8])
@pytest.mark.parametrize("splits", [2, 3, 9, 16])
def test_split_k_forced(dg, variant, splits):
    gu.run_device(dg, 256, 384, 1000, 2.0, 0.5, variant=variant, splits=splits)


def test_tall_skinny_heuristic_fills_the_gpu(dg):
    """M=N=512, K=16384 has 16 output tiles for 148 SMs: the plan must decompose K (stream-K on the TMA kernel,
    split-K on the cp.async kernel) so that (nearly) every SM works."""
    p = dg.plan(512, 512, 16384)
    assert p["ctas"] > 100 and (p["stream_k_tiles"] > 0 or p["splits"] > 1), p
    gu.run_device(dg, 512, 512, 16384, 1.0, 1.0)
    gu.run_device(dg, 512, 512, 16384, 1.0, 1.0, pad=(1, 1, 1))   # odd LD -> cp.async kernel + split-K


@pytest.mark.parametrize("shape", [(2048, 2048, 2048), (1920, 1664, 512), (1152, 1280, 4096), (640, 512, 8192),
                                   (128, 128, 4096), (1000, 1003, 1000), (18944, 128, 256), (2304, 2048, 128)])
def test_stream_k_shapes(dg, shape):
    """Shapes whose tile count is not a multiple of 148: data-parallel waves + stream-K tail, or all stream-K."""
    M, N, K = shape
    pad = (0, 1, 0) if K % 2 else (0, 0, 0)
    gu.run_device(dg, M, N, K, 1.5, 0.5, pad=pad, variant=TMA)


def test_stream_k_is_deterministic_and_close_to_data_parallel(dg):
    M, N, K = 1920, 1664, 768
    A, B, C0 = gu.make_inputs(M, N, K)
    outs = []
    for splits in (0, 0, 1):   # auto (stream-K) twice, then forced pure data-parallel
        dA, dB, dC = gu.DevMat(dg, A), gu.DevMat(dg, B), gu.DevMat(dg, C0)
        dg.dgemm_device(M, N, K, 1.0, dA.ptr, dA.ld, dB.ptr, dB.ld, 1.0, dC.ptr, dC.ld, 0, TMA, splits)
        dg.lib.dgemm_b200_sync(None)
        outs.append(dC.to_host())
        dA.free(); dB.free(); dC.free()
    assert np.array_equal(outs[0], outs[1])                       # run-to-run bit-identical
    gu.check(M, N, K, 1.0, A, B, 1.0, C0, outs[0], Cref=None)     # and inside the bound
    gu.check(M, N, K, 1.0, A, B, 1.0, C0, outs[2], Cref=None)


def test_split_k_is_deterministic(dg):
    A, B, C0 = gu.make_inputs(256, 256, 4096)
    outs = []
    for _ in range(3):
        dA, dB, dC = gu.DevMat(dg, A), gu.DevMat(dg, B), gu.DevMat(dg, C0)
        dg.dgemm_device(256, 256, 4096, 1.0, dA.ptr, dA.ld, dB.ptr, dB.ld, 1.0, dC.ptr, dC.ld, 0, 0, 8)
        dg.lib.dgemm_b200_sync(None)
        outs.append(dC.to_host())
        dA.free(); dB.free(); dC.free()
    assert np.array_equal(outs[0], outs[1]) and np.array_equal(outs[0], outs[2])


def test_variants_agree_bitwise(dg):
    """All loaders feed the same DMMA schedule -> identical bits (only the operand staging differs)."""
    A, B, C0 = gu.make_inputs(520, 390, 777, pad=(0, 1, 0))
    A2, B2, C2 = gu.make_inputs(520, 390, 777, pad=(0, 1, 0))
    outs = []
    for variant in (TMA, CPA8, CPA16):
        dA, dB, dC = gu.DevMat(dg, A), gu.DevMat(dg, B), gu.DevMat(dg, C0)
        dg.dgemm_device(520, 390, 777, 1.0, dA.ptr, dA.ld, dB.ptr, dB.ld, 1.0, dC.ptr, dC.ld, 0, variant, 1)  # pure DP
        dg.lib.dgemm_b200_sync(None)
        outs.append(dC.to_host())
        dA.free(); dB.free(); dC.free()
    assert np.array_equal(outs[0], outs[1]) and np.array_equal(outs[0], outs[2])


@pytest.mark.parametrize("case", [((2200, 2304, 2100), (1, 0, 0), (0, 0, 0)),     # odd LDA only
                                  ((2200, 2304, 2101), (0, 0, 0), (0, 0, 0)),     # odd LDB (= K) only
                                  ((2049, 2051, 2305), (0, 0, 2), (0, 0, 0)),     # odd M -> odd LDA and LDC, odd LDB
                                  ((2304, 2200, 2100), (0, 0, 0), (1, 1, 0))])    # 8-byte-aligned bases
def test_large_unaligned_operands(dg, case, monkeypatch):
    """~2e10 flops and TMA-ineligible: read in place by the parity-split TMA kernel (default), or - what calls with two
    misaligned halves get from 4e11 flops up, forced here through the threshold - D2D realign + plain TMA kernel.
    Same bound, A/B untouched, pad rows untouched."""
    (M, N, K), pad, offset = case
    gu.run_device(dg, M, N, K, 1.5, 0.5, pad=pad, offset=offset)
    monkeypatch.setenv("DGEMM_B200_SPLIT_MAX_FLOPS2", "1e9")
    dg.reload_env()
    gu.run_device(dg, M, N, K, 1.5, 0.5, pad=pad, offset=offset)


@pytest.mark.parametrize("variant", [4, 5])   # ladder rungs: naive, DFMA register tile
@pytest.mark.parametrize("case", [((256, 256, 256), (0, 0, 0), 1.0, 1.0), ((1000, 1003, 997), (0, 0, 0), 2.5, 0.5),
                                  ((130, 70, 17), (3, 1, 5), -1.0, 0.0), ((1, 9, 200), (0, 0, 0), 1.0, 1.0)])
def test_ladder_rungs(dg, variant, case):
    (M, N, K), pad, alpha, beta = case
    gu.run_device(dg, M, N, K, alpha, beta, pad=pad, variant=variant, c_fill=np.nan if beta == 0.0 else None)


@pytest.mark.parametrize("case", [((256, 256, 256), (0, 0, 0), 1.0, 1.0),        # one pair
                                  ((1024, 1024, 1024), (0, 0, 0), 1.0, 1.0),
                                  ((1000, 1003, 998), (0, 0, 0), 2.5, 0.5),       # ragged edges, 8 tile columns
                                  ((640, 384, 4096), (0, 0, 0), -1.0, 0.0),       # ODD number of tile columns: a half-empty pair
                                  ((130, 70, 18), (2, 2, 4), 1.5, 1.0),           # a single tile: the pair's second CTA has no output
                                  ((2304, 2560, 2048), (0, 0, 0), 1.0, 1.0)])     # more pairs than clusters: the pipeline crosses tiles
def test_cluster_multicast_rung(dg, case):
    """Ladder rung 6 (csrc/dgemm_kernels_cluster.cu): 2-CTA clusters, the shared A panel by TMA multicast."""
    (M, N, K), pad, alpha, beta = case
    gu.run_device(dg, M, N, K, alpha, beta, pad=pad, variant=dg.VARIANT_CLUSTER, c_fill=np.nan if beta == 0.0 else None)


def test_cluster_rung_same_bits_as_the_shipped_tile_loop(dg):
    """Same 128x128 tile, same k order, same DMMA schedule: only how A reaches shared memory differs."""
    M, N, K = 1100, 900, 777
    A, B, C0 = gu.make_inputs(M, N, K, pad=(0, 1, 0))
    outs = []
    for variant in (TMA, dg.VARIANT_CLUSTER):
        dA, dB, dC = gu.DevMat(dg, A), gu.DevMat(dg, B), gu.DevMat(dg, C0)
        dg.dgemm_device(M, N, K, 1.0, dA.ptr, dA.ld, dB.ptr, dB.ld, 1.0, dC.ptr, dC.ld, 0, variant, 1)
        dg.lib.dgemm_b200_sync(None)
        outs.append(dC.to_host())
        dA.free(); dB.free(); dC.free()
    assert np.array_equal(outs[0], outs[1])


def test_cluster_rung_rejects_operands_tma_cannot_address(dg):
    A, B, C0 = gu.make_inputs(256, 256, 255)   # odd LDB
    dA, dB, dC = gu.DevMat(dg, A), gu.DevMat(dg, B), gu.DevMat(dg, C0)
    with pytest.raises(dg.DgemmError) as e:
        dg.dgemm_device(256, 256, 255, 1.0, dA.ptr, dA.ld, dB.ptr, dB.ld, 1.0, dC.ptr, dC.ld, 0, dg.VARIANT_CLUSTER, 1)
    assert e.value.status == dg.ERR_BAD_ARG
    dA.free(); dB.free(); dC.free()


@pytest.mark.parametrize("no_wait", [None, "1"])
@pytest.mark.parametrize("splits", [0, 1])
def test_waiting_kernels_survive_a_busy_device(dg, monkeypatch, no_wait, splits):
    """Co-residency guarantee of the configurations whose CTAs wait for each other (k-lockstep of the data-parallel waves,
    parallel stream-K fix-up): another stream holds most of every SM's shared memory for 300 ms while a lockstep-size
    DGEMM is issued.  With the cooperative-launch attribute the grid starts only once all its CTAs fit (after the hog);
    with DGEMM_B200_NO_WAIT=1 the waits are dropped and the kernel shares the device.  Either way: done, and right."""
    import time
    if no_wait:
        monkeypatch.setenv("DGEMM_B200_NO_WAIT", no_wait)
    dg.reload_env()
    M, N, K = 3072, 3072, 4096
    A, B, C0 = gu.make_inputs(M, N, K)
    dA, dB, dC = gu.DevMat(dg, A), gu.DevMat(dg, B), gu.DevMat(dg, C0)
    try:
        assert dg.lib.dgemm_b200_sync(None) == 0
        # 148 CTAs x 150 KB: at most one per SM, and no SM is left with room for a 199 KB DGEMM CTA while they run
        assert dg.lib.dgemm_b200_debug_occupy_sms(148, 150 * 1024, 300.0) == 0, dg.last_error()
        t0 = time.perf_counter()
        dg.dgemm_device(M, N, K, 1.0, dA.ptr, dA.ld, dB.ptr, dB.ld, 1.0, dC.ptr, dC.ld, 0, 0, splits)
        assert dg.lib.dgemm_b200_sync(None) == 0, dg.last_error()
        dt = time.perf_counter() - t0
        assert dt < 5.0, f"{dt:.1f} s: the DGEMM sat behind the hog for far longer than the hog ran"
        gu.check(M, N, K, 1.0, A, B, 1.0, C0, dC.to_host())
        # and half the SMs held for the whole run of the DGEMM: the grid cannot be co-resident until they are released
        assert dg.lib.dgemm_b200_debug_occupy_sms(74, 150 * 1024, 300.0) == 0, dg.last_error()
        dC2 = gu.DevMat(dg, C0)
        dg.dgemm_device(M, N, K, 1.0, dA.ptr, dA.ld, dB.ptr, dB.ld, 1.0, dC2.ptr, dC2.ld, 0, 0, splits)
        assert dg.lib.dgemm_b200_sync(None) == 0, dg.last_error()
        gu.check(M, N, K, 1.0, A, B, 1.0, C0, dC2.to_host())
        dC2.free()
    finally:
        dA.free(); dB.free(); dC.free()
        if no_wait:
            monkeypatch.delenv("DGEMM_B200_NO_WAIT")
        dg.reload_env()


@pytest.mark.parametrize("case", [((256, 256, 256), (0, 0, 0), 1.0, 1.0), ((512, 512, 512), (0, 0, 0), 2.5, 0.5),
                                  ((1000, 1003, 997), (0, 0, 0), -1.0, 1.0),      # odd LDB: parity-split kernel of the build
                                  ((1000, 1003, 997), (7, 7, 7), 1.5, 0.0),       # odd LDA
                                  ((1024, 1024, 1024), (0, 0, 0), 1.0, 1.0), ((300, 700, 64), (2, 0, 0), 1.0, 1.0)])
def test_small_tile_build_at_two_ctas_per_sm(dg, monkeypatch, case):
    """DGEMM_B200_SMALL_CTAS=2: the 64 x 64 kernel compiled for two resident CTAs per SM (csrc/dgemm_kernels_small2.cu),
    grid and stream-K ranges sized for 296 CTA slots.  Same bound; bit-identical to itself run to run."""
    (M, N, K), pad, alpha, beta = case
    monkeypatch.setenv("DGEMM_B200_SMALL_CTAS", "2")
    dg.reload_env()
    try:
        A, B, C0 = gu.make_inputs(M, N, K, pad, c_fill=np.nan if beta == 0.0 else None)
        outs = []
        for _ in range(2):
            dA, dB, dC = gu.DevMat(dg, A), gu.DevMat(dg, B), gu.DevMat(dg, C0)
            dg.dgemm_device(M, N, K, alpha, dA.ptr, dA.ld, dB.ptr, dB.ld, beta, dC.ptr, dC.ld)
            assert dg.lib.dgemm_b200_sync(None) == 0, dg.last_error()
            outs.append(dC.to_host())
            dA.free(); dB.free(); dC.free()
        if beta == 0.0:
            C0[:M, :] = 0.0
        gu.check(M, N, K, alpha, A, B, beta, C0, outs[0])
        assert np.array_equal(outs[0], outs[1])
    finally:
        monkeypatch.delenv("DGEMM_B200_SMALL_CTAS")
        dg.reload_env()


@pytest.mark.parametrize("case", [((128, 64, 16), (0, 0, 0), 1.0, 1.0), ((256, 256, 256), (0, 0, 0), 1.0, 1.0),
                                  ((512, 512, 512), (0, 0, 0), 2.5, 0.5),          # all stream-K, ranges shorter than a tile
                                  ((1000, 1003, 997), (0, 0, 0), -1.0, 1.0),      # odd LDB: parity-split kernel of the build
                                  ((1000, 1003, 997), (7, 7, 7), 1.5, 0.0),       # odd LDA, beta = 0 over NaN
                                  ((1001, 999, 1003), (1, 1, 1), 1.0, 1.0),       # both operands split, odd LDC
                                  ((1024, 1024, 1024), (0, 0, 0), 1.0, 1.0), ((1536, 1536, 1536), (0, 0, 0), 1.0, 1.0),
                                  ((2048, 2048, 512), (0, 0, 0), 1.0, 1.0),       # data-parallel wave + stream-K tail
                                  ((300, 700, 64), (2, 0, 0), 1.0, 1.0), ((129, 65, 4096), (0, 0, 0), 1.0, 0.5),
                                  ((7, 9, 33), (0, 0, 0), 1.0, 1.0)])
def test_mid_tile_build(dg, monkeypatch, case):
    """DGEMM_B200_TILE=mid: the TMA kernel compiled for 128 x 64 CTA tiles with 32 x 32 warp tiles
    (csrc/dgemm_kernels_mid.cu; 24 KB stages, stream-K partials in 16 KB chunks).  Same bound as every other path;
    bit-identical to itself run to run."""
    (M, N, K), pad, alpha, beta = case
    monkeypatch.setenv("DGEMM_B200_TILE", "mid")
    dg.reload_env()
    try:
        A, B, C0 = gu.make_inputs(M, N, K, pad, c_fill=np.nan if beta == 0.0 else None)
        outs = []
        for _ in range(2):
            dA, dB, dC = gu.DevMat(dg, A), gu.DevMat(dg, B), gu.DevMat(dg, C0)
            dg.dgemm_device(M, N, K, alpha, dA.ptr, dA.ld, dB.ptr, dB.ld, beta, dC.ptr, dC.ld)
            assert dg.lib.dgemm_b200_sync(None) == 0, dg.last_error()
            outs.append(dC.to_host())
            dA.free(); dB.free(); dC.free()
        if beta == 0.0:
            C0[:M, :] = 0.0
        gu.check(M, N, K, alpha, A, B, beta, C0, outs[0])
        assert np.array_equal(outs[0], outs[1])
    finally:
        monkeypatch.delenv("DGEMM_B200_TILE")
        dg.reload_env()


@pytest.mark.parametrize("sk", [False, True])
@pytest.mark.parametrize("case", [((64, 32, 16), (0, 0, 0), 1.0, 1.0), ((256, 256, 256), (0, 0, 0), 1.0, 1.0),
                                  ((512, 512, 512), (0, 0, 0), 2.5, 0.5),          # 128 whole tiles, one wave
                                  ((512, 512, 517), (0, 0, 0), 1.0, 1.0),          # odd number of k-tiles: a half-filled ring slot
                                  ((500, 501, 499), (0, 0, 0), -1.0, 1.0),         # odd LDB: parity-split kernel of the build
                                  ((500, 501, 499), (7, 7, 7), 1.5, 0.0),          # odd LDA, beta = 0 over NaN
                                  ((501, 499, 503), (1, 1, 1), 1.0, 1.0),          # both operands split, odd LDC
                                  ((640, 640, 640), (0, 0, 0), 1.0, 1.0),          # 200 tiles: a wave + stream-K tail
                                  ((1000, 1003, 997), (0, 0, 0), 1.0, 1.0),
                                  ((300, 700, 64), (2, 0, 0), 1.0, 1.0), ((65, 33, 4096), (0, 0, 0), 1.0, 0.5),
                                  ((64, 64, 8192), (0, 0, 0), 1.0, 1.0),           # 2 tiles: with TINY_SK many ranges per tile
                                  ((7, 9, 33), (0, 0, 0), 1.0, 1.0)])
def test_tiny_tile_build(dg, monkeypatch, case, sk):
    """DGEMM_B200_TILE=tiny: the TMA kernel compiled for 64 x 32 CTA tiles with 16 x 16 warp tiles and one k-tile per ring
    slot (csrc/dgemm_kernels_tiny.cu).  sk: DGEMM_B200_TINY_SK=1 lets a partial wave of these tiles be cut into stream-K
    ranges (the build's default is one data-parallel wave).  Same bound as every other path; bit-identical run to run."""
    (M, N, K), pad, alpha, beta = case
    monkeypatch.setenv("DGEMM_B200_TILE", "tiny")
    if sk:
        monkeypatch.setenv("DGEMM_B200_TINY_SK", "1")
    dg.reload_env()
    try:
        A, B, C0 = gu.make_inputs(M, N, K, pad, c_fill=np.nan if beta == 0.0 else None)
        outs = []
        for _ in range(2):
            dA, dB, dC = gu.DevMat(dg, A), gu.DevMat(dg, B), gu.DevMat(dg, C0)
            dg.dgemm_device(M, N, K, alpha, dA.ptr, dA.ld, dB.ptr, dB.ld, beta, dC.ptr, dC.ld)
            assert dg.lib.dgemm_b200_sync(None) == 0, dg.last_error()
            outs.append(dC.to_host())
            dA.free(); dB.free(); dC.free()
        if beta == 0.0:
            C0[:M, :] = 0.0
        gu.check(M, N, K, alpha, A, B, beta, C0, outs[0])
        assert np.array_equal(outs[0], outs[1])
    finally:
        monkeypatch.delenv("DGEMM_B200_TILE")
        if sk:
            monkeypatch.delenv("DGEMM_B200_TINY_SK")
        dg.reload_env()


def test_auto_picks_the_tiny_tile_at_512(dg):
    """512^3: 128 tiles of 64 x 32 in one data-parallel wave (no stream-K tiles, 128 CTAs), and the result is the oracle's."""
    p = dg.plan(512, 512, 512)
    assert p["ctas"] == 128 and p["stream_k_tiles"] == 0, p
    assert gu.run_device(dg, 512, 512, 512, 1.0, 1.0) <= gu.BOUND_C


def test_k_lockstep_changes_no_bits(dg, monkeypatch):
    """Large enough for the k-lockstep of the data-parallel waves (2 full waves, operands > 96 MB, K >= 4096): the
    producers' pacing must not change a single bit, twice in a row (the counters are re-zeroed per launch), and the
    result stays inside the bound."""
    M, N, K = 3072, 3072, 4096
    A, B, C0 = gu.make_inputs(M, N, K)
    dA, dB = gu.DevMat(dg, A), gu.DevMat(dg, B)
    outs = []
    for no_lockstep in (None, None, "1"):
        if no_lockstep:
            monkeypatch.setenv("DGEMM_B200_NO_LOCKSTEP", no_lockstep)
            dg.reload_env()
        else:
            monkeypatch.delenv("DGEMM_B200_NO_LOCKSTEP", raising=False)
            dg.reload_env()
        dC = gu.DevMat(dg, C0)
        dg.dgemm_device(M, N, K, 1.0, dA.ptr, dA.ld, dB.ptr, dB.ld, 1.0, dC.ptr, dC.ld, 0, TMA, 0)
        assert dg.lib.dgemm_b200_sync(None) == 0, dg.last_error()
        outs.append(dC.to_host())
        dC.free()
    dA.free(); dB.free()
    assert np.array_equal(outs[0], outs[1]) and np.array_equal(outs[0], outs[2])
    gu.check(M, N, K, 1.0, A, B, 1.0, C0, outs[0])


def _random_cases(n, seed):
    rng = np.random.default_rng(seed)
    cases = []
    for _ in range(n):
        M, N, K = (int(rng.integers(1, 700)) for _ in range(3))
        pad = tuple(int(rng.integers(0, 4)) for _ in range(3))
        alpha = float(rng.choice([1.0, -1.0, 2.5, 0.5]))
        beta = float(rng.choice([0.0, 1.0, 0.5, -2.0]))
        cases.append((M, N, K, pad, alpha, beta))
    return cases


@pytest.mark.parametrize("case", _random_cases(24, 2024), ids=lambda c: "x".join(str(v) for v in c[:3]))
def test_random_shapes_device_and_host_entries(dg, case):
    """Seeded random shapes, paddings and scalars through BOTH entries (device pointers: TMA / realigned / cp.async
    loaders as the alignment dictates; host pointers: planner, staging, deferred C), each against the oracle."""
    M, N, K, pad, alpha, beta = case
    gu.run_device(dg, M, N, K, alpha, beta, pad=pad, c_fill=np.nan if beta == 0.0 else None)
    A, B, C0 = gu.make_inputs(M, N, K, pad)
    Cg = C0.copy(order="F")
    dg.mydgemm_gpu_v20(M, N, K, alpha, A, A.shape[0], B, B.shape[0], beta, Cg, Cg.shape[0])
    gu.check(M, N, K, alpha, A, B, beta, C0, Cg)


# ---- parity-split tensor maps: odd leading dimensions / 8-byte-aligned bases read in place by the TMA kernel ----------
_SPLIT_CASES = [
    ((1000, 1003, 997), (0, 0, 0), (0, 0, 0)),     # BASELINE configs[2]: LDB = 997 odd -> B split by column parity
    ((1000, 1003, 997), (7, 7, 7), (0, 0, 0)),     # LD = dim + 7: LDA = 1007 odd -> A split by k parity
    ((1001, 1003, 999), (0, 0, 0), (0, 0, 0)),     # odd M: LDA, LDC odd; odd K: LDB odd -> both split
    ((640, 512, 8192), (1, 1, 0), (0, 0, 0)),      # all stream-K (parallel fix-up) with both operands split
    ((2200, 2304, 2100), (1, 0, 0), (0, 0, 0)),    # data-parallel wave + stream-K tail, A split
    ((2200, 2304, 2101), (0, 0, 0), (0, 0, 0)),    # the same, B split
    ((2304, 2200, 2100), (0, 0, 0), (1, 1, 0)),    # even LD, bases at 8 mod 16: both parity maps shifted by one row
    ((777, 555, 333), (0, 0, 0), (1, 0, 1)),       # odd LD AND a shifted base for A; C on an 8-byte base too
    ((300, 2, 2), (1, 1, 0), (0, 0, 0)),           # smallest K and N the split maps can take
    ((129, 3, 5), (0, 0, 0), (0, 1, 0)),
    ((7, 130, 17), (0, 0, 0), (0, 0, 0)),
    ((128, 128, 16), (1, 1, 1), (1, 1, 1)),
    ((300, 200, 1), (1, 0, 0), (0, 0, 0)),         # K = 1: no odd-k parity to map -> cp.async kernel
    ((300, 1, 65), (0, 0, 0), (0, 0, 0)),          # N = 1 with LDB = 65 odd: no odd-column parity -> cp.async kernel
]


@pytest.mark.gpu
@pytest.mark.parametrize("case", _SPLIT_CASES, ids=lambda c: "x".join(str(v) for v in c[0]) + "-" + "".join(map(str, c[1] + c[2])))
@pytest.mark.parametrize("ab", [(1.5, 0.5), (-1.0, 0.0)])
def test_parity_split_tensor_maps(dg, case, ab):
    """Operands one tensor map cannot describe run on the TMA kernel through even/odd-column tensor maps (no copy, no
    cp.async): same bound against the oracle, padding rows untouched, NaN-filled C ignored when beta = 0."""
    (M, N, K), pad, offset = case
    gu.run_device(dg, M, N, K, ab[0], ab[1], pad=pad, offset=offset, c_fill=np.nan if ab[1] == 0.0 else None,
                  nan_pad=True)
    if K >= 2 and N >= 2:   # (a parity map needs at least one column; below that VARIANT_AUTO uses the cp.async kernel)
        gu.run_device(dg, M, N, K, ab[0], ab[1], pad=pad, offset=offset, variant=TMA,
                      c_fill=np.nan if ab[1] == 0.0 else None, nan_pad=True)


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["split", "realign", "cpasync"])
def test_odd_ld_modes_agree_with_the_oracle(dg, mode, monkeypatch):
    """DGEMM_B200_ODD_LD selects how TMA-ineligible operands are handled; every mode meets the bound."""
    monkeypatch.setenv("DGEMM_B200_ODD_LD", mode)
    dg.reload_env()
    gu.run_device(dg, 1000, 1003, 997, 2.5, 0.5)
    gu.run_device(dg, 1001, 1003, 999, 1.0, 1.0, offset=(1, 0, 0))


@pytest.mark.gpu
def test_column_parity_split_of_b_is_bitwise_identical(dg):
    """Splitting B by column parity only permutes which shared-memory row holds which column: same DMMA schedule, same
    bits as the single-map kernel on the same values with an even LDB."""
    M, N, K = 520, 390, 777
    A, B, C0 = gu.make_inputs(M, N, K, pad=(0, 1, 0))          # LDB = 778 (even): one tensor map
    B_odd = np.asfortranarray(B[:K, :])                        # same values, LDB = 777 (odd): parity maps
    outs = []
    for Bm in (B, B_odd):
        dA, dB, dC = gu.DevMat(dg, A), gu.DevMat(dg, Bm), gu.DevMat(dg, C0)
        dg.dgemm_device(M, N, K, 1.0, dA.ptr, dA.ld, dB.ptr, dB.ld, 1.0, dC.ptr, dC.ld, 0, TMA, 1)
        assert dg.lib.dgemm_b200_sync(None) == 0, dg.last_error()
        outs.append(dC.to_host())
        dA.free(); dB.free(); dC.free()
    assert np.array_equal(outs[0], outs[1])


# ---- 64 x 64-tile build of the TMA kernel (small problems) against the 128 x 128 one ---------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("shape", [(100, 100, 100), (64, 64, 16), (65, 63, 17), (200, 300, 250), (256, 256, 256),
                                   (500, 501, 498), (512, 512, 512), (640, 384, 2048), (1000, 1002, 996),
                                   (1024, 1024, 1024), (128, 4096, 512), (4096, 128, 512), (31, 7000, 129)])
@pytest.mark.parametrize("ab", [(1.0, 1.0), (2.5, 0.0)])
def test_small_tile_and_big_tile_kernels(dg, shape, ab, monkeypatch):
    """VARIANT_AUTO runs small TMA-eligible problems on the 64 x 64-tile build (all stream-K modes: data-parallel, serial
    and parallel fix-up); DGEMM_B200_NO_SMALL_TILE=1 keeps them on the 128 x 128 tile.  Both against the oracle; padded
    C (odd LDC -> scalar epilogue) included."""
    M, N, K = shape
    for pad in ((0, 0, 0), (2, 2, 1)):
        gu.run_device(dg, M, N, K, ab[0], ab[1], pad=pad, c_fill=np.nan if ab[1] == 0.0 else None, nan_pad=True)
    monkeypatch.setenv("DGEMM_B200_NO_SMALL_TILE", "1")
    dg.reload_env()
    gu.run_device(dg, M, N, K, ab[0], ab[1], c_fill=np.nan if ab[1] == 0.0 else None)


@pytest.mark.gpu
def test_small_tile_is_deterministic(dg):
    M, N, K = 384, 320, 1500
    A, B, C0 = gu.make_inputs(M, N, K)
    outs = []
    for _ in range(3):
        dA, dB, dC = gu.DevMat(dg, A), gu.DevMat(dg, B), gu.DevMat(dg, C0)
        dg.dgemm_device(M, N, K, 1.0, dA.ptr, dA.ld, dB.ptr, dB.ld, 1.0, dC.ptr, dC.ld)
        assert dg.lib.dgemm_b200_sync(None) == 0, dg.last_error()
        outs.append(dC.to_host())
        dA.free(); dB.free(); dC.free()
    assert np.array_equal(outs[0], outs[1]) and np.array_equal(outs[0], outs[2])
