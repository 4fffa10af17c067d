"""CPU: bench.py's host-side pieces — thread choice for kernel19, the --impl reference line."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def test_pick_threads_gives_even_bands():
    for M in (2048, 4096, 8192, 1000, 6):
        for cores in (1, 3, 7, 8, 16, 96):
            n = bench.pick_threads(M, cores)
            assert 1 <= n <= cores and (M % (2 * n) == 0 or n == 1)
    assert bench.pick_threads(8192, 16) == 16 and bench.pick_threads(8192, 7) == 4


def test_reference_arm_prints_contract_line():
    env = dict(os.environ, BENCH_REF_SAMPLE="512")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "2",
                          "--warmup", "1"], capture_output=True, text=True, env=env, check=True).stdout
    line = json.loads(out.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["unit"] == "TFLOP/s" and line["value"] > 0
    assert line["e2e"]["h2d_bytes_per_step"] == 0 and line["e2e"]["d2h_bytes_per_step"] == 0
    assert line["cpu_baseline"]["kind"] in ("reference", "port") and line["cpu_baseline"]["cores"] >= 1
    assert line["steps"] == 2 and line["higher_is_better"] is True


def test_reference_arm_other_ranks_exit_quietly():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", BENCH_REF_SAMPLE="512")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2"],
                       capture_output=True, text=True, env=env)
    assert r.returncode == 0 and r.stdout.strip() == ""


def test_bench_generator_matches_oracle_and_spot_check_catches_errors():
    import numpy as np
    import oracle_lib as ol
    x = ol.fill_signed(7, 5, 7, 2, row0=3, col0=65530)
    r, c = np.meshgrid(np.arange(3, 10), np.arange(65530, 65535), indexing="ij")
    assert np.array_equal(x, bench.gen_signed(2, r.ravel(), c.ravel()).reshape(7, 5))
    # spot_check: exact result passes, a 1e-9 relative error fails
    M, N, K = 40, 24, 300
    A = ol.fill_signed(M, K, M, 1); B = ol.fill_signed(K, N, K, 2, col0=128); C0 = ol.fill_signed(M, N, M, 3, col0=128)
    C = C0.copy(order="F")
    ol.oracle_k18(M, N, K, 1.0, A, B, 1.0, C)
    idx = [(0, 0), (39, 23), (17, 5)]
    before = [C0[i, j] for i, j in idx]; after = [C[i, j] for i, j in idx]
    assert bench.spot_check(before, after, idx, M, K, 128, 1.0, 1.0) <= 1.0
    assert bench.spot_check(before, [a * (1 + 1e-9) + 1e-9 for a in after], idx, M, K, 128, 1.0, 1.0) > 1.0


def test_parity_blocks_catch_one_wrong_entry_and_cover_the_seams():
    """bench.py's block check of sizes the oracle cannot run whole: blocks of 192 x 16 outputs through
    oracle_dgemm_k18 on compact copies.  A correct panel passes with ratio ~0, one entry off by 1e-9 fails; the positions
    include the corners, the stream-K tail region and a block across every seam."""
    import numpy as np
    import bench
    import oracle_lib as ol
    M, N, K, n0 = 700, 300, 513, 40
    A, B = ol.fill_signed(M, K, M, 1), ol.fill_signed(K, N, K, 2, col0=n0)
    C0 = ol.fill_signed(M, N, M, 3, col0=n0)
    Cm = C0.copy(order="F")
    ol.oracle_k18(M, N, K, 1.5, A, B, 0.5, Cm)
    blocks = bench.parity_block_positions(M, N, [128], 3)
    assert (0, 0, 192, 16) in blocks and (M - 192, N - 16, 192, 16) in blocks
    assert any(j0 < 128 < j0 + w for _i, j0, _h, w in blocks), "no block across the seam at column 128"
    cut = lambda X: (lambda b: X[b[0]: b[0] + b[2], b[1]: b[1] + b[3]])
    assert bench.parity_blocks(blocks, K, n0, 1.5, 0.5, cut(C0), cut(Cm)) <= 2.0
    i0, j0, _h, _w = blocks[5]
    Cm[i0 + 3, j0 + 2] += 1e-9
    assert bench.parity_blocks(blocks, K, n0, 1.5, 0.5, cut(C0), cut(Cm)) > 2.0
    big = bench.parity_block_positions(65536, 8192, [2048, 4096, 6144], 1)
    assert len(big) >= 16 and all(0 <= i and i + h <= 65536 and 0 <= j and j + w <= 8192 for i, j, h, w in big)
    for seam in (2048, 4096, 6144):
        assert any(j < seam < j + w for _i, j, _h, w in big)
    assert bench.parity_block_positions(100, 7, [], 1) == [(0, 0, 100, 7)]
