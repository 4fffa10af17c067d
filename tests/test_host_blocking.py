"""Out-of-core blocking of the one-GPU host path (dgemm_b200_host_blocking, csrc/dgemm_api.cu: choose_host_blocking): pure
host logic, no GPU.  The reference's counterpart is its outer N / K blocking (include/kernel18.h:652-669)."""
import pytest


def footprint(Mb, Nb, Kb, need_ab=True, num_sms=148):
    ldm, ldk = (Mb + 1) & ~1, (Kb + 1) & ~1
    b = 2 * ldm * Nb * 8
    if need_ab:
        b += (ldm * Kb + ldk * Nb) * 8
    return b + num_sms * 2 * 128 * 128 * 8


@pytest.mark.parametrize("M,N,K,budget", [(8192, 8192, 8192, 256 << 20), (8192, 8192, 8192, 1 << 30), (3000, 3000, 3000, 100 << 20),
                                          (100000, 100000, 100000, int(160e9)), (200000, 150000, 100000, int(160e9)),
                                          (4096, 4096, 1000000, 1 << 30), (1000, 1003, 997, 50 << 20),
                                          (300, 300, 4000, 43_000_000), (70000, 128, 70000, 1 << 30)])
def test_blocks_fit_the_budget_and_cover_the_problem(dg, M, N, K, budget):
    r = dg.host_blocking(M, N, K, budget)
    Mb, Nb, Kb = r["Mb"], r["Nb"], r["Kb"]
    assert 0 < Mb <= M and 0 < Nb <= N and 0 < Kb <= K
    assert footprint(Mb, Nb, Kb) <= budget
    assert Mb == M or Mb % 128 == 0
    assert Nb == N or Nb % 128 == 0
    assert Kb == K or Kb % 16 == 0
    mn, nn, kn = -(-M // Mb), -(-N // Nb), -(-K // Kb)
    want = 8.0 * (M * K * nn + K * N * mn + M * N * 2 * kn)
    assert r["pcie_bytes"] == pytest.approx(want)
    # K is only cut when a block with the whole of K would be smaller than 128 x 128 ... or moves more bytes
    if Kb < K:
        whole_k_fits = footprint(min(M, 128), min(N, 128), K) <= budget
        if whole_k_fits:
            alt = dg.host_blocking(M, N, K, budget)   # deterministic
            assert alt == r


def test_a_call_that_fits_is_not_cut(dg):
    r = dg.host_blocking(2048, 2048, 2048, 1 << 30)
    assert (r["Mb"], r["Nb"], r["Kb"]) == (2048, 2048, 2048)
    assert r["pcie_bytes"] == 8.0 * 2048 * 2048 * 4
    r = dg.host_blocking(2048, 2048, 2048, 1 << 30, beta=0.0)
    assert r["pcie_bytes"] == 8.0 * 2048 * 2048 * 3
    r = dg.host_blocking(2048, 2048, 2048, 1 << 30, alpha=0.0)   # C = beta * C: A and B are not sent
    assert r["pcie_bytes"] == 8.0 * 2048 * 2048 * 2


def test_more_budget_never_moves_more_bytes(dg):
    last = None
    for mb in (64, 128, 256, 512, 1024, 2048, 4096):
        r = dg.host_blocking(8192, 8192, 8192, mb << 20)
        if last is not None:
            assert r["pcie_bytes"] <= last * 1.0001
        last = r["pcie_bytes"]


def test_a_budget_below_the_smallest_block_is_an_error(dg):
    with pytest.raises(dg.DgemmError) as e:
        dg.host_blocking(8192, 8192, 8192, 16 << 20)
    assert e.value.status == dg.ERR_OOM
    with pytest.raises(dg.DgemmError):
        dg.host_blocking(10, 10, 10, 0)
