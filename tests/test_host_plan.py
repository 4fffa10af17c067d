"""Host-side schedule of mydgemm_gpu_v20 (csrc/dgemm_hostplan.cpp, dgemm_b200_host_plan): pure host logic, no GPU.

The planner replaces the reference's N -> K -> M blocking loops (include/kernel18.h:652-669) for operands that start
in host memory; these tests check that every schedule is a valid decomposition of the DGEMM: each (slab, panel)
block exactly once, aligned boundaries that cover N and K, and an order in which a K-slab of a panel never
overtakes an earlier one (so `beta` is applied exactly once, by the first call of the panel)."""
import itertools

import pytest

import dgemm_b200 as dg

SHAPES = [(16384, 16384, 16384), (8192, 8192, 8192), (4096, 4096, 4096), (2048, 2048, 2048), (3000, 3000, 3000),
          (1000, 1003, 997), (1001, 1003, 997), (512, 512, 16384), (16384, 256, 512), (256, 16384, 512), (128, 128, 128),
          (100, 100, 100), (1, 1, 1), (7, 9, 8), (65536, 8192, 65536), (300, 5000, 17), (5000, 300, 40000)]


@pytest.mark.parametrize("shape", SHAPES)
@pytest.mark.parametrize("beta", [0.0, 1.0])
def test_plan_is_a_valid_decomposition(shape, beta):
    M, N, K = shape
    nb, kb, order, ms = dg.host_plan(M, N, K, 1.0, beta)
    assert nb[0] == 0 and nb[-1] == N and kb[0] == 0 and kb[-1] == K
    assert all(b > a for a, b in zip(nb, nb[1:])) and all(b > a for a, b in zip(kb, kb[1:]))
    assert all(b % 128 == 0 for b in nb[1:-1]), nb      # panel starts on CTA-tile columns
    assert all(b % 16 == 0 for b in kb[1:-1]), kb       # slab starts on k-tile rows (and 16-byte aligned B blocks)
    P, S = len(nb) - 1, len(kb) - 1
    cells = [(s, p) for (s0, p, s1) in order for s in range(s0, s1)]      # a call covers slabs [s0, s1) of panel p
    assert all(0 <= s0 < s1 <= S and 0 <= p < P for (s0, p, s1) in order)
    assert sorted(cells) == sorted(itertools.product(range(S), range(P)))   # every (slab, panel) block exactly once
    assert len(order) <= 448
    assert ms > 0


@pytest.mark.parametrize("shape", SHAPES[:8])
def test_order_streams_operands_in_index_order(shape):
    """The H2D copies are queued when a call first needs them: slabs and panels must become resident in ascending
    index order (the lead-in pieces are the low indices), and the tapered panels must be the last to finish."""
    M, N, K = shape
    nb, kb, order, _ = dg.host_plan(M, N, K)
    first_slab, first_panel, last_of_panel = {}, {}, {}
    for i, (s0, p, s1) in enumerate(order):
        for s in range(s0, s1):
            first_slab.setdefault(s, i)
        first_panel.setdefault(p, i)
        last_of_panel[p] = i
    for p in first_panel:   # within a panel the K-slabs are applied in ascending order (beta only with the first call)
        seq = [s0 for (s0, pp, s1) in order if pp == p]
        assert seq == sorted(seq)
    assert [s for s, _ in sorted(first_slab.items(), key=lambda kv: (kv[1], kv[0]))] == sorted(first_slab)
    assert [p for p, _ in sorted(first_panel.items(), key=lambda kv: kv[1])] == sorted(first_panel)
    assert max(last_of_panel, key=last_of_panel.get) == len(nb) - 2        # the last panel finishes last
    assert order[-1][1] == len(nb) - 2


def test_degenerate_calls_are_one_piece():
    for (M, N, K, alpha) in [(512, 512, 0, 1.0), (512, 512, 512, 0.0), (0, 512, 512, 1.0), (512, 0, 512, 1.0)]:
        nb, kb, order, _ = dg.host_plan(M, N, K, alpha, 0.5)
        assert order == [(0, 0, 1)] and nb == [0, N] and kb == [0, K]


def test_large_square_plan_has_lead_in_and_taper():
    """16384^3 (the bench workload): small first pieces (short exposed first copy), narrow last panels (short exposed
    last D2H), and a model estimate within 8 % of the kernel-only time at 36.4 TFLOP/s."""
    nb, kb, order, ms = dg.host_plan(16384, 16384, 16384)
    widths = [b - a for a, b in zip(nb, nb[1:])]
    S = len(kb) - 1
    assert order[-1] == (0, len(nb) - 2, S)            # the last panels take the whole K in one call (A is resident)
    assert widths[0] <= 1024 and kb[1] <= 1024
    assert widths[-1] <= 256 and max(widths) >= 2048
    ideal_ms = 2.0 * 16384 ** 3 / 36.4e12 * 1e3
    assert ideal_ms < ms < 1.08 * ideal_ms


def test_plan_rejects_bad_arguments():
    with pytest.raises(dg.DgemmError):
        dg.host_plan(-1, 4, 4)
    with pytest.raises(dg.DgemmError):
        dg.host_plan(16384, 16384, 16384, cap=16)


def test_deferred_c_only_when_c_is_read():
    """beta != 0 on a large problem: the start-up panels are opened without their C (it is added after their last
    call), so the first device call waits for one small slab of A and one B block only.  beta == 0: nothing to defer."""
    nb, kb, order, ms1 = dg.host_plan(16384, 16384, 16384, 1.0, 1.0)
    d = dg.host_plan.deferred_panels
    assert 1 <= d < len(nb) - 1
    assert all(p < d for (_, p, _) in order[:d])          # the first calls open the deferred panels
    dg.host_plan(16384, 16384, 16384, 1.0, 0.0)
    assert dg.host_plan.deferred_panels == 0


@pytest.mark.parametrize("shape", [(128, 4000000, 128), (128, 128, 4000000), (64, 200000, 200000), (8, 3670017, 16)])
def test_planner_always_returns_a_schedule_for_huge_thin_shapes(dg, shape):
    """N or K in the millions with the other dimensions thin: even the widest pieces exceed the item budget; the planner
    must widen them instead of returning an empty plan (which used to abort mydgemm_gpu_v20 with std::length_error)."""
    M, N, K = shape
    nb, kb, order, ms = dg.host_plan(M, N, K, cap=1 << 16)
    assert nb[0] == 0 and nb[-1] == N and kb[0] == 0 and kb[-1] == K and order and ms < 1e12
    assert (len(nb) - 1) * (len(kb) - 1) <= 448
    seen = set()
    for s, p, e in order:
        for x in range(s, e):
            assert (x, p) not in seen
            seen.add((x, p))
    assert len(seen) == (len(nb) - 1) * (len(kb) - 1)
