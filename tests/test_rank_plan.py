"""CPU tests of the multi-GPU rank schedule (csrc/dgemm_rankplan.cpp through dgemm_b200_rank_plan; no GPU, no compute in
the library): structural invariants of every rank's plan, and a numpy emulation of a whole world of ranks executing their
plans against each other's exchange buffers -- the way dgemm_b200_rank_pipeline / mydgemm_gpu_v21 run them on GPUs -- checked
against the oracle.  The reference's counterpart is partition_m_dim + the barrier structure of mydgemm_cpu_v19
(include/kernel19.h:636-656, 695-745)."""
import numpy as np
import pytest

import oracle_lib as ol
from dgemm_b200.sharding import partition


def plans(dg, M, N, K, G, need_c_in=True, p1=0):
    nb, kb = partition(N, G, 128), partition(K, G, 16)
    return nb, kb, [dg.rank_plan(M, nb[r + 1] - nb[r], K, G, r, kb, need_c_in, p1) for r in range(G)]


SHAPES = [(65536, 65536, 65536, 8), (65536, 65536, 65536, 2), (32768, 32768, 32768, 4), (4096, 8192, 4096, 8),
          (1000, 1003, 997, 2), (1000, 1003, 997, 3), (512, 640, 4096, 4), (2048, 1024, 128, 8), (300, 20000, 48, 3),
          (4096, 2048, 4096, 1), (64, 128, 16, 1), (16384, 16384, 16384, 5), (16384, 16384, 16384, 7)]


@pytest.mark.parametrize("M,N,K,G", SHAPES)
@pytest.mark.parametrize("p1", [0, 1, 2, 99])
def test_rank_plan_invariants(dg, M, N, K, G, p1):
    nb, kb, pl = plans(dg, M, N, K, G, True, p1)
    for r, p in enumerate(pl):
        ncols = nb[r + 1] - nb[r]
        calls, copies = p["calls"], p["copies"]
        # (c) copies: every own piece once, in ascending order, covering the own slab exactly
        a_copies = [c for c in copies if c["kind"] == 0]
        assert [c["piece"] for c in a_copies] == list(range(len(a_copies)))
        pos = kb[r]
        for c in a_copies:
            assert c["k0"] == pos and c["kc"] > 0
            pos += c["kc"]
        assert pos == kb[r + 1]
        if ncols == 0:
            assert not calls
            continue
        # (a) per sub-panel (the column ranges of the calls that finish one): the calls that touch it contain it whole, their
        # K ranges tile [0, K) exactly once; first / last flags sit on the right calls
        subs = sorted({(c["c0"], c["nc"]) for c in calls if c["last"]})
        assert subs[0][0] == 0 and subs[-1][0] + subs[-1][1] == ncols
        for (c0, nc), nxt in zip(subs, subs[1:] + [(ncols, 0)]):
            assert c0 + nc == nxt[0] and nc > 0
            mine = [c for c in calls if c["c0"] < c0 + nc and c0 < c["c0"] + c["nc"]]
            assert all(c["c0"] <= c0 and c0 + nc <= c["c0"] + c["nc"] for c in mine)
            assert all((c["c0"], c["nc"]) in ((c0, nc), (0, ncols)) for c in mine)
            cover = np.zeros(K, dtype=np.int32)
            for c in mine:
                cover[c["k0"]: c["k0"] + c["kc"]] += 1
            assert (cover == 1).all()
            assert [c["first"] for c in mine] == [1] + [0] * (len(mine) - 1)
            assert [c["last"] for c in mine] == [0] * (len(mine) - 1) + [1]
            assert mine[-1]["wait_x_copy"] >= 0 and copies[mine[-1]["wait_x_copy"]]["kind"] == 2
        # (b) pulls: every foreign column exactly once, complete before its first use, need_piece inside the owner's pieces
        pulled = np.zeros(K, dtype=np.int32)
        have = np.zeros(K, dtype=bool)
        have[kb[r]: kb[r + 1]] = True
        for c in calls:
            assert have[c["k0"]: c["k0"] + c["kc"]].all(), "a call multiplies columns of A that have not been pulled yet"
            if c["f_owner"] >= 0:
                g = c["f_owner"]
                assert g != r and kb[g] <= c["f_k0"] and c["f_k0"] + c["f_kc"] <= kb[g + 1] and c["f_kc"] > 0
                assert c["f_k0"] % 2 == 0 and c["f_kc"] % 2 == 0 or c["f_k0"] + c["f_kc"] == kb[g + 1]
                pulled[c["f_k0"]: c["f_k0"] + c["f_kc"]] += 1
                n_pieces = len([x for x in pl[g]["copies"] if x["kind"] == 0])
                assert 1 <= c["f_need_piece"] <= n_pieces
                # the piece named must really contain the last pulled column
                ends = np.cumsum([x["kc"] for x in pl[g]["copies"] if x["kind"] == 0]) + kb[g]
                assert ends[c["f_need_piece"] - 1] >= c["f_k0"] + c["f_kc"]
                assert c["f_need_piece"] == 1 or ends[c["f_need_piece"] - 2] < c["f_k0"] + c["f_kc"]
            # a pull is complete when its call is (stream order): usable from the NEXT call on
            if c["f_owner"] >= 0:
                have[c["f_k0"]: c["f_k0"] + c["f_kc"]] = True
        foreign = np.ones(K, dtype=bool)
        foreign[kb[r]: kb[r + 1]] = False
        assert (pulled[foreign] == 1).all() and (pulled[~foreign] == 0).all()
        # (c) every call waits for its own B block, and its A piece (own slab) was queued before that block
        for c in calls:
            b = copies[c["wait_copy"]]
            assert b["kind"] == 1 and (b["k0"], b["kc"], b["c0"], b["nc"]) == (c["k0"], c["kc"], c["c0"], c["nc"])
            if kb[r] <= c["k0"] < kb[r + 1]:
                assert any(x["kind"] == 0 and x["k0"] <= c["k0"] and c["k0"] + c["kc"] <= x["k0"] + x["kc"]
                           for x in copies[: c["wait_copy"]])


def test_rank_plan_rejects_an_empty_own_slab(dg):
    kb = partition(32, 4, 16)   # 16, 16, 0, 0 columns
    with pytest.raises(dg.DgemmError):
        dg.rank_plan(256, 128, 32, 4, 3, kb)
    assert dg.rank_plan(256, 128, 32, 4, 0, kb)["calls"]   # a rank with a slab is fine; empty slabs are skipped


def emulate(dg, M, N, K, G, alpha, beta, p1=0, seed=5):
    """All ranks execute their plans in lock-step rounds of one call each; a pull reads the OWNER's exchange buffer as it is at
    that moment (NaN where the owner has not uploaded yet) and must therefore respect f_need_piece."""
    nb, kb, pl = plans(dg, M, N, K, G, beta != 0.0, p1)
    A, B, C0 = ol.fill_signed(M, K, M, seed), ol.fill_signed(K, N, K, seed + 1), ol.fill_signed(M, N, M, seed + 2)
    xbuf = [np.full((M, K), np.nan) for _ in range(G)]          # exchange buffers
    landed = [0] * G                                           # sequence words
    Cd = [np.full((M, nb[r + 1] - nb[r]), np.nan) for r in range(G)]
    Xd = [np.full((M, nb[r + 1] - nb[r]), np.nan) for r in range(G)]
    Bd = [np.full((K, nb[r + 1] - nb[r]), np.nan) for r in range(G)]
    out = np.full((M, N), np.nan, order="F")
    next_copy = [0] * G

    def run_copies(r, through):
        while next_copy[r] <= through:
            c = pl[r]["copies"][next_copy[r]]
            if c["kind"] == 0:
                xbuf[r][:, c["k0"]: c["k0"] + c["kc"]] = A[:M, c["k0"]: c["k0"] + c["kc"]]
                landed[r] = c["piece"] + 1
            elif c["kind"] == 1:
                Bd[r][c["k0"]: c["k0"] + c["kc"], c["c0"]: c["c0"] + c["nc"]] = \
                    B[c["k0"]: c["k0"] + c["kc"], nb[r] + c["c0"]: nb[r] + c["c0"] + c["nc"]]
            else:
                Xd[r][:, c["c0"]: c["c0"] + c["nc"]] = C0[:M, nb[r] + c["c0"]: nb[r] + c["c0"] + c["nc"]]
            next_copy[r] += 1

    pos = [0] * G
    stalled = 0
    while any(pos[r] < len(pl[r]["calls"]) or next_copy[r] < len(pl[r]["copies"]) for r in range(G)):
        progressed = False
        for r in range(G):
            if pos[r] >= len(pl[r]["calls"]):
                if next_copy[r] < len(pl[r]["copies"]):
                    run_copies(r, len(pl[r]["copies"]) - 1)
                    progressed = True
                continue
            c = pl[r]["calls"][pos[r]]
            run_copies(r, max(c["wait_copy"], c["wait_x_copy"]))
            if c["f_owner"] >= 0 and landed[c["f_owner"]] < c["f_need_piece"]:
                continue   # the wait kernel holds the compute stream; the owner advances in its own turn
            sl = slice(c["k0"], c["k0"] + c["kc"])
            cs = slice(c["c0"], c["c0"] + c["nc"])
            acc = alpha * (xbuf[r][:, sl] @ Bd[r][sl, cs])
            Cd[r][:, cs] = acc if c["first"] else Cd[r][:, cs] + acc
            if c["f_owner"] >= 0:
                fs = slice(c["f_k0"], c["f_k0"] + c["f_kc"])
                xbuf[r][:, fs] = xbuf[c["f_owner"]][:, fs]
            if c["last"]:
                if beta != 0.0:
                    Cd[r][:, cs] += beta * Xd[r][:, cs]
                out[:, nb[r] + c["c0"]: nb[r] + c["c0"] + c["nc"]] = Cd[r][:, cs]
            pos[r] += 1
            progressed = True
        stalled = 0 if progressed else stalled + 1
        assert stalled < 3, "deadlock: every rank waits for a piece nobody uploads"
    want = C0.copy(order="F")
    if beta == 0.0:
        want[:] = 0.0
    ol.oracle_k18(M, N, K, alpha, A, B, beta, want)
    ratio, _ = ol.bound_ratio(M, N, K, alpha, A, B, beta, C0 if beta != 0.0 else np.zeros_like(C0), out, want)
    return ratio


@pytest.mark.parametrize("M,N,K,G,p1", [(96, 1024, 256, 2, 0), (64, 1536, 512, 3, 0), (64, 2048, 1024, 4, 0),
                                        (32, 4096, 2048, 8, 0), (64, 2048, 1024, 4, 1), (64, 2048, 1024, 4, 99),
                                        (50, 1003, 997, 3, 0), (40, 7000, 640, 5, 2), (64, 300, 4096, 4, 0)])
@pytest.mark.parametrize("alpha,beta", [(1.0, 1.0), (2.5, 0.0), (-1.0, 0.5)])
def test_world_of_ranks_emulated_against_the_oracle(dg, M, N, K, G, p1, alpha, beta):
    """A pull that ran ahead of its owner's upload would copy NaN; a call on columns not yet pulled would multiply NaN."""
    assert emulate(dg, M, N, K, G, alpha, beta, p1) <= 2.0
