"""CPU: the multi-GPU sharding layer's host logic (partition, slab order, beta-once semantics), single process
and with world_size-2/3 gloo process groups standing in for NCCL.  The GEMM stand-in is the oracle — here it is
the checker AND the arithmetic, because what is under test is the sharding/stitching, not the CUDA kernel."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle_lib as ol

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_partition_properties(dg):
    from dgemm_b200.sharding import partition
    for total in (0, 1, 100, 1000, 1003, 16384, 65536):
        for parts in (1, 2, 3, 4, 8):
            for align in (16, 128):
                b = partition(total, parts, align)
                assert b[0] == 0 and b[-1] == total and len(b) == parts + 1
                assert all(b[i] <= b[i + 1] for i in range(parts))
                assert all(x % align == 0 for x in b[:-1] if x < total)
    assert partition(65536, 8, 128) == [i * 8192 for i in range(9)]


def test_slab_order_starts_at_own_slab(dg):
    from dgemm_b200.sharding import slab_order
    for world in (1, 2, 4, 8):
        for r in range(world):
            o = slab_order(r, world)
            assert o[0] == r and sorted(o) == list(range(world))


def _sharded_reference(M, N, K, alpha, beta, world, A, B, C0):
    """Runs sharded_dgemm_step for every rank in one process; returns the stitched C."""
    from dgemm_b200.sharding import partition, sharded_dgemm_step
    nb, kb = partition(N, world, 128), partition(K, world, 16)
    C = C0.copy(order="F")
    for rank in range(world):
        n0, n1 = nb[rank], nb[rank + 1]
        if n1 <= n0:
            continue
        Bp = np.asfortranarray(B[:, n0:n1]); Cp = np.asfortranarray(C[:, n0:n1])
        calls = []

        def gemm(kbeg, kc, b_eff):
            calls.append(b_eff)
            Ak = np.asfortranarray(A[:, kbeg:kbeg + kc]) if kc else ol.colmajor(A.shape[0], 1)
            Bk = np.asfortranarray(Bp[kbeg:kbeg + kc, :]) if kc else ol.colmajor(1, n1 - n0)
            ol.oracle_k18(M, n1 - n0, kc, alpha, Ak, Bk, b_eff, Cp)

        sharded_dgemm_step(rank, world, kb, beta, lambda g: g, lambda h: None, gemm)
        assert calls[0] == beta and all(b == 1.0 for b in calls[1:])  # beta applied exactly once
        C[:, n0:n1] = Cp
    return C


@pytest.mark.parametrize("world", [1, 2, 3, 8])
@pytest.mark.parametrize("shape", [(64, 300, 100), (50, 1000, 37), (32, 130, 0), (16, 40, 16)])
def test_sharded_equals_unsharded(dg, world, shape):
    M, N, K = shape
    A = ol.fill_signed(M, K, M, 1) if K else ol.colmajor(M, 1)
    B = ol.fill_signed(K, N, max(K, 1), 2)
    C0 = ol.fill_signed(M, N, M, 3)
    want = C0.copy(order="F")
    ol.oracle_k18(M, N, K, 1.5, A, B, 0.5, want)
    got = _sharded_reference(M, N, K, 1.5, 0.5, world, A, B, C0)
    # same per-entry products; only the K-slab association differs -> within the north-star bound, not bitwise
    ratio, _ = ol.bound_ratio(M, N, K, 1.5, A, B, 0.5, C0, got, want)
    assert ratio <= 2.0


@pytest.mark.parametrize("world", [1, 2, 3, 8])
@pytest.mark.parametrize("K", [0, 16, 37, 200])
def test_fused_order_never_uses_a_slab_before_it_was_fetched(dg, world, K):
    """sharded_dgemm_step_fused: slab i+1 is pulled by the call that multiplies slab i; beta is applied once; every
    non-empty slab is multiplied exactly once; the last call fetches nothing."""
    from dgemm_b200.sharding import partition, sharded_dgemm_step_fused
    kb = partition(K, world, 16)
    for rank in range(world):
        present = {rank}
        used, betas, fetches = [], [], []

        def gemm(kbeg, kc, b_eff, fetch):
            if kc:
                g = max(i for i in range(world) if kb[i] == kbeg and kb[i + 1] > kb[i])
                assert g in present, f"rank {rank} multiplied slab {g} before it arrived"
                used.append(g)
            betas.append(b_eff)
            fetches.append(fetch)
            if fetch is not None:
                owner, fk, fc = fetch
                assert owner != rank and fk == kb[owner] and fc == kb[owner + 1] - kb[owner] and fc > 0
                present.add(owner)

        sharded_dgemm_step_fused(rank, world, kb, 0.5, gemm)
        assert sorted(used) == [g for g in range(world) if kb[g + 1] > kb[g]]
        assert betas[0] == 0.5 and all(b == 1.0 for b in betas[1:])
        assert fetches[-1] is None


def _gloo_worker(rank, world, port, M, N, K, alpha, beta, out_dir):
    sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
    import dgemm_b200  # noqa: F401
    from dgemm_b200.sharding import partition, sharded_dgemm_step
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    nb, kb = partition(N, world, 128), partition(K, world, 16)
    n0, n1 = nb[rank], nb[rank + 1]
    # every rank generates ONLY its own K-slab of A (others start as NaN and must arrive by broadcast)
    A = torch.full((K, M), float("nan"), dtype=torch.float64)      # row-major (K, M) == column-major M x K
    own = ol.fill_signed(M, kb[rank + 1] - kb[rank], M, 1, col0=kb[rank])
    A[kb[rank]:kb[rank + 1]] = torch.from_numpy(np.ascontiguousarray(own.T))
    Bp = ol.fill_signed(K, n1 - n0, max(K, 1), 2, col0=n0)
    Cp = ol.fill_signed(M, n1 - n0, M, 3, col0=n0)

    def gemm(kbeg, kc, b_eff):
        Ak = np.asfortranarray(A[kbeg:kbeg + kc].numpy().T)
        assert not np.isnan(Ak).any(), "slab used before it arrived"
        ol.oracle_k18(M, n1 - n0, kc, alpha, Ak, np.asfortranarray(Bp[kbeg:kbeg + kc, :]), b_eff, Cp)

    sharded_dgemm_step(rank, world, kb, beta,
                       lambda g: dist.broadcast(A[kb[g]:kb[g + 1]], src=g, async_op=True),
                       lambda w: w.wait(), gemm)
    np.save(os.path.join(out_dir, f"c{rank}.npy"), Cp)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_gloo_multi_rank_panels_stitch(dg, tmp_path, world):
    M, N, K, alpha, beta = 48, 384, 200, 2.0, 0.5
    port = 29500 + (os.getpid() % 2000) + world
    mp.spawn(_gloo_worker, args=(world, port, M, N, K, alpha, beta, str(tmp_path)), nprocs=world, join=True)
    from dgemm_b200.sharding import partition
    nb = partition(N, world, 128)
    A = ol.fill_signed(M, K, M, 1); B = ol.fill_signed(K, N, K, 2); C0 = ol.fill_signed(M, N, M, 3)
    want = C0.copy(order="F")
    ol.oracle_k18(M, N, K, alpha, A, B, beta, want)
    got = np.concatenate([np.load(tmp_path / f"c{r}.npy") for r in range(world) if nb[r + 1] > nb[r]], axis=1)
    ratio, _ = ol.bound_ratio(M, N, K, alpha, A, B, beta, C0, np.asfortranarray(got), want)
    assert ratio <= 2.0
