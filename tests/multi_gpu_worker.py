"""Worker for the -m gpu multi-GPU parity test: run under torchrun with N ranks (one per GPU).
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 --master-port P tests/multi_gpu_worker.py M N K alpha beta [fused|nccl]
Each rank owns a column panel of B/C and a K-slab of A (generated on the device), runs ONE sharded DGEMM step
(NVLink exchange of A + the CUDA DGEMM per slab) and checks its C panel against the oracle on the host."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np
import torch
import torch.distributed as dist

import dgemm_b200 as dg
import oracle_lib as ol
from dgemm_b200.sharding import partition, sharded_dgemm_step, sharded_dgemm_step_fused


def main():
    M, N, K = (int(x) for x in sys.argv[1:4])
    alpha, beta = float(sys.argv[4]), float(sys.argv[5])
    mode = sys.argv[6] if len(sys.argv) > 6 else "fused"
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    sptr = torch.cuda.current_stream().cuda_stream
    nb, kb = partition(N, world, 128), partition(K, world, 16)
    if mode == "pipeline":
        return pipeline_mode(M, N, K, alpha, beta, rank, world, dev, nb, kb)
    n0, n1 = nb[rank], nb[rank + 1]
    ncols = n1 - n0
    k0, k1 = kb[rank], kb[rank + 1]
    # foreign slabs must ARRIVE, not pre-exist: the whole of A starts as NaN, only the own slab is generated
    if mode == "fused":
        A_ptr = dg.malloc(8 * M * max(K, 1))
        nanfill = torch.full((M * max(K, 1),), float("nan"), dtype=torch.float64, device=dev)
        dg._check(dg.lib.dgemm_b200_sync(None))
        torch.cuda.synchronize()
        # device-to-device copy of the NaN pattern into the IPC-able buffer via the fused-fetch path of a K=0 call
        dg.dgemm_device_fetch(0, 0, 0, 0.0, 0, 1, 0, 1, 0.0, 0, 1, sptr, nanfill.data_ptr(), A_ptr, 8 * M * max(K, 1))
        torch.cuda.synchronize()
        handles = [None] * world
        dist.all_gather_object(handles, dg.ipc_export(A_ptr))
        peers = {g: dg.ipc_open(handles[g]) for g in range(world) if g != rank}
    else:
        A = torch.full((M * max(K, 1),), float("nan"), dtype=torch.float64, device=dev)
        A_ptr = A.data_ptr()
    B = torch.empty(max(K, 1) * max(ncols, 1), dtype=torch.float64, device=dev)
    Cm = torch.empty(M * max(ncols, 1), dtype=torch.float64, device=dev)
    dg.fill_signed(A_ptr + 8 * k0 * M, M, k1 - k0, M, 1, 0, k0, sptr)
    dg.fill_signed(B.data_ptr(), K, ncols, max(K, 1), 2, 0, n0, sptr)
    dg.fill_signed(Cm.data_ptr(), M, ncols, M, 3, 0, n0, sptr)
    torch.cuda.synchronize()
    dist.barrier()
    ldb = max(K, 1)
    if mode == "fused":
        def gemm_fetch(kbeg, kc, b_eff, fetch):
            src = dst = nbytes = 0
            if fetch is not None:
                owner, fk, fc = fetch
                src, dst, nbytes = peers[owner] + 8 * fk * M, A_ptr + 8 * fk * M, 8 * fc * M
            dg.dgemm_device_fetch(M, ncols, kc, alpha, A_ptr + 8 * kbeg * M, M, B.data_ptr() + 8 * kbeg, ldb, b_eff,
                                  Cm.data_ptr(), M, sptr, src, dst, nbytes)
        sharded_dgemm_step_fused(rank, world, kb, beta, gemm_fetch)
    else:
        sharded_dgemm_step(
            rank, world, kb, beta,
            lambda g: dist.broadcast(A[kb[g] * M: kb[g + 1] * M], src=g, async_op=True),
            lambda w: w.wait(),
            lambda kbeg, kc, b_eff: dg.dgemm_device(M, ncols, kc, alpha, A_ptr + 8 * kbeg * M, M,
                                                    B.data_ptr() + 8 * kbeg, ldb, b_eff, Cm.data_ptr(), M, sptr))
    torch.cuda.synchronize()
    ok = True
    if ncols > 0:
        got = np.asfortranarray(Cm.cpu().numpy().reshape(ncols, M).T)
        Ah = ol.fill_signed(M, K, M, 1) if K else ol.colmajor(M, 1)
        Bh = ol.fill_signed(K, ncols, max(K, 1), 2, col0=n0)
        C0 = ol.fill_signed(M, ncols, M, 3, col0=n0)
        want = C0.copy(order="F")
        ol.oracle_k18(M, ncols, K, alpha, Ah, Bh, beta, want)
        ratio, max_abs = ol.bound_ratio(M, ncols, K, alpha, Ah, Bh, beta, C0, got, want)
        ok = ratio <= 2.0
        print(f"rank {rank}: panel [{n0},{n1}) max |err|/bound = {ratio:.3g} (max abs {max_abs:.3g}) {'OK' if ok else 'FAIL'}",
              flush=True)
    flag = torch.tensor([0 if ok else 1], device=dev)
    dist.all_reduce(flag)
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(1 if flag.item() else 0)


def pipeline_mode(M, N, K, alpha, beta, rank, world, dev, nb, kb):
    """dgemm_b200_rank_pipeline from HOST operands, real peers: every rank uploads its own K-slab, pulls the others over
    NVLink behind their sequence words, and its whole C panel is checked against the oracle.  Two passes (pass numbers 0, 1)."""
    host = dist.new_group(backend="gloo")
    n0, ncols = nb[rank], nb[rank + 1] - nb[rank]
    k0, kc = kb[rank], kb[rank + 1] - kb[rank]
    x = dg.exchange_alloc(M, K)
    handles = [None] * world
    dist.all_gather_object(handles, dg.ipc_export(x), group=host)
    peers = {g: dg.ipc_open(handles[g]) for g in range(world) if g != rank}
    xbuf = [x if g == rank else peers[g] for g in range(world)]
    Ah = ol.fill_signed(M, K, M, 1)
    Bh = ol.fill_signed(K, max(ncols, 1), max(K, 1), 2, col0=n0)
    C0 = ol.fill_signed(M, max(ncols, 1), M, 3, col0=n0)
    A_slab = np.asfortranarray(Ah[:, k0:k0 + kc])
    ok = True
    Cin = C0
    for p in range(2):
        Cg = Cin.copy(order="F")
        dg.rank_pipeline(M, ncols, K, alpha, A_slab, M, Bh, Bh.shape[0], beta, Cg, M, rank, world, kb, xbuf, p)
        dist.barrier(group=host)
        if ncols > 0:
            want = Cin.copy(order="F")
            ol.oracle_k18(M, ncols, K, alpha, Ah, Bh, beta, want)
            ratio, max_abs = ol.bound_ratio(M, ncols, K, alpha, Ah, Bh, beta, Cin, Cg, want)
            good = ratio <= 2.0
            ok = ok and good
            print(f"rank {rank} pass {p}: panel [{n0},{n0 + ncols}) max |err|/bound = {ratio:.3g} (max abs {max_abs:.3g}) "
                  f"{'OK' if good else 'FAIL'}", flush=True)
        Cin = Cg
    flag = torch.tensor([0 if ok else 1], device=dev)
    dist.all_reduce(flag)
    dist.barrier(group=host)
    for q in peers.values():
        dg.ipc_close(q)
    dist.barrier(group=host)
    dg.free(x)
    dist.destroy_process_group()
    sys.exit(1 if flag.item() else 0)


if __name__ == "__main__":
    main()
