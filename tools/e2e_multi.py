"""Multi-GPU end-to-end experiments (torchrun, one rank per GPU): the rank pipeline under different schedules, each with its
per-rank timeline, plus the aggregate host->GPU bandwidth of the box.  One JSON line per experiment on rank 0.

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/e2e_multi.py \
        [--size 65536] [--variants default,ring,p1=2] [--passes 2] [--out gpurun_out/e2e_multi.jsonl]
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import torch
import torch.distributed as dist

import bench
import dgemm_b200 as dg
from dgemm_b200.sharding import partition


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--size", type=int, default=65536)
    ap.add_argument("--variants", default="default,ring")
    ap.add_argument("--passes", type=int, default=2)
    ap.add_argument("--out", default=None)
    ap.add_argument("--timeline-dir", default=None)
    args = ap.parse_args()
    rank, N = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    numa = bench.bind_to_gpu_numa(torch, local) if os.environ.get("BENCH_NO_NUMA") != "1" else "off"
    dist.init_process_group("nccl", device_id=dev)
    host = dist.new_group(backend="gloo")
    M = Nn = K = args.size
    nb, kb = partition(Nn, N, 128), partition(K, N, 16)
    n0, ncols = nb[rank], nb[rank + 1] - nb[rank]
    k0, kc = kb[rank], kb[rank + 1] - kb[rank]
    out = []

    def emit(rec):
        if rank == 0:
            print(json.dumps(rec), flush=True)
            out.append(rec)

    # operands: generated on the device, copied out to pinned host buffers
    hA = torch.empty(M * kc, dtype=torch.float64).pin_memory()
    hB = torch.empty(K * ncols, dtype=torch.float64).pin_memory()
    hC = torch.empty(M * ncols, dtype=torch.float64).pin_memory()
    tmp = torch.empty(max(M * kc, K * ncols, M * ncols), dtype=torch.float64, device=dev)
    for h, rows, cols, seed, r0, c0 in ((hA, M, kc, 1, 0, k0), (hB, K, ncols, 2, 0, n0), (hC, M, ncols, 3, 0, n0)):
        dg.fill_signed(tmp.data_ptr(), rows, cols, rows, seed, r0, c0, 0)
        torch.cuda.synchronize()
        h.copy_(tmp[: rows * cols])
    # ---- aggregate H2D / D2H bandwidth: every rank moves the same buffer at the same time
    nbytes = min(8 * M * ncols, 4 << 30)
    for direction in ("h2d", "d2h", "both"):
        dist.barrier(group=host)
        torch.cuda.synchronize()
        s2 = torch.cuda.Stream()
        t0 = time.perf_counter()
        if direction in ("h2d", "both"):
            tmp[: nbytes // 8].copy_(hC[: nbytes // 8], non_blocking=True)
        if direction in ("d2h", "both"):
            with torch.cuda.stream(s2):
                hB[: nbytes // 8].copy_(tmp[nbytes // 8: 2 * (nbytes // 8)] if tmp.numel() >= 2 * (nbytes // 8) else tmp[: nbytes // 8],
                                        non_blocking=True)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        t = torch.tensor([dt], dtype=torch.float64, device=dev)
        allt = [torch.zeros_like(t) for _ in range(N)]
        dist.all_gather(allt, t)
        per = [nbytes / float(x.item()) / 1e9 * (2 if direction == "both" else 1) for x in allt]
        emit({"probe": f"simultaneous_{direction}", "gbytes_per_rank": nbytes / 1e9, "gb_per_s_per_rank": [round(p, 1) for p in per],
              "aggregate_gb_per_s": round(sum(per), 1), "numa": numa})
    # B was clobbered by the D2H probe: regenerate
    dg.fill_signed(tmp.data_ptr(), K, ncols, K, 2, 0, n0, 0)
    torch.cuda.synchronize()
    hB.copy_(tmp[: K * ncols])
    del tmp
    torch.cuda.empty_cache()

    x = dg.exchange_alloc(M, K)
    handles = [None] * N
    dist.all_gather_object(handles, dg.ipc_export(x), group=host)
    peers = {g: dg.ipc_open(handles[g]) for g in range(N) if g != rank}
    xbuf = [x if g == rank else peers[g] for g in range(N)]
    pass_no = 0
    import numpy as np
    for variant in args.variants.split(","):
        os.environ.pop("DGEMM_B200_RANK_P1", None)
        os.environ.pop("DGEMM_B200_RANK_SUBPANEL", None)
        if variant == "ring":
            os.environ["DGEMM_B200_RANK_P1"] = "99"
        elif variant.startswith("p1="):
            os.environ["DGEMM_B200_RANK_P1"] = variant[3:]
        elif variant.startswith("sub="):
            os.environ["DGEMM_B200_RANK_SUBPANEL"] = variant[4:]
        times, tl = [], []
        rng = np.random.default_rng(1234 + rank)
        idx = [(int(rng.integers(M)), int(rng.integers(ncols))) for _ in range(4)] + [(M - 1, ncols - 1), (0, 0)]
        check = 0.0
        for p in range(args.passes + 2):   # pass 0 warms up (buffers grow), the last one carries the timeline and the check
            last = p == args.passes + 1
            dist.barrier(group=host)
            before = [float(hC[j * M + i]) for i, j in idx]
            t0 = time.perf_counter()
            tl = dg.rank_pipeline(M, ncols, K, 1.0, hA.data_ptr(), M, hB.data_ptr(), K, 1.0, hC.data_ptr(), M, rank, N, kb, xbuf,
                                  pass_no, 30000 if last else 0)
            dt = time.perf_counter() - t0
            pass_no += 1
            dist.barrier(group=host)
            if last:
                after = [float(hC[j * M + i]) for i, j in idx]
                check = bench.spot_check(before, after, idx, M, K, n0, 1.0, 1.0)
                mine = bench.summarise_timeline(tl)
                full_tl = tl
            elif p > 0:
                times.append(dt)
        t = torch.tensor([max(times), min(times), check], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        allr = [None] * N
        dist.all_gather_object(allr, mine, group=host)
        if args.timeline_dir:
            allt = [None] * N
            dist.all_gather_object(allt, full_tl, group=host)
            if rank == 0:
                os.makedirs(args.timeline_dir, exist_ok=True)
                with open(os.path.join(args.timeline_dir, f"e2e_timeline_n{N}_{variant.replace('=', '')}.json"), "w") as f:
                    json.dump({"n_gpus": N, "size": M, "variant": variant,
                               "kinds": {"0": "h2d done for call", "1": "call start", "2": "call end", "3": "d2h of sub-panel done"},
                               "ranks": allt}, f)
        worst = float(t[0].item())
        emit({"variant": variant, "n_gpus": N, "size": M, "e2e_ms_worst_rank_worst_pass": round(1e3 * worst, 2),
              "e2e_tflops": round(2.0 * M * Nn * K / worst / 1e12, 2), "spot_check": float(t[2].item()),
              "plan": {"p1": dg.rank_plan(M, ncols, K, N, rank, kb)["p1"], "calls": len(dg.rank_plan(M, ncols, K, N, rank, kb)["calls"])},
              "timeline_per_rank": allr})
        if not float(t[2].item()) <= 1.0:
            raise SystemExit("WRONG RESULT")
    dist.barrier(group=host)
    for p in peers.values():
        dg.ipc_close(p)
    dist.barrier(group=host)
    dg.free(x)
    if rank == 0 and args.out:
        with open(args.out, "a") as f:
            for rec in out:
                f.write(json.dumps(rec) + "\n")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
