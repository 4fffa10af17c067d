#!/usr/bin/env python
"""bench.py — the DGEMM hot path on N B200s of one node, one JSON line on rank 0.

    python bench.py --gpus 1 --steps K --warmup W                  (N = 1)
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W                      (N > 1, one rank per GPU, NCCL)
    python bench.py --impl reference ...                            the reference's own CPU kernel19 on the host cores

metric   : FP64 DGEMM TFLOPS = 2*M*N*K / t  (same formula as the reference harness, test.c:49)
workload : N = 1 -> M=N=K=16384 (BASELINE.json metric / configs[1] top size), alpha = beta = 1, column-major, LD = dim
           N > 1 -> M=N=K=65536 sharded by column panels of B and C, A replicated over NVLink (BASELINE configs[3]);
                    strong scaling.  --size overrides both.
step     : one pass C <- alpha*A*B + beta*C over the whole problem.
value    : kernel-only, operands resident in HBM (N > 1: every rank holds its K-slab of A, its B and C panels;
           the NVLink replication of A is INSIDE the timed region).  CUDA events on the launching stream.
e2e      : the same metric through the reference-facing host-pointer entry (mydgemm_gpu_v20 at N = 1), pinned host
           buffers, H2D of A, B, C and D2H of C inside the timed region.
PyTorch is plumbing only: device memory, streams, events, torch.distributed.  The math is libdgemm_b200.so.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

FLOP_PER_CLK_PER_SM = 128.0  # 64 FP64 FMA lanes per SM; confirmed by csrc/peak_probe.cu (127.5 measured with DMMA)
UNIT = "TFLOP/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=None)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--size", type=int, default=None, help="M=N=K override")
    ap.add_argument("--exchange", default="fused", choices=["fused", "nccl"],
                    help="N > 1: how K-slabs of A reach the other GPUs (fused = pulled over NVLink by the GEMM kernel "
                         "itself; nccl = ncclBroadcast per slab, the baseline)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-cublas", action="store_true")
    return ap.parse_args()


# ------------------------------------------------------------------------------------------------------------------
# clocks: nvidia-smi sampled DURING the timed region (B200_PROFILING.md recipe)
# ------------------------------------------------------------------------------------------------------------------
class ClockSampler:
    QUERY = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.rows = []
        self.proc = None
        self.gpu = gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.gpu), f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, smax, power, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            f = [x.strip() for x in r.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); smax.append(float(f[1])); power.append(float(f[2]))
            except ValueError:
                continue
            for name, val in zip(names, f[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(smax), "power_w_max": max(power),
                "samples": len(sm), "reasons": sorted(reasons)}


# ------------------------------------------------------------------------------------------------------------------
# CPU baseline: the reference's own kernel19 (oracle/_ref, compiled from /root/reference in the build container)
# ------------------------------------------------------------------------------------------------------------------
def pick_threads(M, cores):
    """kernel19 needs an even row band per thread: largest nthr <= cores with M % (2*nthr) == 0."""
    for n in range(cores, 0, -1):
        if M % (2 * n) == 0:
            return n
    return 1


def cpu_reference_step(n, nthr, state):
    """One bounded sample: kernel19 on an n^3 problem (n <= 9216: SURVEY.md F2). Returns seconds."""
    import oracle_lib as ol
    if "A" not in state:
        state["A"] = ol.fill_signed(n, n, n, 11)
        state["B"] = ol.fill_signed(n, n, n, 12)
        state["C"] = ol.fill_signed(n, n, n, 13)
        state["kind"] = "reference" if ol.have_ref() else "port"
        if state["kind"] == "reference":
            ol.ref().ref_set_num_threads(nthr)
    t0 = time.perf_counter()
    if state["kind"] == "reference":
        ol.call_gemm(ol.ref().mydgemm_cpu_v19, n, n, n, 1.0, state["A"], state["B"], 1.0, state["C"])
    else:
        ol.oracle_k19(n, n, n, 1.0, state["A"], state["B"], 1.0, state["C"], nthr)
    return time.perf_counter() - t0


def cpu_model():
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("model name"):
                    return line.split(":", 1)[1].strip()
    except OSError:
        pass
    return "unknown"


def run_cpu_baseline(sample_n, reps):
    cores = os.cpu_count() or 1
    nthr = pick_threads(sample_n, cores)
    os.environ.setdefault("OMP_PROC_BIND", "close")
    os.environ.setdefault("OMP_PLACES", "cores")
    state = {}
    cpu_reference_step(sample_n, nthr, state)  # warm-up (first touch, thread team start)
    best = min(cpu_reference_step(sample_n, nthr, state) for _ in range(reps))
    # and the serial kernel18 beside it (north_star: both reference kernels on the box's own cores), one core, 2048^3
    k18 = None
    try:
        import oracle_lib as ol
        if ol.have_ref():
            m = 2048
            a, b, c = ol.fill_signed(m, m, m, 21), ol.fill_signed(m, m, m, 22), ol.fill_signed(m, m, m, 23)
            ol.call_gemm(ol.ref().mydgemm_cpu_v18, m, m, m, 1.0, a, b, 1.0, c)
            t0 = time.perf_counter()
            ol.call_gemm(ol.ref().mydgemm_cpu_v18, m, m, m, 1.0, a, b, 1.0, c)
            dt = time.perf_counter() - t0
            k18 = {"value": 2.0 * m ** 3 / dt / 1e12, "unit": UNIT, "cores": 1, "sample": f"M=N=K={m}, reference kernel18"}
    except Exception as e:  # noqa: BLE001
        k18 = {"value": None, "sample": repr(e)}
    return {"value": 2.0 * sample_n ** 3 / best / 1e12, "unit": UNIT, "cores": nthr, "kind": state["kind"], "kernel18": k18,
            "sample": f"M=N=K={sample_n} alpha=1 beta=1, reference kernel19 (OpenMP, -O0 -march=skylake-avx512 as "
                      f"reference Makefile:3) best of {reps} after 1 warm-up, {nthr} threads on {cores} host cores "
                      f"({cpu_model()}); same 2MNK/t formula",
            "seconds": best}


def reference_arm(args):
    """--impl reference: the reference's CPU implementation of the path, timed on this box's host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    N = args.gpus
    size = args.size or (16384 if N == 1 else 65536)
    sample_n = int(os.environ.get("BENCH_REF_SAMPLE", 8192 if (os.cpu_count() or 1) >= 8 else 4096))
    steps = args.steps or 3
    cores = os.cpu_count() or 1
    nthr = pick_threads(sample_n, cores)
    os.environ.setdefault("OMP_PROC_BIND", "close")
    os.environ.setdefault("OMP_PLACES", "cores")
    state = {}
    for _ in range(max(1, min(args.warmup, 1))):  # each sample is seconds of CPU work; one warm-up is enough
        cpu_reference_step(sample_n, nthr, state)
    t = [cpu_reference_step(sample_n, nthr, state) for _ in range(steps)]
    total = sum(t)
    value = 2.0 * sample_n ** 3 * steps / total / 1e12
    sample = (f"each step = reference kernel19 on M=N=K={sample_n} (a {sample_n}^3 sub-problem of the {size}^3 "
              f"workload; kernel19 cannot take N > 9216, SURVEY.md F2), {nthr} threads on {cores} host cores "
              f"({cpu_model()})")
    line = {"impl": "reference", "metric": "fp64_dgemm_tflops", "value": value, "unit": UNIT, "n_gpus": N,
            "steps": steps, "warmup": 1, "ms_per_step": 1e3 * total / steps, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"M=N=K={size} column-major DGEMM alpha=1 beta=1", "sample": sample},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": nthr, "kind": state["kind"], "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit(line)


# ------------------------------------------------------------------------------------------------------------------
# the B200 arm
# ------------------------------------------------------------------------------------------------------------------
_REAL_STDOUT = None


def claim_stdout():
    """Rank 0 must print exactly ONE line on stdout.  Libraries write there too (NCCL prints its version banner on
    stdout when NCCL_DEBUG=VERSION is set on the box), so file descriptor 1 is pointed at stderr for the whole run and
    the JSON line goes to a private duplicate of the original stdout."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


def main():
    args = parse()
    claim_stdout()
    if args.impl == "reference":
        reference_arm(args)
        return

    import torch
    import torch.distributed as dist
    import dgemm_b200 as dg
    from dgemm_b200.sharding import partition, sharded_dgemm_step, sharded_dgemm_step_fused

    N = args.gpus
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if N > 1 and world != N:
        raise SystemExit(f"--gpus {N} needs torchrun with {N} ranks (WORLD_SIZE={world})")
    if not torch.cuda.is_available() or dg.device_count() < 1:
        raise SystemExit("bench.py: no B200 visible — the DGEMM path is CUDA-only (no CPU fallback)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if N > 1:
        dist.init_process_group("nccl", device_id=dev)

    size = args.size or (16384 if N == 1 else 65536)
    M = Nn = K = size
    alpha, beta = 1.0, 1.0
    steps = args.steps or (10 if size <= 16384 else 3)
    warmup = max(args.warmup, 0)
    stream = torch.cuda.current_stream()
    sptr = stream.cuda_stream

    # ---- operands resident in HBM: every rank owns a column panel of B and C and a K-slab of A ----------------
    nb = partition(Nn, N, 128)
    kb = partition(K, N, 16)
    n0, n1 = nb[rank], nb[rank + 1]
    ncols = n1 - n0
    fused = N > 1 and args.exchange == "fused"
    ctx = {"peers": {}}
    if fused:
        # A lives in a plain cudaMalloc buffer (IPC-exportable); peers map it and the GEMM kernel pulls from it
        A_ptr = dg.malloc(8 * M * K)
        handles = [None] * N
        dist.all_gather_object(handles, dg.ipc_export(A_ptr))
        ctx["peers"] = {g: dg.ipc_open(handles[g]) for g in range(N) if g != rank}
        A = None
    else:
        A = torch.empty(M * K, dtype=torch.float64, device=dev)      # full A (M x K, ld = M) on every rank
        A_ptr = A.data_ptr()
    B = torch.empty(K * max(ncols, 1), dtype=torch.float64, device=dev)
    Cm = torch.empty(M * max(ncols, 1), dtype=torch.float64, device=dev)
    k0, k1 = kb[rank], kb[rank + 1]
    # only the rank's own K-slab of A is initialised locally; the rest arrives over NVLink inside each step
    dg.fill_signed(A_ptr + 8 * k0 * M, M, k1 - k0, M, 1, 0, k0, sptr)
    dg.fill_signed(B.data_ptr(), K, ncols, K, 2, 0, n0, sptr)
    dg.fill_signed(Cm.data_ptr(), M, ncols, M, 3, 0, n0, sptr)
    torch.cuda.synchronize()
    if N > 1:
        dist.barrier()

    def gemm_fetch(kbeg, kc, b_eff, fetch):
        src = dst = nbytes = 0
        if fetch is not None:
            owner, fk, fc = fetch
            src, dst, nbytes = ctx["peers"][owner] + 8 * fk * M, A_ptr + 8 * fk * M, 8 * fc * M
        dg.dgemm_device_fetch(M, ncols, kc, alpha, A_ptr + 8 * kbeg * M, M, B.data_ptr() + 8 * kbeg, K, b_eff,
                              Cm.data_ptr(), M, sptr, src, dst, nbytes)

    def step():
        """One pass over this rank's panel. N > 1: the own K-slab of A needs no transfer and goes first; every GEMM
        call pulls the next slab over NVLink with its spare warps (fused) or waits for its ncclBroadcast (nccl)."""
        if N == 1:
            dg.dgemm_device(M, Nn, K, alpha, A_ptr, M, B.data_ptr(), K, beta, Cm.data_ptr(), M, sptr)
        elif fused:
            sharded_dgemm_step_fused(rank, N, kb, beta, gemm_fetch)
        else:
            sharded_dgemm_step(
                rank, N, kb, beta,
                lambda g: dist.broadcast(A[kb[g] * M: kb[g + 1] * M], src=g, async_op=True),
                lambda work: work.wait(),  # stream-level wait on the NCCL stream
                lambda kbeg, kc, b_eff: dg.dgemm_device(M, ncols, kc, alpha, A_ptr + 8 * kbeg * M, M,
                                                        B.data_ptr() + 8 * kbeg, K, b_eff, Cm.data_ptr(), M, sptr))

    def barrier():
        if N > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(warmup):
        step()
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
        time.sleep(0.15)
    dg.reset_launch_count()
    barrier()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
    ev[0].record(stream)
    for i in range(steps):
        step()
        ev[i + 1].record(stream)
    barrier()
    launches = dg.launch_count()
    clocks = sampler.stop() if rank == 0 else None
    # spot check of the timed path: after n passes with alpha = beta = 1, C = C0 + n * A*B exactly up to rounding
    import numpy as np
    n_pass = warmup + steps
    rng = np.random.default_rng(99 + rank)
    idx = [(int(rng.integers(M)), int(rng.integers(ncols))) for _ in range(3)] + [(M - 1, ncols - 1)]
    got = [float(Cm[j * M + i].item()) for i, j in idx]
    c0 = [float(gen_signed(3, np.array([i]), np.array([n0 + j]))[0]) for i, j in idx]
    # n passes of (C <- A*B + C): same as one pass with alpha = n on the original C (the bound scales with alpha)
    value_check = spot_check(c0, got, idx, M, K, n0, float(n_pass), 1.0)
    if N > 1:
        vc = torch.tensor([value_check], dtype=torch.float64, device=dev)
        dist.all_reduce(vc, op=dist.ReduceOp.MAX)
        value_check = float(vc.item())
    if not value_check <= 1.0:
        raise SystemExit(f"bench.py: timed result is WRONG: max |err|/bound = {value_check}")
    total_ms = ev[0].elapsed_time(ev[steps])
    step_ms = [ev[i].elapsed_time(ev[i + 1]) for i in range(steps)]
    if N > 1:
        t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms = float(t.item())
        lt = torch.tensor([launches], dtype=torch.int64, device=dev)
        dist.all_reduce(lt, op=dist.ReduceOp.SUM)
        launches = int(lt.item())
    flops = 2.0 * M * Nn * K
    value = flops * steps / (total_ms * 1e-3) / 1e12

    # ---- end to end through the reference-facing host-pointer entry (pinned host buffers) -----------------------
    e2e = None
    if not args.no_e2e:
        e2e = run_e2e(torch, dist, dg, N, rank, dev, M, Nn, K, alpha, beta, nb, kb, A_ptr, A, B, Cm,
                      max(2, min(steps, 5)), fused, step, ctx["peers"])

    if rank != 0:
        if N > 1:
            dist.barrier()
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel (dgemm_tma_ws_kernel): FP64 tensor pipe -------------------------------
    kernels_per_step = max(1, launches // (steps * N))
    kern_ms = statistics.mean(step_ms) / 1.0  # one rank's step; at N = 1 a step is exactly one kernel launch
    achieved = (flops / N) / (kern_ms * 1e-3) / 1e12 if N == 1 else value / N
    sm_mhz = (clocks or {}).get("sm_mhz") or 1965.0
    n_sms = torch.cuda.get_device_properties(dev).multi_processor_count
    peak = n_sms * FLOP_PER_CLK_PER_SM * sm_mhz * 1e6 / 1e12
    probe = load_profile_json("profiles/fp64_peak_probe.json")
    traffic = load_profile_json("profiles/ncu_traffic.json")
    roofline = {"bound": "tensor", "pipe": "fp64 DMMA (mma.sync.m8n8k4.f64)", "achieved": achieved, "peak": peak,
                "unit": UNIT, "frac": achieved / peak,
                "peak_source": f"{n_sms} SMs x 128 flop/clk/SM x {sm_mhz:.0f} MHz sampled by nvidia-smi during the "
                               "timed region (MEASURED_PEAKS.json has no FP64 entry; csrc/peak_probe.cu measures "
                               "127.5 flop/clk/SM with DMMA alone)",
                "dmma_probe_tflops": (probe or {}).get("dmma_tflops"),
                "frac_of_dmma_probe": (achieved / probe["dmma_tflops"]) if probe and probe.get("dmma_tflops") else None,
                "traffic": (traffic or {}).get(str(size)),
                "algorithmic_bytes": 8.0 * (M * K + K * Nn + 2 * M * Nn),
                "kernel": "dg::dgemm_tma_ws_kernel<6>", "kernels_per_step": kernels_per_step}

    cpu_baseline = None
    if not args.no_cpu_baseline:
        try:
            cpu_baseline = run_cpu_baseline(8192 if (os.cpu_count() or 1) >= 8 else 4096, 2)
        except Exception as e:  # noqa: BLE001 - the baseline is a report, never a reason to lose the GPU number
            cpu_baseline = {"value": None, "unit": UNIT, "cores": 0, "kind": "unavailable", "sample": repr(e)}

    cublas = None
    if not args.no_cublas and N == 1:
        cublas = cublas_ceiling(torch, A, B, Cm, M, Nn, K)

    line = {
        "metric": "fp64_dgemm_tflops", "value": value, "unit": UNIT, "n_gpus": N, "steps": steps, "warmup": warmup,
        "ms_per_step": total_ms / steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"M=N=K={size} column-major DGEMM C=alpha*A*B+beta*C, alpha=1 beta=1, LD=dim",
                   "sharding": "single GPU" if N == 1 else
                   f"column panels of B/C over {N} GPUs; K-slabs of A replicated over NVLink inside the step, " +
                   ("pulled from peer HBM by spare warps of the DGEMM kernel (fused)" if fused else
                    "one ncclBroadcast per slab (baseline)"),
                   "l2": "operands (3 x %.1f GB) far exceed the 126 MB L2; no flush needed" % (8.0 * size * size / 1e9),
                   "inputs": "U(-1,1) 53-bit counter-based, generated on device (dgemm_b200_fill_signed)",
                   "timing": "CUDA events on the launching stream, max over ranks"},
        "clocks": clocks, "e2e": e2e, "gpu_launches": launches, "roofline": roofline, "cpu_baseline": cpu_baseline,
        "cublas_ceiling": cublas,
        "spot_check_max_err_over_bound": value_check,
        "pct_of_fp64_peak": 100.0 * value / (N * peak),
    }
    emit(line)
    if N > 1:
        dist.barrier()
        dist.destroy_process_group()


def gen_signed(seed, rows, cols):
    """numpy copy of the counter-based input generator (dgemm_b200_fill_signed): entry (i, j) = f(seed, i, j).
    rows / cols are integer arrays of equal length; returns float64 values.  Used only to spot-check bench results."""
    import numpy as np
    with np.errstate(over="ignore"):
        x = np.uint64(seed) * np.uint64(0xD1342543DE82EF95) + (rows.astype(np.uint64) + (cols.astype(np.uint64) << np.uint64(32)))
        x = x + np.uint64(0x9E3779B97F4A7C15)
        x = (x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        x = (x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        x = x ^ (x >> np.uint64(31))
    return (x >> np.uint64(11)).astype(np.int64).astype(np.float64) * (1.0 / 4503599627370496.0) - 1.0


def spot_check(before, after, idx, M, K, n0, alpha, beta):
    """after[i,j] must equal alpha * A[i,:].B[:,j] + beta * before[i,j] within the north-star bound; A and B rows /
    columns are regenerated on the host from the counter-based generator.  Returns max |err| / bound."""
    import numpy as np
    worst = 0.0
    kk = np.arange(K, dtype=np.int64)
    for (i, j), b, a in zip(idx, before, after):
        arow = gen_signed(1, np.full(K, i, dtype=np.int64), kk).astype(np.longdouble)
        bcol = gen_signed(2, kk, np.full(K, n0 + j, dtype=np.int64)).astype(np.longdouble)
        dot = float(np.sum(arow * bcol))
        mag = float(np.sum(np.abs(arow * bcol)))
        want = alpha * dot + beta * b
        bound = 2.0 * K * 2.220446049250313e-16 * (abs(alpha) * mag + abs(beta * b))
        worst = max(worst, abs(a - want) / bound)
    return worst


def load_profile_json(rel):
    try:
        with open(os.path.join(ROOT, rel)) as f:
            return json.load(f)
    except (OSError, ValueError):
        return None


def cublas_ceiling(torch, A, B, Cm, M, Nn, K):
    """cuBLAS cublasDgemm on the same device buffers — ceiling reference only, never on the product path."""
    a = A.view(K, M).t()   # column-major M x K == row-major K x M transposed
    b = B.view(Nn, K).t()
    for _ in range(2):
        c = a @ b
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
        c = a @ b
    e1.record()
    torch.cuda.synchronize()
    del c
    ms = e0.elapsed_time(e1) / 3
    return {"tflops": 2.0 * M * Nn * K / ms / 1e9, "ms": ms, "note": "torch.matmul(float64) -> cublasDgemm; ceiling only"}


def run_e2e(torch, dist, dg, N, rank, dev, M, Nn, K, alpha, beta, nb, kb, A_ptr, A, B, Cm, reps, fused, step, peers):
    """Same metric through the public host-pointer API with HOST (pinned) buffers, copies inside the timed region.
    N = 1: mydgemm_gpu_v20 (the reference's 11-argument signature).  N > 1: each rank stages its own K-slab of A and
    its B/C panels from pinned memory, runs the sharded step (A exchanged over NVLink), copies its C panel back."""
    n0, n1 = nb[rank], nb[rank + 1]
    ncols = n1 - n0
    k0, k1 = kb[rank], kb[rank + 1]
    hA = torch.empty(M * (k1 - k0), dtype=torch.float64).pin_memory()
    hB = torch.empty(K * ncols, dtype=torch.float64).pin_memory()
    hC = torch.empty(M * ncols, dtype=torch.float64).pin_memory()
    dg._check(dg.lib.dgemm_b200_d2h(hA.data_ptr(), M, A_ptr + 8 * k0 * M, M, M, k1 - k0, None))
    dg._check(dg.lib.dgemm_b200_sync(None))
    hB.copy_(B[: K * ncols])
    hC.copy_(Cm[: M * ncols])
    torch.cuda.synchronize()
    h2d = 8 * (hA.numel() + hB.numel() + hC.numel())
    d2h = 8 * hC.numel()
    stream = torch.cuda.current_stream()
    sptr = stream.cuda_stream
    token = torch.zeros(1, device=dev)

    def one_simple():
        dg._check(dg.lib.dgemm_b200_h2d(A_ptr + 8 * k0 * M, M, hA.data_ptr(), M, M, k1 - k0, sptr))
        dist.all_reduce(token)  # stream-ordered barrier: every owner's K-slab is in its HBM before anyone pulls it
        dg._check(dg.lib.dgemm_b200_h2d(B.data_ptr(), K, hB.data_ptr(), K, K, ncols, sptr))
        dg._check(dg.lib.dgemm_b200_h2d(Cm.data_ptr(), M, hC.data_ptr(), M, M, ncols, sptr))
        step()
        dist.all_reduce(token)  # ... and nobody still reads a slab when its owner overwrites it in the next pass
        dg._check(dg.lib.dgemm_b200_d2h(hC.data_ptr(), M, Cm.data_ptr(), M, M, ncols, sptr))
        stream.synchronize()

    copy_s, d2h_s = torch.cuda.Stream(), torch.cuda.Stream()
    from dgemm_b200.sharding import partition, slab_order
    # column sub-panels of this rank's panel: four main pieces and two narrow last ones (short exposed last D2H)
    tails = [384, 128] if ncols >= 4096 else []
    sp = partition(ncols - sum(tails), 4 if ncols - sum(tails) >= 4 * 256 else 1, 128)
    for t in tails:
        sp.append(sp[-1] + t)
    SP = len(sp) - 1
    order = [g for g in slab_order(rank, N) if kb[g + 1] > kb[g]] if N > 1 else []
    # the own K-slab goes in as sub-slabs so that the first call waits for a quarter of it
    Q = 4 if (k1 - k0) >= 4 * 512 else 1
    qb = [k0 + x for x in partition(k1 - k0, Q, 16)]
    X = torch.empty(M * max(ncols, 1), dtype=torch.float64, device=dev) if (N > 1 and beta != 0.0) else None

    def one_pipelined():
        """Rank-local pipeline on three streams.  Compute: (K-slab, column sub-panel) calls in slab-major order, the
        own slab first and in Q sub-slabs.  H2D, in need-order: own A sub-slabs with their B blocks, the B blocks of
        the other slabs, and LAST the C sub-panels — C is deferred: every sub-panel starts with beta = 0 and gets
        beta*C_in added (dgemm_b200_add_scaled) after its last slab, so the start of the pipeline waits for a quarter
        slab of A and one B block only.  Every C sub-panel goes back as soon as it is finished.  The NVLink pull of
        slab i+1 rides on a call of slab i (the LAST call of the own slab for the first pull: by then every owner's
        slab has reached its HBM, which the stream-ordered all_reduce in front of that call guarantees)."""
        cptr = copy_s.cuda_stream
        copy_s.wait_stream(stream)
        ev_b = {}
        for q in range(Q):
            ka, kc = qb[q], qb[q + 1] - qb[q]
            if kc <= 0:
                continue
            dg._check(dg.lib.dgemm_b200_h2d(A_ptr + 8 * ka * M, M, hA.data_ptr() + 8 * (ka - k0) * M, M, M, kc, cptr))
            for j in range(SP):
                c0, nc = sp[j], sp[j + 1] - sp[j]
                off = 8 * (c0 * K + ka)
                dg._check(dg.lib.dgemm_b200_h2d(B.data_ptr() + off, K, hB.data_ptr() + off, K, kc, nc, cptr))
                ev_b[(0, q, j)] = copy_s.record_event()
        ev_a = copy_s.record_event()
        for i, g in enumerate(order):
            if i == 0:
                continue
            for j in range(SP):
                c0, nc = sp[j], sp[j + 1] - sp[j]
                off = 8 * (c0 * K + kb[g])
                dg._check(dg.lib.dgemm_b200_h2d(B.data_ptr() + off, K, hB.data_ptr() + off, K, kb[g + 1] - kb[g], nc, cptr))
                ev_b[(i, 0, j)] = copy_s.record_event()
        ev_x = []
        if X is not None:
            for j in range(SP):
                c0, nc = sp[j], sp[j + 1] - sp[j]
                dg._check(dg.lib.dgemm_b200_h2d(X.data_ptr() + 8 * c0 * M, M, hC.data_ptr() + 8 * c0 * M, M, M, nc, cptr))
                ev_x.append(copy_s.record_event())
        last_i = len(order) - 1
        js = [j for j in range(SP) if sp[j + 1] > sp[j]]
        for i, g in enumerate(order):
            subs = [(qb[q], qb[q + 1] - qb[q], q) for q in range(Q)] if i == 0 else [(kb[g], kb[g + 1] - kb[g], 0)]
            subs = [t for t in subs if t[1] > 0]
            for si, (ka, kc, q) in enumerate(subs):
                for j in js:
                    c0, nc = sp[j], sp[j + 1] - sp[j]
                    stream.wait_event(ev_b[(i, q, j)])
                    src = dst = nbytes = 0
                    if i + 1 < len(order):
                        first_pull = i == 0 and si == len(subs) - 1 and j == js[-1]
                        if first_pull:
                            stream.wait_event(ev_a)
                            dist.all_reduce(token)  # every owner's K-slab is in its HBM before anyone pulls it
                        if first_pull or (i > 0 and j == js[0]):
                            nx = order[i + 1]
                            src, dst = peers[nx] + 8 * kb[nx] * M, A_ptr + 8 * kb[nx] * M
                            nbytes = 8 * (kb[nx + 1] - kb[nx]) * M
                    first_call = i == 0 and si == 0
                    b_eff = (0.0 if X is not None else beta) if first_call else 1.0
                    dg.dgemm_device_fetch(M, nc, kc, alpha, A_ptr + 8 * ka * M, M, B.data_ptr() + 8 * (c0 * K + ka), K,
                                          b_eff, Cm.data_ptr() + 8 * c0 * M, M, sptr, src, dst, nbytes)
                    if i == last_i and si == len(subs) - 1:
                        if X is not None:
                            stream.wait_event(ev_x[j])
                            dg.add_scaled(M, nc, beta, X.data_ptr() + 8 * c0 * M, M, Cm.data_ptr() + 8 * c0 * M, M, sptr)
                        d2h_s.wait_event(stream.record_event())
                        dg._check(dg.lib.dgemm_b200_d2h(hC.data_ptr() + 8 * c0 * M, M, Cm.data_ptr() + 8 * c0 * M, M, M, nc,
                                                        d2h_s.cuda_stream))
        if len(order) == 1:
            stream.wait_event(ev_a)
        dist.all_reduce(token)  # nobody still reads a slab when its owner overwrites it in the next pass
        d2h_s.synchronize()
        stream.synchronize()

    def one():
        if N == 1:
            dg.mydgemm_gpu_v20(M, Nn, K, alpha, hA.data_ptr(), M, hB.data_ptr(), K, beta, hC.data_ptr(), M)
        elif fused and os.environ.get("BENCH_E2E_SIMPLE") != "1":
            one_pipelined()
        else:
            one_simple()

    one()  # warm-up (allocates the library's staging buffers)
    # spot check of the end-to-end path itself: one more pass, a few entries recomputed on the host
    import numpy as np
    rng = np.random.default_rng(1234 + rank)
    idx = [(int(rng.integers(M)), int(rng.integers(ncols))) for _ in range(6)] + [(M - 1, ncols - 1), (0, 0)]
    before = [float(hC[j * M + i]) for i, j in idx]
    one()
    after = [float(hC[j * M + i]) for i, j in idx]
    check = spot_check(before, after, idx, M, K, n0, alpha, beta)
    if N > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        one()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    if N > 1:
        t = torch.tensor([dt], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt = float(t.item())
        b = torch.tensor([h2d, d2h], dtype=torch.int64, device=dev)
        dist.all_reduce(b, op=dist.ReduceOp.SUM)
        h2d, d2h = int(b[0].item()), int(b[1].item())
        ck = torch.tensor([check], dtype=torch.float64, device=dev)
        dist.all_reduce(ck, op=dist.ReduceOp.MAX)
        check = float(ck.item())
    if not check <= 1.0:
        raise SystemExit(f"bench.py: end-to-end result is WRONG: max |err|/bound = {check}")
    return {"value": 2.0 * M * Nn * K * reps / dt / 1e12, "spot_check_max_err_over_bound": check,
            "spot_check": f"{len(idx)} entries per rank recomputed on the host in long double", "unit": UNIT, "h2d_bytes_per_step": h2d,
            "d2h_bytes_per_step": d2h, "ms_per_step": 1e3 * dt / reps, "reps": reps,
            "api": "mydgemm_gpu_v20(M,N,K,alpha,A,LDA,B,LDB,beta,C,LDC) with pinned host pointers" if N == 1 else
                   "per rank: dgemm_b200_h2d (own K-slab of A, B/C panels) + sharded step ("
                   + ("dgemm_b200_device_fetch, fused NVLink pull; (slab, sub-panel) calls pipelined against H2D/D2H "
                      "on three streams, own slab in sub-slabs, C deferred (dgemm_b200_add_scaled after the last slab)"
                      if fused else "ncclBroadcast + dgemm_b200_device")
                   + ") + dgemm_b200_d2h, pinned host buffers"}


if __name__ == "__main__":
    main()
