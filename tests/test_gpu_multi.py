"""-m gpu, needs >= 2 B200s (skipped otherwise): the column-panel sharded DGEMM under torchrun/NCCL, and the
single-process multi-GPU C-ABI entry mydgemm_gpu_v21."""
import os
import subprocess
import sys

import numpy as np
import pytest

import gpu_util as gu

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ngpu():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:  # noqa: BLE001
        return 0


@pytest.mark.skipif(_ngpu() < 2, reason="needs >= 2 GPUs")
@pytest.mark.parametrize("mode", ["fused", "nccl"])
@pytest.mark.parametrize("shape", [(1024, 2048, 1024), (1000, 1003, 997), (512, 640, 4096), (2048, 1024, 8)])
def test_torchrun_sharded_dgemm(shape, mode):
    M, N, K = shape
    n = min(_ngpu(), 4)
    port = 29700 + os.getpid() % 200
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr",
           "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "tests", "multi_gpu_worker.py"), str(M), str(N),
           str(K), "1.5", "0.5", mode]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "FAIL" not in r.stdout


@pytest.mark.skipif(_ngpu() < 2, reason="needs >= 2 GPUs")
@pytest.mark.parametrize("shape", [(1024, 2048, 1024), (1000, 1003, 997), (4096, 8192, 4096), (2048, 1024, 64)])
def test_torchrun_rank_pipeline_from_host_operands(shape):
    """dgemm_b200_rank_pipeline with REAL peers (tests/test_gpu_rank_pipeline.py covers it with emulated peers on one GPU)."""
    M, N, K = shape
    n = min(_ngpu(), 4)
    port = 29500 + os.getpid() % 200
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr",
           "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "tests", "multi_gpu_worker.py"), str(M), str(N),
           str(K), "1.5", "0.5", "pipeline"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "FAIL" not in r.stdout


@pytest.mark.skipif(_ngpu() < 2, reason="needs >= 2 GPUs")
@pytest.mark.parametrize("shape,pad,alpha,beta", [((1024, 2048, 1024), (0, 0, 0), 1.0, 1.0),
                                                  ((4096, 6144, 4096), (0, 0, 0), 1.0, 1.0),
                                                  ((1000, 1003, 997), (3, 1, 5), 2.5, 0.5),
                                                  ((300, 700, 0), (0, 0, 0), 1.0, 0.5),
                                                  ((512, 900, 333), (0, 0, 0), -1.0, 0.0)])
def test_v21_single_process_multi_gpu(dg, shape, pad, alpha, beta):
    M, N, K = shape
    A, B, C0 = gu.make_inputs(M, N, K, pad)
    Cg = C0.copy(order="F")
    dg.mydgemm_gpu_v21(M, N, K, alpha, A, A.shape[0], B, B.shape[0], beta, Cg, Cg.shape[0])
    gu.check(M, N, K, alpha, A, B, beta, C0, Cg)


@pytest.mark.skipif(_ngpu() < 2, reason="needs >= 2 GPUs")
def test_v21_nccl_exchange_mode(shape=(1024, 2048, 1024)):
    """The NCCL-broadcast exchange (baseline) stays selectable: run it in a fresh process with the env switch."""
    code = ("import sys; sys.path[:0]=[%r,%r]; import dgemm_b200 as dg, gpu_util as gu; "
            "A,B,C0=gu.make_inputs(1024,2048,1024); Cg=C0.copy(order='F'); "
            "dg.mydgemm_gpu_v21(1024,2048,1024,1.5,A,1024,B,1024,0.5,Cg,1024); "
            "print('ratio', gu.check(1024,2048,1024,1.5,A,B,0.5,C0,Cg))") % (ROOT, os.path.join(ROOT, "tests"))
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300,
                       env=dict(os.environ, DGEMM_B200_EXCHANGE="nccl"))
    assert r.returncode == 0 and "ratio" in r.stdout, r.stdout[-1500:] + r.stderr[-1500:]


def test_v21_falls_back_to_one_gpu_for_narrow_problems(dg):
    """N < 2 tiles: sharding columns is pointless; v21 must still compute (on one GPU), not fail."""
    A, B, C0 = gu.make_inputs(256, 100, 128)
    Cg = C0.copy(order="F")
    dg.mydgemm_gpu_v21(256, 100, 128, 1.0, A, 256, B, 128, 1.0, Cg, 256)
    gu.check(256, 100, 128, 1.0, A, B, 1.0, C0, Cg)
