"""CPU: the C-ABI library loads, exports every symbol include/*.h declares, and fails LOUDLY without a GPU
(no CPU fallback, no routing through the oracle)."""
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "optimizing-dgemm-on-intel-cpus-with-avx512f_b200")


def declared_functions():
    text = open(os.path.join(ROOT, "include", "dgemm_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(mydgemm_gpu_v2[01]|dgemm_b200_[a-z0-9_]+)\s*\(", text)))


def test_header_and_binding_agree(dg):
    assert declared_functions() == sorted(dg.EXPORTS)


def test_library_exports_every_declared_symbol(dg):
    out = subprocess.run(["nm", "-D", "--defined-only", dg.LIB_PATH], capture_output=True, text=True, check=True).stdout
    exported = set(line.split()[-1] for line in out.splitlines() if " T " in line)
    for name in declared_functions():
        assert name in exported, f"{name} declared in include/dgemm_b200.h but not exported"
        assert hasattr(dg.lib, name)


def test_library_is_sm100a_cuda_not_cpu(dg):
    """The product .so carries sm_100a SASS with the FP64 tensor-core and TMA instructions."""
    sass = subprocess.run(["cuobjdump", "-sass", dg.LIB_PATH], capture_output=True, text=True).stdout
    if not sass:
        pytest.skip("cuobjdump not available")
    assert "sm_100a" in sass
    assert "DMMA.8x8x4" in sass        # mma.sync.m8n8k4.f64
    assert "UTMALDG" in sass           # cp.async.bulk.tensor (TMA)
    assert "LDGSTS" in sass            # cp.async loader for TMA-ineligible operands
    assert "SYNCS" in sass             # mbarrier
    assert "UBLKCP" in sass            # cp.async.bulk: stream-K partials come in through the TMA ring
    assert "dgemm_tma_ws_split_kernel" in sass   # parity-split variants for odd LD / 8-byte-aligned operands
    for ns in ("N2dg", "N4dg64", "N6dg64x2", "N8dg128x64", "N7dg64x32"):   # every tile build of the kernel body is in the library
        assert f"_Z{ns}19dgemm_tma_ws_kernel" in sass, ns
    assert "dgemm_tma_cluster_kernel" in sass      # ladder rung 6: 2-CTA clusters + TMA multicast
    assert "UTMALDG.2D.MULTICAST" in sass or "MULTICAST" in sass


def test_product_does_not_link_the_oracle(dg):
    out = subprocess.run(["ldd", dg.LIB_PATH], capture_output=True, text=True).stdout
    assert "oracle" not in out and "libref" not in out
    for dirpath, _, files in os.walk(PKG):
        for f in files:
            if f.endswith((".cu", ".cuh", ".h", ".c", ".py")):
                src = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "oracle_lib" not in src and "liboracle" not in src and "libref" not in src, f


def test_kernel20_header_is_plain_c():
    """include/kernel20.h + include/dgemm_b200.h compile as C (what a kernels.h-style aggregator needs)."""
    src = '#include "kernel20.h"\nint main(void){return (int)(sizeof(&test_mydgemm_v20)+sizeof(&test_mydgemm_v21)) == 0;}\n'
    r = subprocess.run(["/usr/bin/gcc", "-std=c99", "-Wall", "-Werror", "-fsyntax-only", "-I", os.path.join(ROOT, "include"),
                        "-x", "c", "-"], input=src, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr


def test_no_gpu_means_loud_failure(dg):
    if dg.device_count() > 0:
        pytest.skip("a B200 is visible; the no-device path cannot be exercised here")
    A = np.asfortranarray(np.ones((4, 4)))
    C0 = np.asfortranarray(np.full((4, 4), 3.0))
    with pytest.raises(dg.DgemmError) as e:
        dg.mydgemm_gpu_v20(4, 4, 4, 1.0, A, 4, A, 4, 1.0, C0, 4)
    assert e.value.status == dg.ERR_NO_DEVICE
    assert np.all(C0 == 3.0), "C must be left untouched when the call fails"
    with pytest.raises(dg.DgemmError):
        dg.mydgemm_gpu_v21(4, 4, 4, 1.0, A, 4, A, 4, 1.0, C0, 4)
    with pytest.raises(dg.DgemmError):
        dg.dgemm_device(4, 4, 4, 1.0, 16, 4, 16, 4, 1.0, 16, 4)
    assert dg.lib.dgemm_b200_malloc(64) is None


def test_bad_arguments_are_rejected_before_touching_a_device(dg):
    A = np.asfortranarray(np.ones((4, 4)))
    for args in [(-1, 4, 4, 4, 4, 4), (4, 4, 4, 3, 4, 4), (4, 4, 4, 4, 3, 4), (4, 4, 4, 4, 4, 3)]:
        M, N, K, lda, ldb, ldc = args
        with pytest.raises(dg.DgemmError) as e:
            dg.mydgemm_gpu_v20(M, N, K, 1.0, A, lda, A, ldb, 1.0, A, ldc)
        assert e.value.status in (dg.ERR_BAD_ARG, dg.ERR_NO_DEVICE)


def test_harness_binary_refuses_without_gpu(dg):
    exe = os.path.join(PKG, "bin", "dgemm_b200")
    if not os.path.exists(exe):
        pytest.skip("harness not built")
    r = subprocess.run([exe], capture_output=True, text=True)
    assert r.returncode != 0 and "select a kernel" in r.stdout          # reference test.c:8-11 behaviour
    r = subprocess.run([exe, "7"], capture_output=True, text=True)
    assert r.returncode != 0 and "valid kernel number" in r.stdout      # reference test.c:16-19 behaviour
    if dg.device_count() == 0:
        r = subprocess.run([exe, "20", "100"], capture_output=True, text=True)
        assert r.returncode != 0 and "no CPU fallback" in r.stdout


def test_cblas_entry_validates_codes_without_a_device(dg):
    """Layout / transpose codes are checked before the device is touched; a valid call without a GPU fails loudly."""
    A = np.zeros((4, 4), order="F"); B = np.zeros((4, 4), order="F"); Cm = np.ones((4, 4), order="F")
    with pytest.raises(dg.DgemmError) as e:
        dg.cblas_dgemm(7, dg.NO_TRANS, dg.NO_TRANS, 4, 4, 4, 1.0, A, 4, B, 4, 0.0, Cm, 4)
    assert e.value.status == dg.ERR_BAD_ARG
    if dg.device_count() == 0:
        with pytest.raises(dg.DgemmError) as e:
            dg.cblas_dgemm(dg.ROW_MAJOR, dg.TRANS, dg.NO_TRANS, 4, 4, 4, 1.0, A, 4, B, 4, 0.0, Cm, 4)
        assert e.value.status == dg.ERR_NO_DEVICE
        assert np.all(Cm == 1.0)   # C untouched: no CPU fallback
