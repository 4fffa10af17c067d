import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(autouse=True)
def _fresh_library_env(request):
    """The library reads its DGEMM_B200_* switches once; a test that flips one with monkeypatch calls dg.reload_env() itself,
    and this fixture (set up before monkeypatch, so torn down after its undo) re-reads the restored environment afterwards."""
    yield
    if "monkeypatch" in request.fixturenames:
        mod = sys.modules.get("dgemm_b200")
        if mod is not None:
            mod.reload_env()


@pytest.fixture(scope="session")
def oracle():
    import oracle_lib
    oracle_lib.build_oracle() if not os.path.exists(os.path.join(ROOT, "oracle", "liboracle.so")) else None
    return oracle_lib


@pytest.fixture(scope="session")
def dg():
    """The product library through its Python/ctypes mirror. Import fails loudly if the .so is missing."""
    import dgemm_b200
    return dgemm_b200
