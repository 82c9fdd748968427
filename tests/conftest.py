import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")
NORI_REF = os.path.join(ROOT, "oracle", "_ref", "nori_ref")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def nb():
    import norib200
    return norib200


@pytest.fixture(scope="session")
def built_lib(nb):
    """The in-tree libnorib200.so; built here if a checkout has none yet (nvcc cross-compiles)."""
    if not os.path.exists(nb.LIB_PATH):
        import importlib.util
        spec = importlib.util.spec_from_file_location("nb_build", os.path.join(ROOT, "optix-renderer_b200", "build.py"))
        m = importlib.util.module_from_spec(spec); spec.loader.exec_module(m)
        m.build_all()
    return nb.load_library()


@pytest.fixture(scope="session")
def _session_ctx(nb, built_lib):
    c = nb.Context(0)
    yield c
    c.close()


@pytest.fixture
def ctx(_session_ctx):
    """One device context for the whole session, options back to their defaults per test."""
    _session_ctx.set_option("strategy", 0)
    _session_ctx.set_option("profile", 0)
    _session_ctx.set_option("flat_scan", -1)
    _session_ctx.set_option("walk", 1)
    _session_ctx.set_option("ordered", 0)
    _session_ctx.set_option("wide", 1)
    _session_ctx.set_option("trace_wide", 0)
    _session_ctx.set_option("trace_fast", 0)
    return _session_ctx


def golden(name):
    return os.path.join(GOLDEN, name)


def have_oracle_binary():
    return os.path.exists(NORI_REF) and os.access(NORI_REF, os.X_OK)


def run_oracle(*args):
    """oracle/_ref/nori_ref (the reference's own CPU code) — test infrastructure only."""
    r = subprocess.run([NORI_REF] + list(args), stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    assert r.returncode == 0, r.stderr[-2000:]
    return r.stdout


def rel_rmse(a, b):
    """sqrt(mean((A-B)^2 / (B^2 + 1e-2)))"""
    a = np.asarray(a, dtype=np.float64); b = np.asarray(b, dtype=np.float64)
    return float(np.sqrt(np.mean((a - b) ** 2 / (b ** 2 + 1e-2))))
