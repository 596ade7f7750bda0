import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

# The suite pins the per-step kernels it names (variant = ...); the resident whole-loop kernel, which the default
# "auto" setting would pick on these small grids, has tests of its own that ask for it (tests/test_gpu_resident.py).
os.environ.setdefault("PSVWAVE_RESIDENT", "0")

GOLDEN = os.path.join(ROOT, "tests", "golden")
C1_INI = os.path.join(ROOT, "configs", "default_testing_configuration.ini")
SMALL_INI = os.path.join(GOLDEN, "small_multishot.ini")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_collection_modifyitems(config, items):
    # GPU tests are selected with `-m gpu`; when they are collected on a box without a
    # device (plain `pytest tests/`), skip them instead of failing.
    # A missing/broken extension is NOT a reason to skip: the GPU tests then fail loudly on import
    # (there is no CPU fallback to fall back to).
    try:
        import psvwave_b200
    except ImportError:
        return
    try:
        have_gpu = psvwave_b200.device_count() > 0
    except Exception:
        have_gpu = False
    if have_gpu:
        return
    skip = pytest.mark.skip(reason="no CUDA device on this machine")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


def rel_l2(a, b):
    """Relative L2 difference ||a-b|| / ||b|| (0 when both vanish)."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    nb = np.linalg.norm(b)
    d = np.linalg.norm(a - b)
    return 0.0 if d == 0.0 else d / (nb if nb > 0 else 1.0)


@pytest.fixture(scope="session")
def golden():
    def load(name):
        return np.load(os.path.join(GOLDEN, name + ".npz"))
    return load


@pytest.fixture(scope="session", autouse=True)
def _built_oracle():
    from oracle import oracle
    if not os.path.exists(oracle.ORC_LIB):
        oracle.build()
    return True


@pytest.fixture()
def in_tmp(tmp_path, monkeypatch):
    monkeypatch.chdir(tmp_path)
    return tmp_path
