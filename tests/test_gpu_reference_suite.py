"""The reference's own Python tests, unmodified, against this library (plan item 3).

They ``import psvWave`` and open ``tests/test_configurations/default_testing_configuration.ini`` relative to the working
directory (tests/python/test_copy.py:1-24, test_descent.py:1-137 of larsgeb/psvWave).  This repository ships that import
name (psvWave/, __psvWave_cpp.py) and that path, so the files run as they are with the repository root as cwd.  Only the
INVOCATION lives here: the files are taken from a reference checkout -- ``$PSVWAVE_REFERENCE_TESTS`` or
/root/reference/tests/python -- and the test is skipped where there is none (the GPU boxes of this project have no
checkout; tests/test_gpu_api.py holds the same checks re-expressed, which run everywhere).
"""
import os
import subprocess
import sys

import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu

REFERENCE_TESTS = os.environ.get("PSVWAVE_REFERENCE_TESTS", "/root/reference/tests/python")


def test_import_name_and_relative_configuration_path(monkeypatch):
    """What those files rely on, checked everywhere: the import name, the top-level extension name, the relative path."""
    monkeypatch.chdir(ROOT)
    import __psvWave_cpp
    import psvWave
    import psvwave_b200
    assert psvWave.fdModel is psvwave_b200.fdModel is __psvWave_cpp.fdModel
    model = psvWave.fdModel("tests/test_configurations/default_testing_configuration.ini")
    twin = model.copy()
    model.set_model_vector(model.get_model_vector() + 1)
    assert (twin.get_model_vector() != model.get_model_vector()).any()


@pytest.mark.skipif(not os.path.isdir(REFERENCE_TESTS), reason="no reference checkout on this machine")
@pytest.mark.parametrize("name", ["test_copy.py", "test_descent.py"])
def test_reference_python_test_file_runs_unmodified(name):
    env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""))
    r = subprocess.run([sys.executable, "-m", "pytest", "-q", "-p", "no:cacheprovider", os.path.join(REFERENCE_TESTS, name)],
                       cwd=ROOT, env=env, capture_output=True, text=True, timeout=1800)
    assert r.returncode == 0, r.stdout[-4000:] + r.stderr[-2000:]
