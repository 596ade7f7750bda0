"""examples/c_api_determinism.c: the reference's C++ tests (tests/test_file_constructor.cpp:13-73,
tests/test_copy_constructor.cpp:14-69) re-expressed in plain C over the C ABI -- the binding a C / C++ user of fdModel
would write.  Without a GPU: it compiles with gcc against include/psvwave_b200.h, links against the shared library
alone and reports the missing device through the ABI's status code.  With a GPU: it runs, its determinism and copy checks
hold, and its number equals the sum the Python binding produces."""
import os
import subprocess

import numpy as np
import pytest

from conftest import ROOT, SMALL_INI


@pytest.fixture(scope="module")
def demo(tmp_path_factory):
    import psvwave_b200
    exe = str(tmp_path_factory.mktemp("c_example") / "c_api_determinism")
    libdir = os.path.dirname(psvwave_b200.LIBRARY_PATH)
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"),
                    os.path.join(ROOT, "examples", "c_api_determinism.c"), "-L", libdir, "-lpsvwave_b200",
                    f"-Wl,-rpath,{libdir}", "-o", exe], check=True)
    return exe


def test_c_example_builds_and_fails_loudly_without_a_device(demo):
    import psvwave_b200
    if psvwave_b200.device_count() > 0:
        pytest.skip("a CUDA device is present: covered by the GPU test")
    needed = subprocess.run(["ldd", demo], capture_output=True, text=True).stdout
    assert "libpsvwave_b200.so" in needed and "torch" not in needed and "python" not in needed
    r = subprocess.run([demo, SMALL_INI], capture_output=True, text=True)
    assert r.returncode == 2 and "no CPU fallback" in r.stderr.replace("There is no", "no")


@pytest.mark.gpu
def test_c_example_runs_and_matches_the_python_binding(demo, tmp_path):
    import psvwave_b200
    r = subprocess.run([demo, SMALL_INI], capture_output=True, text=True, cwd=tmp_path)
    assert r.returncode == 0, r.stderr
    value = float([l for l in r.stdout.splitlines() if l.startswith("sum ")][0].split()[1])
    m = psvwave_b200.fdModel(SMALL_INI)
    for s in range(m.n_shots):
        m.forward_simulate(s, False)
    ux = m.get_synthetic_data()[0]
    acc = 0.0
    for v in ux.ravel():  # the example's (and the reference test's) serial sum
        acc += float(v)
    assert value == acc and np.isfinite(value) and value != 0.0
