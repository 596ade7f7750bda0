"""SURVEY 8(f) rows 1 and 4 on the GPU: the per-shot gradient loop (psv_run_gradient) and the
device-resident steepest-descent loop (psv_descent_step / psv_gradient_descent), against the
reference's own call sequences run on the CPU oracle:
  * run_model (src/fdModel.cpp:1291-1308) for the gradient,
  * the loop of tests/python/test_descent.py:106-129 (m -= step * g; set_model_vector(m); forward;
    misfit; adjoint sources; reset; adjoint; map) for the descent.
"exact" arithmetic: bit-identical; default arithmetic: <= 1e-10 relative L2.
"""
import numpy as np
import pytest

from conftest import rel_l2
from helpers import gradient_workflow, ragged_config, random_medium
from oracle.oracle import OracleModel

pytestmark = pytest.mark.gpu
TOL = 1e-10


@pytest.fixture(scope="module")
def case(tmp_path_factory):
    ini = str(tmp_path_factory.mktemp("descent") / "ragged.ini")
    ragged_config(ini)
    true = OracleModel(ini)
    true.set_parameter_fields(*random_medium(61, 47, seed=21, amp=0.03))
    for s in range(true.n_shots):
        true.forward_simulate(s, False)
    obs = tuple(np.array(a) for a in true.get_synthetic_data())
    o = OracleModel(ini)
    o.set_observed_data(*obs)
    for s in range(o.n_shots):
        o.forward_simulate(s, True)
    gradient_workflow(o)
    g0, misfit0 = o.get_gradient_vector().copy(), o.misfit
    step = 1e-3 * np.abs(o.get_model_vector()).max() / np.abs(g0).max()
    m = o.get_model_vector()
    misfits, models = [], []
    for _ in range(3):  # tests/python/test_descent.py:106-129
        g = o.get_gradient_vector()
        m = m - step * g
        o.set_model_vector(m)
        models.append(m.copy())
        for s in range(o.n_shots):
            o.forward_simulate(s, True)
        gradient_workflow(o)
        misfits.append(o.misfit)
    # (the model re-read from the grid is the block MEAN of the filled value, not the running m itself)
    return dict(ini=ini, obs=obs, g0=g0, misfit0=misfit0, step=float(step), misfits=misfits,
                models=[o.get_model_vector().copy()], vp_end=np.array(o.get_parameter_fields()[0]),
                g_end=o.get_gradient_vector().copy())


def fd(ini, **options):
    import psvwave_b200
    m = psvwave_b200.fdModel(ini)
    for k, v in options.items():
        m._set_option(k, v)
    return m


@pytest.mark.parametrize("exact", [1, 0])
@pytest.mark.parametrize("release", [True, False])
def test_run_gradient_is_run_model(case, exact, release):
    m = fd(case["ini"], exact=exact)
    m.set_observed_data(*case["obs"])
    misfit = m.run_gradient(release)
    assert misfit == m.misfit
    if exact:
        assert np.array_equal(m.get_gradient_vector(), case["g0"])
        assert abs(misfit - case["misfit0"]) <= 1e-14 * case["misfit0"]  # summation order differs
    else:
        assert rel_l2(m.get_gradient_vector(), case["g0"]) <= TOL
        assert abs(misfit - case["misfit0"]) <= TOL * case["misfit0"]
    if release:
        with pytest.raises(RuntimeError):
            m.adjoint_simulate(0)  # the snapshots are gone
    else:
        m.adjoint_simulate(0)


def test_descent_step_equals_host_vector_update(case):
    a, b = fd(case["ini"], exact=1), fd(case["ini"], exact=1)
    for m in (a, b):
        m.set_parameter_fields(*random_medium(61, 47, seed=5))  # not block-constant: means matter
        m.set_observed_data(*case["obs"])
        m.run_gradient()
    a.descent_step(case["step"])
    v = b.get_model_vector()
    v -= case["step"] * b.get_gradient_vector()
    b.set_model_vector(v)
    for fa, fb in zip(a.get_parameter_fields(), b.get_parameter_fields()):
        assert np.array_equal(fa, fb)
    assert np.array_equal(a._get_array(5), b._get_array(5))  # lm follows (update_from_velocity)


@pytest.mark.parametrize("exact", [1, 0])
def test_gradient_descent_follows_the_reference_loop(case, exact):
    m = fd(case["ini"], exact=exact)
    m.set_observed_data(*case["obs"])
    m.run_gradient()
    misfits = m.gradient_descent(3, case["step"])
    assert misfits.shape == (3,)
    if exact:
        assert np.array_equal(m.get_model_vector(), case["models"][-1])
        assert np.array_equal(m.get_parameter_fields()[0], case["vp_end"])
        assert np.array_equal(m.get_gradient_vector(), case["g_end"])
        assert np.allclose(misfits, case["misfits"], rtol=1e-13, atol=0)
    else:
        assert rel_l2(m.get_model_vector(), case["models"][-1]) <= TOL
        assert rel_l2(m.get_gradient_vector(), case["g_end"]) <= TOL
        assert np.allclose(misfits, case["misfits"], rtol=TOL, atol=0)
    assert misfits[-1] < case["misfit0"]  # it descends
