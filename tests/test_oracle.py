"""Pin the oracle (oracle/psv_oracle.c, our C restatement):

1. against the committed golden vectors produced by the UNMODIFIED reference core
   (tests/golden/make_golden.py) -- bit for bit;
2. against the known answers printed in the reference's own notebook 2
   (notebooks/Notebook 2 ...ipynb:240 "Data misfit: 1920401.82", :443 "|proj g|= 7.71590D+01");
3. where oracle/_ref/libpsv_ref.so exists (dev container, or shipped prebuilt), directly
   against the reference on fresh ragged inputs -- bit for bit.
"""
import numpy as np
import pytest

from helpers import C1_INI, gradient_workflow, notebook2_models, ragged_config, random_medium
from oracle.oracle import OracleModel, RefModel, ref_available

from conftest import SMALL_INI


def test_small_multishot_bit_exact_vs_reference_golden(golden):
    g = golden("small_multishot")
    m = OracleModel(SMALL_INI)
    for name in ("taper", "stf", "moment"):
        assert np.array_equal(m.array(name), g[name]), name
    m.set_parameter_fields(g["vp"], g["vs"], g["rho"])
    for name in ("lm", "la", "mu", "b_vx"):
        assert np.array_equal(m.array(name), g[name]), name
    for s in range(m.n_shots):
        m.forward_simulate(s, True)
    ux, uz = m.get_synthetic_data()
    assert np.array_equal(ux, g["ux"]) and np.array_equal(uz, g["uz"])
    for f in ("vx", "vz", "txx", "tzz", "txz"):
        assert np.array_equal(m.array("accu_" + f)[:, [0, 17, 60]], g["snap_" + f]), f
        assert np.array_equal(m.array(f), g["fwd_end_" + f]), f
    assert np.array_equal(m.get_model_vector(), g["model_vector"])
    m.set_observed_data(g["obs_ux"], g["obs_uz"])
    gradient_workflow(m)
    assert m.misfit == float(g["misfit"])
    for name in ("a_stf_ux", "a_stf_uz", "lambda_kernel", "mu_kernel", "density_l_kernel", "vp_kernel",
                 "vs_kernel", "density_v_kernel"):
        assert np.array_equal(m.array(name), g[name]), name
    assert np.array_equal(m.get_gradient_vector(), g["gradient"])
    assert np.array_equal(m.array("vx"), g["adj_end_vx"])
    assert np.array_equal(m.array("txx"), g["adj_end_txx"])


@pytest.fixture(scope="module")
def c1_oracle_run():
    """Default configuration, homogeneous forward + notebook-2 gradient (about 10 s of CPU)."""
    m = OracleModel(C1_INI)
    m.forward_simulate(0, True)
    homog = dict(ux=m.get_synthetic_data()[0], uz=m.get_synthetic_data()[1],
                 snap_vx=m.array("accu_vx")[0, 400].copy(), snap_txz=m.array("accu_txz")[0, 400].copy(),
                 final_vx=m.array("vx").copy(), final_tzz=m.array("tzz").copy(),
                 taper=m.array("taper").copy(), stf=m.array("stf").copy(), moment=m.array("moment").copy(),
                 snapnorms={(f, k): np.linalg.norm(m.array("accu_" + f)[0, k])
                            for f in ("vx", "vz", "txx", "tzz", "txz") for k in (100, 400, 799)})
    start, target = notebook2_models(m)
    m.set_parameter_fields(*target)
    m.forward_simulate(0, True)
    obs = m.get_synthetic_data()
    m.set_observed_data(*obs)
    m.set_parameter_fields(*start)
    gradient_workflow(m)
    return homog, obs, m


def test_c1_homogeneous_vs_reference_golden(golden, c1_oracle_run):
    g = golden("c1_homogeneous")
    h = c1_oracle_run[0]
    assert np.array_equal(h["ux"][:, :, ::8], g["ux_dec8"])
    assert np.array_equal(h["uz"][:, :, ::8], g["uz_dec8"])
    assert h["ux"].sum() == float(g["sum_ux"]) and h["uz"].sum() == float(g["sum_uz"])
    assert np.linalg.norm(h["ux"]) == float(g["norm_ux"])
    assert np.array_equal(h["snap_vx"][::3, ::3], g["snap_vx_400_dec"])
    assert np.array_equal(h["snap_txz"][::3, ::3], g["snap_txz_400_dec"])
    assert np.array_equal(h["final_vx"][::3, ::3], g["final_vx_dec"])
    assert np.array_equal(h["final_tzz"][::3, ::3], g["final_tzz_dec"])
    assert np.array_equal(h["taper"][:, 75], g["taper_row"])
    assert np.array_equal(h["stf"][:, ::16], g["stf_dec"])
    assert np.array_equal(h["moment"], g["moment"])
    for (f, k), v in h["snapnorms"].items():
        assert v == float(g[f"snapnorm_{f}_{k}"])


def test_c1_survey_known_scalars(c1_oracle_run):
    # scalars probed from the compiled reference on the default ini
    h = c1_oracle_run[0]
    assert h["ux"].sum() == pytest.approx(-21.249427595741352, rel=1e-13)
    assert h["uz"].sum() == pytest.approx(50.68531309390501, rel=1e-13)
    assert np.linalg.norm(h["ux"]) == pytest.approx(150626.52785852677, rel=1e-14)
    assert np.linalg.norm(h["uz"]) == pytest.approx(99534.43907778655, rel=1e-14)


def test_c1_notebook2_known_answers(golden, c1_oracle_run):
    g = golden("c1_notebook2")
    _, obs, m = c1_oracle_run
    # what the reference's notebook prints
    assert f"{m.misfit:.2f}" == "1920401.82"
    grad = m.get_gradient_vector()
    assert f"{np.abs(grad).max():.5E}" == "7.71590E+01"
    # what the compiled reference gives, bit for bit
    assert m.misfit == float(g["misfit"])
    assert np.array_equal(grad[::7], g["g_dec7"])
    assert np.linalg.norm(grad) == float(g["g_norm"])
    assert np.array_equal(obs[0][:, :, ::8], g["obs_ux_dec8"])
    assert np.array_equal(obs[1][:, :, ::8], g["obs_uz_dec8"])
    for name in ("lambda_kernel", "mu_kernel", "density_l_kernel"):
        assert np.array_equal(m.array(name)[::3, ::3], g[name + "_dec"])
    assert np.array_equal(m.array("vs_kernel")[::3, ::3], g["vs_kernel_dec"])
    absmax = [np.abs(k).max() for k in m.get_kernels()]
    assert np.array_equal(np.array(absmax), g["kernel_absmax"])


@pytest.mark.skipif(not ref_available(), reason="oracle/_ref/libpsv_ref.so not built on this machine")
def test_ragged_case_bit_exact_vs_live_reference(tmp_path):
    ini = str(tmp_path / "ragged.ini")
    ragged_config(ini)
    r, o = RefModel(ini), OracleModel(ini)
    vp, vs, rho = random_medium(r.nx, r.nz, seed=7)
    for m in (r, o):
        m.set_parameter_fields(vp, vs, rho)
        for s in range(m.n_shots):
            m.forward_simulate(s, True)
    obs = [a * f for a, f in zip(r.get_synthetic_data(), (1.05, 0.95))]
    for m in (r, o):
        m.set_observed_data(*obs)
        gradient_workflow(m)
    for name in ("taper", "stf", "moment", "rtf_ux", "rtf_uz", "accu_vx", "accu_vz", "accu_txx", "accu_tzz",
                 "accu_txz", "lambda_kernel", "mu_kernel", "density_l_kernel", "vp_kernel", "vs_kernel",
                 "density_v_kernel", "vx", "vz", "txx", "tzz", "txz"):
        assert np.array_equal(r.array(name), o.array(name)), name
    assert r.misfit == o.misfit
    assert np.array_equal(r.get_gradient_vector(), o.get_gradient_vector())
    assert np.array_equal(r.get_model_vector(), o.get_model_vector())
    v = o.get_model_vector() * 1.01
    r.set_model_vector(v)
    o.set_model_vector(v)
    for name in ("vp", "vs", "rho", "lm", "la", "mu", "b_vx"):
        assert np.array_equal(r.array(name), o.array(name)), name
