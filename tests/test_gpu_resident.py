"""The resident whole-loop kernel (psvwave_b200/csrc/resident.cuh): ONE launch per forward_simulate / adjoint_simulate
call for grids that fit on chip -- the reference's default configuration (BASELINE.json configs[0], loops
src/fdModel.cpp:552-751 and :814-958).  Same bar as every other kernel: "exact" arithmetic is BIT-IDENTICAL to the
reference's golden vectors / the CPU oracle on every array of the path, default arithmetic <= 1e-10 relative L2 and
bit-identical to the per-step kernels (same point-wise code).  Geometry is swept: rows per CTA from 2 (halos reach two
neighbours) to one CTA for the whole grid (61 rows of the ragged case).
"""
import numpy as np
import pytest

from conftest import SMALL_INI, rel_l2
from helpers import C1_INI, DYNAMIC, gpu_array, gradient_workflow, notebook2_models, ragged_config, random_medium
from oracle.oracle import OracleModel

pytestmark = pytest.mark.gpu

TOL = 1e-10
KERNELS = ("lambda_kernel", "mu_kernel", "density_l_kernel", "vp_kernel", "vs_kernel", "density_v_kernel")


def fd(ini, **options):
    import psvwave_b200
    m = psvwave_b200.fdModel(ini)
    m._set_option("resident", 1)
    for k, v in options.items():
        m._set_option(k, v)
    return m


def close(a, b, what, exact):
    if exact:
        assert np.array_equal(a, b), f"{what}: not bit-identical (rel L2 {rel_l2(a, b):.3e})"
    else:
        e = rel_l2(a, b)
        assert e <= TOL, f"{what}: rel L2 {e:.3e} > {TOL}"


@pytest.mark.parametrize("opts", [dict(exact=1), dict(exact=0), dict(exact=1, resident_rows=3), dict(exact=1, resident_rows=7),
                                  dict(exact=1, resident_rows=40)],
                         ids=lambda o: "-".join(f"{k}{v}" for k, v in o.items()))
def test_small_multishot_vs_reference_golden(golden, opts):
    g = golden("small_multishot")
    exact = bool(opts["exact"])
    m = fd(SMALL_INI, **opts)
    m.set_parameter_fields(g["vp"], g["vs"], g["rho"])
    for s in range(m.n_shots):
        m.forward_simulate(s, True)
        assert m._get_stat("resident_ctas") >= 1 and m._get_stat("launches") <= 2
    ux, uz = m.get_synthetic_data()
    close(ux, g["ux"], "ux", exact)
    close(uz, g["uz"], "uz", exact)
    for f, a in zip(DYNAMIC, m.get_snapshots()):
        close(a[:, [0, 17, 60]], g["snap_" + f], "snapshot " + f, exact)
        close(gpu_array(m, f), g["fwd_end_" + f], "final " + f, exact)
    m.set_observed_data(g["obs_ux"], g["obs_uz"])
    gradient_workflow(m)
    assert m._get_stat("resident_ctas") >= 1 and m._get_stat("launches") == 1
    assert abs(m.misfit - float(g["misfit"])) <= (1e-14 if exact else TOL) * float(g["misfit"])
    for name in ("a_stf_ux", "a_stf_uz") + KERNELS:
        close(gpu_array(m, name), g[name], name, exact)
    close(m.get_gradient_vector(), g["gradient"], "gradient", exact)
    close(gpu_array(m, "vx"), g["adj_end_vx"], "adjoint final vx", exact)
    close(gpu_array(m, "txx"), g["adj_end_txx"], "adjoint final txx", exact)


@pytest.mark.parametrize("rows", [0, 2, 5, 11, 61])
def test_ragged_case_vs_oracle_and_per_step_kernels(tmp_path, rows):
    """Odd grid sizes, 3 shots, coinciding receivers (adjoint injection in two rounds), a receiver next to a source,
    stacked sources: every array of the path, against the oracle (exact) and the per-step default kernel (default
    arithmetic; FMA contraction may differ between the two kernels, so the tolerance applies)."""
    import psvwave_b200
    ini = str(tmp_path / "ragged.ini")
    ragged_config(ini)
    o = OracleModel(ini)
    models = [o, fd(ini, exact=1, resident_rows=rows), fd(ini, exact=0, resident_rows=rows), psvwave_b200.fdModel(ini)]
    assert models[3]._get_option("resident") == 0  # the suite's default: per-step launches
    vp, vs, rho = random_medium(o.nx, o.nz, seed=7)
    for x in models:
        x.set_parameter_fields(vp, vs, rho)
        for s in range(x.n_shots):
            x.forward_simulate(s, True)
    obs = [a * f for a, f in zip(o.get_synthetic_data(), (1.05, 0.95))]
    for x in models:
        x.set_observed_data(*obs)
        gradient_workflow(x)
    _, me, mf, ms = models
    assert me._get_stat("resident_ctas") >= 1 and ms._get_stat("resident_ctas") == 0
    for name in ("rtf_ux", "rtf_uz", "a_stf_ux", "a_stf_uz", "accu_vx", "accu_vz", "accu_txx", "accu_tzz", "accu_txz",
                 "vx", "vz", "txx", "tzz", "txz") + KERNELS:
        close(gpu_array(me, name), o.array(name), name, True)
        close(gpu_array(mf, name), gpu_array(ms, name), name + " (default arithmetic vs per-step kernel)", False)
        close(gpu_array(mf, name), o.array(name), name, False)
    close(me.get_gradient_vector(), o.get_gradient_vector(), "gradient", True)
    assert me.misfit == pytest.approx(o.misfit, rel=1e-14)


def test_default_configuration_auto_mode_and_notebook2_answers(golden):
    """configs[0]: "resident" = -1 (the library default) picks the loop kernel on the 250 x 150 grid by itself; the
    homogeneous forward run and the notebook-2 gradient reproduce the reference's golden vectors and printed answers."""
    import psvwave_b200
    m = psvwave_b200.fdModel(C1_INI)
    m._set_option("resident", -1)
    m.forward_simulate(0, True)
    assert m._get_stat("resident_ctas") >= 60 and m._get_stat("launches") == 2
    g = golden("c1_homogeneous")
    ux, uz = m.get_synthetic_data()
    close(ux[:, :, ::8], g["ux_dec8"], "ux", False)
    close(uz[:, :, ::8], g["uz_dec8"], "uz", False)
    assert np.linalg.norm(ux) == pytest.approx(150626.52785852677, rel=TOL)
    snaps = m.get_snapshots()
    close(snaps[0][0, 400][::3, ::3], g["snap_vx_400_dec"], "snapshot vx[400]", False)
    close(snaps[4][0, 400][::3, ::3], g["snap_txz_400_dec"], "snapshot txz[400]", False)
    del snaps
    close(gpu_array(m, "tzz")[::3, ::3], g["final_tzz_dec"], "final tzz", False)
    start, target = notebook2_models(m)
    m.set_parameter_fields(*target)
    m.forward_simulate(0, True)
    m.set_observed_data(*m.get_synthetic_data())
    m.set_parameter_fields(*start)
    gradient_workflow(m)
    assert m._get_stat("resident_ctas") >= 60
    assert f"{m.misfit:.2f}" == "1920401.82"  # notebooks/Notebook 2 ...ipynb:240
    grad = m.get_gradient_vector()
    assert f"{np.abs(grad).max():.5E}" == "7.71590E+01"  # ...ipynb:443
    g2 = golden("c1_notebook2")
    close(grad[::7], g2["g_dec7"], "gradient", False)
    for name in ("lambda_kernel", "mu_kernel", "density_l_kernel", "vs_kernel"):
        close(gpu_array(m, name)[::3, ::3], g2[name + "_dec"], name, False)


def test_what_the_loop_kernel_refuses(tmp_path):
    import psvwave_b200
    ini = str(tmp_path / "r.ini")
    ragged_config(ini)
    m = fd(ini)
    m._set_option("profile", 1)
    with pytest.raises(ValueError):
        m.forward_simulate(0, True)  # resident = 1 turns "does not apply" into an error
    m._set_option("profile", 0)
    m._set_option("fp32", 1)
    with pytest.raises(ValueError):
        m.forward_simulate(0, True)
    m._set_option("fp32", 0)
    with pytest.raises(ValueError):
        m._set_option("resident_rows", 1)
    m.forward_simulate(0, True)
    assert m._get_stat("resident_ctas") >= 1
    # wavefield dumps need the host between steps: the per-step path runs, silently, as before
    (tmp_path / "snapshots").mkdir()
    import os
    cwd = os.getcwd()
    os.chdir(tmp_path)
    try:
        m.forward_simulate(0, False, False, True)
    finally:
        os.chdir(cwd)
    assert m._get_stat("resident_ctas") == 0


def test_narrow_grid_with_more_receivers_than_threads_in_a_strip(tmp_path):
    """A narrow grid launches fewer threads per CTA than the compiled maximum (nz = 24 -> pitch 32 -> 96 threads for
    2-row strips) and the loop kernel gives each receiver / source of a strip to one thread: 110 receivers on one grid
    row do not fit.  The plan has to see that -- automatic mode falls back to the per-step kernels (same seismograms
    and kernels as resident = 0, bit for bit), resident = 1 says why -- instead of dropping receivers 96..109.  90
    receivers on that row do fit and run in the loop kernel, identical to the per-step result."""
    import psvwave_b200

    def run(nr, resident):
        ini = str(tmp_path / f"narrow{nr}.ini")
        ragged_config(ini, nt=120, nx_inner=40, nz_inner=10, nx_inner_boundary=3, nz_inner_boundary=2, n_sources=2,
                      ix_sources=[12, 30], iz_sources=[4, 6], moment_angles=[20.0, 70.0], n_shots=1,
                      which_source_to_fire_in_which_shot=[[0, 1]], nr=nr, ix_receivers=[9] * nr,
                      iz_receivers=[k % 10 for k in range(nr)], npx=1, npz=1)
        m = psvwave_b200.fdModel(ini)
        m._set_option("exact", 1)
        m._set_option("resident", resident)
        vp, vs, rho = random_medium(m.nx, m.nz, seed=3)
        m.set_parameter_fields(vp, vs, rho)
        m.forward_simulate(0, True)
        fwd_ctas = m._get_stat("resident_ctas")
        ux, uz = m.get_synthetic_data()
        m.set_observed_data(ux * 1.1, uz * 0.9)
        gradient_workflow(m)
        return m, fwd_ctas, m._get_stat("resident_ctas")

    ref, c0, c1 = run(110, 0)
    assert c0 == 0 and c1 == 0
    auto, c0, c1 = run(110, -1)
    assert c0 == 0 and c1 == 0  # too many point operations for 96 threads: per-step launches
    few_ref, _, _ = run(90, 0)
    few, c0, c1 = run(90, -1)
    assert c0 >= 1 and c1 >= 1  # fits: the loop kernel runs
    for a, b in ((auto, ref), (few, few_ref)):
        ux, uz = a.get_synthetic_data()
        rx, rz = b.get_synthetic_data()
        assert np.all(np.abs(rx).max(axis=2) > 0)  # every trace recorded
        assert np.array_equal(ux, rx) and np.array_equal(uz, rz)
        for name in KERNELS:
            close(gpu_array(a, name), gpu_array(b, name), name, True)
    with pytest.raises(ValueError, match="too many sources / receivers"):
        run(110, 1)
