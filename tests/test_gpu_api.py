"""The reference's own tests re-run against the drop-in (tests/python/test_copy.py,
tests/python/test_descent.py, tests/test_file_constructor.cpp, tests/test_copy_constructor.cpp),
the C ABI called directly through ctypes, text I/O round trips, and size-independent
properties at a grid too large for the CPU oracle."""
import ctypes as C
import os

import numpy as np
import pytest

from conftest import SMALL_INI, rel_l2
from helpers import C1_INI, array_ids, gpu_array, gradient_workflow, make_config, ragged_config, random_medium
from oracle.oracle import OracleModel

pytestmark = pytest.mark.gpu


def fd(ini, **options):
    import psvwave_b200
    m = psvwave_b200.fdModel(ini)
    for k, v in options.items():
        m._set_option(k, v)
    return m


# ---- reference tests, re-expressed ------------------------------------------------------
def test_copy_is_deep(tmp_path):
    """tests/python/test_copy.py:5-24 + tests/test_copy_constructor.cpp:14-69."""
    ini = str(tmp_path / "r.ini")
    ragged_config(ini)
    a = fd(ini)
    a.forward_simulate(0, True)
    b = a.copy()
    va = a.get_model_vector()
    b.set_model_vector(va * 1.1)
    assert np.array_equal(a.get_model_vector(), va)
    assert np.allclose(b.get_model_vector(), va * 1.1, rtol=1e-15)
    # the copy carries seismograms and snapshots
    assert np.array_equal(a.get_synthetic_data()[0], b.get_synthetic_data()[0])
    for x, y in zip(a.get_snapshots(), b.get_snapshots()):
        assert np.array_equal(x, y)
    # same medium -> same seismogram sum (copy constructor test)
    c = a.copy()
    c.forward_simulate(0, False)
    assert c.get_synthetic_data()[0].sum() == a.get_synthetic_data()[0].sum()
    del a
    c.forward_simulate(1, True)  # the copy owns its memory


def test_descent_reduces_misfit(tmp_path):
    """tests/python/test_descent.py:5-137 (smaller grid): three steepest-descent steps."""
    ini = str(tmp_path / "d.ini")
    make_config(ini, nt=600, nx_inner=80, nz_inner=50, nx_inner_boundary=8, nz_inner_boundary=6, np_boundary=12,
                np_factor=0.03, n_sources=2, n_shots=2, ix_sources=[10, 70], iz_sources=[8, 8],
                moment_angles=[90.0, 0.0], which_source_to_fire_in_which_shot=[[0], [1]], nr=8,
                ix_receivers=[5, 15, 25, 35, 45, 55, 65, 75], iz_receivers=[45] * 8, snapshot_interval=5,
                peak_frequency=60.0, delay_cycles_per_shot=0)
    m = fd(ini)
    IX, IZ = m.get_coordinates(True)
    vp, vs, rho = m.get_parameter_fields()
    circle = ((IX - IX.mean()) ** 2 + (IZ - IZ.mean()) ** 2) ** 0.5 < 10
    target = m.copy()
    target.set_parameter_fields(vp * (1 + 0.1 * circle), vs * (1 - 0.1 * circle), rho)
    for s in range(target.n_shots):
        target.forward_simulate(s, False)
    m.set_observed_data(*target.get_synthetic_data())
    x = m.get_model_vector()
    misfits = []
    step = None
    for _ in range(4):
        m.set_model_vector(x)
        gradient_workflow(m)
        misfits.append(m.misfit)
        g = m.get_gradient_vector()
        if step is None:
            step = 0.02 * np.abs(x).max() / np.abs(g).max()
        x = x - step * g
    assert misfits[0] > misfits[-1] > 0, misfits


def test_gradient_matches_finite_difference(tmp_path):
    """Directional derivative of the misfit along a random block direction."""
    ini = str(tmp_path / "g.ini")
    ragged_config(ini, nt=300)
    m = fd(ini)
    o = OracleModel(ini)
    for s in range(o.n_shots):
        o.forward_simulate(s, False)
    obs = [a * 1.2 for a in o.get_synthetic_data()]
    m.set_observed_data(*obs)
    x0 = m.get_model_vector()
    gradient_workflow(m)
    g = m.get_gradient_vector()
    rng = np.random.default_rng(3)
    d = rng.standard_normal(x0.size) * x0 * 1e-2

    def misfit(x):
        m.set_model_vector(x)
        for s in range(m.n_shots):
            m.forward_simulate(s, False)
        m.calculate_l2_misfit()
        return m.misfit

    eps = 1e-3
    fdiff = (misfit(x0 + eps * d) - misfit(x0 - eps * d)) / (2 * eps)
    # discretise-then-optimise mismatch of the scheme is a few percent; sign and size must agree
    assert fdiff * g.dot(d) > 0
    assert abs(fdiff - g.dot(d)) <= 0.25 * abs(fdiff)


# ---- the C ABI, straight through ctypes -------------------------------------------------
def test_c_abi_direct(golden):
    import psvwave_b200
    g = golden("small_multishot")
    lib = C.CDLL(psvwave_b200.LIBRARY_PATH)
    lib.psv_last_error.restype = C.c_char_p
    h = C.c_void_p()
    assert lib.psv_create_from_ini(SMALL_INI.encode(), C.byref(h)) == 0, lib.psv_last_error()
    ids = array_ids()
    dp = lambda a: a.ctypes.data_as(C.POINTER(C.c_double))  # noqa: E731
    n = C.c_size_t()
    assert lib.psv_array_size(h, ids["vp"], C.byref(n)) == 0 and n.value == g["vp"].size
    vp, vs, rho = (np.ascontiguousarray(g[k]) for k in ("vp", "vs", "rho"))
    assert lib.psv_set_parameter_fields(h, dp(vp), dp(vs), dp(rho), C.c_size_t(vp.size)) == 0
    assert lib.psv_set_option(h, b"exact", C.c_longlong(1)) == 0
    for s in range(2):
        assert lib.psv_forward_simulate(h, s, 1, 0, 0) == 0, lib.psv_last_error()
    ux = np.empty_like(g["ux"])
    assert lib.psv_get_array(h, ids["rtf_ux"], dp(ux), C.c_size_t(ux.size)) == 0
    assert np.array_equal(ux, g["ux"])
    # errors: bad shot -> PSV_ERR_INVALID_ARGUMENT, wrong size -> PSV_ERR_INVALID_ARGUMENT
    assert lib.psv_forward_simulate(h, 9, 1, 0, 0) == 1 and b"Shot" in lib.psv_last_error()
    assert lib.psv_get_array(h, ids["rtf_ux"], dp(ux), C.c_size_t(3)) == 1
    obs_x, obs_z = np.ascontiguousarray(g["obs_ux"]), np.ascontiguousarray(g["obs_uz"])
    assert lib.psv_set_array(h, ids["rtf_ux_true"], dp(obs_x), C.c_size_t(obs_x.size)) == 0
    assert lib.psv_set_array(h, ids["rtf_uz_true"], dp(obs_z), C.c_size_t(obs_z.size)) == 0
    misfit = C.c_double()
    assert lib.psv_run_model(h, 0, 1) == 0, lib.psv_last_error()
    assert lib.psv_get_real(h, 11, C.byref(misfit)) == 0  # PSV_REAL_MISFIT
    assert abs(misfit.value - float(g["misfit"])) <= 1e-14 * float(g["misfit"])
    grad = np.empty_like(g["gradient"])
    assert lib.psv_get_gradient_vector(h, dp(grad), C.c_size_t(grad.size)) == 0
    assert np.array_equal(grad, g["gradient"])
    launches = C.c_double()
    assert lib.psv_get_stat(h, b"launches", C.byref(launches)) == 0 and launches.value > 0
    assert lib.psv_destroy(h) == 0


# ---- text I/O (src/fdModel.cpp:971-1126) ---------------------------------------------------
def test_receiver_files_round_trip(tmp_path, in_tmp):
    ini = str(tmp_path / "r.ini")
    ragged_config(ini, nt=60, snapshot_interval=6, npx=1, npz=1)
    m = fd(ini)
    for s in range(m.n_shots):
        m.forward_simulate(s, False)
    m.write_receivers()
    m.write_sources()
    assert os.path.exists("rtf_ux0.txt") and os.path.exists("sources_shot_2.txt")
    m.load_receivers()
    for a, b in zip(m.get_synthetic_data(), m.get_observed_data()):
        assert np.array_equal(a, b)  # digits10+10 significant digits round-trip exactly
    m.calculate_l2_misfit()
    assert m.misfit == 0.0
    # file layout: each receiver = newline + nt values separated by blanks
    rows = open("rtf_ux1.txt").read().split("\n")
    assert rows[0] == "" and len(rows) == m.nr + 1 and len(rows[1].split()) == m.nt
    v = m.get_model_vector()
    np.savetxt("vec.txt", v)
    assert np.array_equal(m.load_vector("vec.txt", False), v)


# ---- properties at a size the CPU oracle is not run at ---------------------------------------
@pytest.fixture(scope="module")
def big(tmp_path_factory):
    d = tmp_path_factory.mktemp("big")
    ini = str(d / "big.ini")
    nxi, nzi = 1200, 700
    make_config(ini, nt=300, nx_inner=nxi, nz_inner=nzi, nx_inner_boundary=20, nz_inner_boundary=20,
                np_boundary=40, np_factor=0.0075, dx=2.5, dz=2.5, dt=2.5e-4, peak_frequency=12.0,
                n_sources=4, n_shots=3, ix_sources=[300, 600, 300, 600], iz_sources=[30, 350, 30, 350],
                moment_angles=[90.0, 45.0, 90.0, 45.0], which_source_to_fire_in_which_shot=[[0], [1], [2, 3]],
                nr=40, ix_receivers=list(range(100, 1100, 25)), iz_receivers=[20] * 40,
                snapshot_interval=20, delay_cycles_per_shot=0.0, npx=1, npz=1)
    return ini


def test_big_variants_agree_and_superpose(big):
    import psvwave_b200
    ms = []
    for variant in range(psvwave_b200.max_variant() + 1):
        m = fd(big, variant=variant, exact=1)
        rng = np.random.default_rng(11)
        smooth = 1 + 0.1 * np.sin(np.arange(m.nx)[:, None] / 60.0) * np.cos(np.arange(m.nz)[None, :] / 45.0)
        m.set_parameter_fields(2500.0 * smooth, 1300.0 * smooth, 2000.0 + 0 * smooth + rng.random((m.nx, m.nz)))
        for s in range(3):
            m.forward_simulate(s, s == 2)
        ms.append(m)
    ux0, uz0 = ms[0].get_synthetic_data()
    assert np.linalg.norm(ux0) > 0
    for m in ms[1:]:  # all kernel variants are the same arithmetic -> bit-identical
        ux, uz = m.get_synthetic_data()
        assert np.array_equal(ux, ux0) and np.array_equal(uz, uz0)
        for f in ("vx", "txz"):
            assert np.array_equal(gpu_array(m, f), gpu_array(ms[0], f))
    # linearity: shot 2 fires copies of sources 0 and 1 together (no delay) = shot 0 + shot 1
    # (the reference requires every source to be listed exactly once, src/fdModel.cpp:433-441)
    assert rel_l2(ux0[0] + ux0[1], ux0[2]) < 1e-12
    assert rel_l2(uz0[0] + uz0[1], uz0[2]) < 1e-12
    # default arithmetic vs IEEE-order arithmetic at scale
    f = fd(big)
    vp, vs, rho = ms[0].get_parameter_fields()
    f.set_parameter_fields(vp, vs, rho)
    f.forward_simulate(2, True)
    assert rel_l2(f.get_synthetic_data()[0][2], ux0[2]) < 1e-10
    # adjoint at scale: variants agree on the kernels
    ks = []
    for m in ms + [f]:
        m.set_observed_data(ux0 * 0.9, uz0 * 1.1)
        m.calculate_l2_adjoint_sources()
        m.reset_kernels()
        m.adjoint_simulate(2)
        m.map_kernels_to_velocity()
        ks.append(m.get_kernels())
    for k in ks[1:-1]:
        for a, b in zip(k, ks[0]):
            assert np.array_equal(a, b)
    for a, b in zip(ks[-1], ks[0]):
        assert rel_l2(a, b) < 1e-10
    # kernels vanish outside the free-parameter window (src/fdModel.cpp:821-825)
    lam = gpu_array(ms[0], "lambda_kernel")
    w = np.zeros_like(lam, dtype=bool)
    w[60:-60, 60:-60] = True
    assert np.all(lam[~w] == 0) and np.abs(lam[w]).max() > 0
