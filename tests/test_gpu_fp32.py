"""Optional FP32 wavefield mode (north_star: "within a stated tolerance for an optional FP32 mode";
the reference's counterpart is the compile-time `real_simulation` typedef, src/fdModel.h:17-21).
`_set_option("fp32", 1)` keeps the wavefield, the medium the stencil reads and the snapshots in single
precision; seismograms, sensitivity kernels and everything outside the time loop stay double.
STATED TOLERANCE against the FP64 CPU oracle: 1e-3 relative L2 on synthetics, kernels and gradient.
"""
import numpy as np
import pytest

from conftest import SMALL_INI, rel_l2
from helpers import DYNAMIC, gpu_array, gradient_workflow, ragged_config, random_medium
from oracle.oracle import OracleModel

pytestmark = pytest.mark.gpu
TOL32 = 1e-3


def fd(ini, **options):
    import psvwave_b200
    m = psvwave_b200.fdModel(ini)
    for k, v in options.items():
        m._set_option(k, v)
    return m


def test_fp32_small_multishot_against_reference_golden(golden):
    g = golden("small_multishot")
    m = fd(SMALL_INI, fp32=1)
    assert m._get_option("fp32") == 1
    m.set_parameter_fields(g["vp"], g["vs"], g["rho"])
    for s in range(m.n_shots):
        m.forward_simulate(s, True)
    ux, uz = m.get_synthetic_data()
    worst = max(rel_l2(ux, g["ux"]), rel_l2(uz, g["uz"]))
    snaps = m.get_snapshots()
    for f, a in zip(DYNAMIC, snaps):
        assert a.shape == (m.n_shots, m.snapshots, m.nx, m.nz)
        worst = max(worst, rel_l2(a[:, [0, 17, 60]], g["snap_" + f]), rel_l2(gpu_array(m, f), g["fwd_end_" + f]))
    m.set_observed_data(g["obs_ux"], g["obs_uz"])
    gradient_workflow(m)
    assert abs(m.misfit - float(g["misfit"])) <= TOL32 * float(g["misfit"])
    for name in ("lambda_kernel", "mu_kernel", "density_l_kernel", "vp_kernel", "vs_kernel", "density_v_kernel"):
        worst = max(worst, rel_l2(gpu_array(m, name), g[name]))
    worst = max(worst, rel_l2(m.get_gradient_vector(), g["gradient"]))
    print(f"fp32 mode: worst relative L2 against the FP64 reference = {worst:.2e}")
    assert worst <= TOL32
    assert 1e-9 < worst  # it really ran in single precision


def test_fp32_ragged_against_oracle_and_switching_back(tmp_path):
    ini = str(tmp_path / "ragged.ini")
    ragged_config(ini)
    med = random_medium(61, 47, seed=4)
    o = OracleModel(ini)
    o.set_parameter_fields(*med)
    for s in range(o.n_shots):
        o.forward_simulate(s, True)
    o.set_observed_data(*[0.9 * np.array(a) for a in o.get_synthetic_data()])
    gradient_workflow(o)
    m = fd(ini)
    m.set_parameter_fields(*med)
    m._set_option("fp32", 1)  # after the medium was set: the float images follow
    for s in range(m.n_shots):
        m.forward_simulate(s, True)
    m.set_observed_data(*o.get_observed_data())
    gradient_workflow(m)
    assert rel_l2(m.get_synthetic_data()[0], o.get_synthetic_data()[0]) <= TOL32
    assert rel_l2(m.get_gradient_vector(), o.get_gradient_vector()) <= TOL32
    # the copy keeps the mode and the state
    c = m.copy()
    assert c._get_option("fp32") == 1
    assert np.array_equal(gpu_array(c, "vx"), gpu_array(m, "vx"))
    # back to double: bit-identical to the oracle again in exact arithmetic
    m._set_option("fp32", 0)
    m._set_option("exact", 1)
    for s in range(m.n_shots):
        m.forward_simulate(s, True)
    gradient_workflow(m)
    assert np.array_equal(m.get_synthetic_data()[0], o.get_synthetic_data()[0])
    assert np.array_equal(m.get_gradient_vector(), o.get_gradient_vector())


def test_fp32_refuses_what_it_does_not_cover(tmp_path):
    import psvwave_b200
    ini = str(tmp_path / "ragged.ini")
    ragged_config(ini)
    m = fd(ini, fp32=1, variant=1)
    with pytest.raises(ValueError):
        m.forward_simulate(0)
    a, b = fd(ini, fp32=1), fd(ini)
    with pytest.raises(ValueError):
        psvwave_b200.forward_batch([a, b], 0, False)
    b._set_option("fp32", 1)
    psvwave_b200.forward_batch([a, b], 0, True)
    c = fd(ini, fp32=1)
    c.forward_simulate(0, True)
    assert np.array_equal(a.get_synthetic_data()[0][0], c.get_synthetic_data()[0][0])


@pytest.mark.parametrize("n_slabs,native", [(2, True), (3, True), (2, False)])
def test_fp32_on_a_split_grid_equals_the_whole_grid(golden, n_slabs, native):
    """The FP32 mode on x-slabs (halo rows travel as floats): same kernels, same arithmetic on the same inputs, so the
    split run is bit-identical to the whole-grid FP32 run -- and within the stated tolerance of the FP64 reference."""
    import psvwave_b200
    g = golden("small_multishot")
    devices = list(range(min(psvwave_b200.device_count(), n_slabs))) if native else None
    d = psvwave_b200.SlabDomain(SMALL_INI, n_slabs=n_slabs, halo_steps=2, options=dict(fp32=1), native=native,
                                devices=devices)
    w = fd(SMALL_INI, fp32=1)
    for x in (d, w):
        x.set_parameter_fields(g["vp"], g["vs"], g["rho"])
        x.set_observed_data(g["obs_ux"], g["obs_uz"])
        gradient_workflow(x)
    assert np.array_equal(d.get_synthetic_data()[0], w.get_synthetic_data()[0])
    assert np.array_equal(d.get_synthetic_data()[1], w.get_synthetic_data()[1])
    for a, b in zip(d.get_kernels(), w.get_kernels()):
        assert np.array_equal(a, b)
    assert np.array_equal(d.get_gradient_vector(), w.get_gradient_vector())
    assert rel_l2(d.get_gradient_vector(), g["gradient"]) <= TOL32
