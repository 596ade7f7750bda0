"""Domain split (BASELINE.json configs[4], SURVEY 8e "large single grids"): the grid cut into
x-slabs with halo exchange must reproduce the whole-grid run.  Here the slabs live in one process:
"native" = the whole march inside the library (psv_march_slabs: peer-to-peer pulls between the
slabs' state arrays, CUDA events between their streams), on one GPU or one slab per GPU when the
box has several; "staged" = pack kernel -> device copy -> unpack kernel, the buffers and kernels
the NCCL transport of the one-process-per-GPU layout uses.  The bar is the strictest one: with
"exact" arithmetic every array is BIT-IDENTICAL to the whole-grid CUDA run and to the CPU oracle;
default arithmetic <= 1e-10.
"""
import numpy as np
import pytest

from conftest import rel_l2
from helpers import DYNAMIC, array_ids, gpu_array, make_config, random_medium
from oracle.oracle import OracleModel

pytestmark = pytest.mark.gpu

TOL = 1e-10


def slab_config(path, **extra):
    """Long in x (so that 2-4 slabs each own >= 16 rows), odd nz, sources/receivers near the cuts."""
    base = dict(nt=150, nx_inner=130, nz_inner=33, nx_inner_boundary=4, nz_inner_boundary=3,
                np_boundary=7, np_factor=0.04, n_sources=4, n_shots=2,
                ix_sources=[8, 64, 65, 100], iz_sources=[6, 7, 20, 5],
                moment_angles=[0.0, 45.0, 90.0, 135.0],
                which_source_to_fire_in_which_shot=[[1, 0], [2, 3]],
                nr=8, ix_receivers=[3, 63, 64, 65, 66, 29, 30, 120], iz_receivers=[3, 7, 7, 30, 16, 9, 9, 11],
                snapshot_interval=6, npx=2, npz=9, delay_cycles_per_shot=2, dt=0.00025)
    base.update(extra)
    return make_config(path, **base)


def workflow(m, medium, obs):
    m.set_parameter_fields(*medium)
    for s in range(m.n_shots):
        m.forward_simulate(s, True)
    m.set_observed_data(*obs)
    m.calculate_l2_misfit()
    m.calculate_l2_adjoint_sources()
    m.reset_kernels()
    for s in range(m.n_shots):
        m.adjoint_simulate(s)
    m.map_kernels_to_velocity()


def whole(ini, **opts):
    import psvwave_b200
    m = psvwave_b200.fdModel(ini)
    for k, v in opts.items():
        m._set_option(k, v)
    return m


@pytest.fixture(scope="module")
def case(tmp_path_factory):
    ini = str(tmp_path_factory.mktemp("slab") / "slab.ini")
    slab_config(ini)
    nx, nz = 144, 47
    medium = random_medium(nx, nz, seed=11)
    o = OracleModel(ini)
    o.set_parameter_fields(*random_medium(nx, nz, seed=12))
    for s in range(o.n_shots):
        o.forward_simulate(s, False)
    obs = tuple(np.array(a) for a in o.get_synthetic_data())
    ref = OracleModel(ini)
    workflow(ref, medium, obs)
    return dict(ini=ini, medium=medium, obs=obs, oracle=ref, shape=(nx, nz))


def gpus():
    import psvwave_b200
    return psvwave_b200.device_count()


@pytest.mark.parametrize("transport", ["native", "staged", "native-multi-gpu"])
@pytest.mark.parametrize("n_slabs,halo_steps,variant", [(2, 1, 3), (2, 4, 3), (3, 2, 3), (4, 4, 2), (3, 1, 2), (1, 4, 3)])
def test_split_grid_is_bit_identical_to_whole_grid_and_oracle(case, n_slabs, halo_steps, variant, transport):
    import psvwave_b200
    opts = dict(variant=variant, exact=1)
    devices = None
    if transport == "native-multi-gpu":
        if gpus() < 2:
            pytest.skip("needs 2 CUDA devices")
        devices = list(range(min(gpus(), n_slabs)))  # more slabs than GPUs: dealt round-robin
    d = psvwave_b200.SlabDomain(case["ini"], n_slabs=n_slabs, halo_steps=halo_steps, options=opts,
                                native=transport != "staged", devices=devices)
    if devices and n_slabs > 1:
        assert [int(m.device) for m in d.slabs] == [devices[k % len(devices)] for k in range(n_slabs)]
    assert (d.nx, d.nz) == case["shape"]
    owned = [d.owned_rows(k)[0] for k in range(n_slabs)]
    assert owned[0][0] == 0 and owned[-1][1] == d.nx and all(a[1] == b[0] for a, b in zip(owned, owned[1:]))
    workflow(d, case["medium"], case["obs"])
    w = whole(case["ini"], **opts)
    workflow(w, case["medium"], case["obs"])
    o = case["oracle"]
    ux, uz = d.get_synthetic_data()
    for a, b, c in zip((ux, uz), w.get_synthetic_data(), o.get_synthetic_data()):
        assert np.array_equal(a, b) and np.array_equal(a, c)
    assert d.calculate_l2_misfit() == w.misfit
    for a, b, c in zip(d.get_kernels(), w.get_kernels(), o.get_kernels()):
        assert np.array_equal(a, b) and np.array_equal(a, c)
    assert np.array_equal(d.get_gradient_vector(), w.get_gradient_vector())
    assert np.array_equal(d.get_gradient_vector(), o.get_gradient_vector())
    # final adjoint state on the rows each slab owns
    ids = array_ids()
    for f in DYNAMIC:
        got = d.gather_grid(lambda m, f=f: m._get_array(ids[f]))
        assert np.array_equal(got, gpu_array(w, f)), f
    if n_slabs > 1:
        assert d.exchanges == (d.nt - 1) // halo_steps


def test_split_grid_default_arithmetic_within_tolerance(case):
    import psvwave_b200
    d = psvwave_b200.SlabDomain(case["ini"], n_slabs=3, halo_steps=4)
    workflow(d, case["medium"], case["obs"])
    o = case["oracle"]
    assert rel_l2(d.get_synthetic_data()[0], o.get_synthetic_data()[0]) <= TOL
    assert rel_l2(d.get_gradient_vector(), o.get_gradient_vector()) <= TOL
    for a, c in zip(d.get_kernels(), o.get_kernels()):
        assert rel_l2(a, c) <= TOL


def test_slab_handles_refuse_what_needs_the_whole_grid(case):
    import psvwave_b200
    with pytest.raises(ValueError):
        psvwave_b200.fdModel._slab(case["ini"], 0, 2, 6)  # halo not a multiple of 4
    with pytest.raises(ValueError):
        psvwave_b200.fdModel._slab(case["ini"], 0, 12, 16)  # slabs narrower than the halo
    with pytest.raises(ValueError):
        psvwave_b200.fdModel._slab(case["ini"], 2, 2, 16)
    s = psvwave_b200.fdModel._slab(case["ini"], 1, 2, 8)
    info = s._slab_info()
    assert info["is_slab"] and info["nx_global"] == 144 and info["x_offset"] == 72 - 8 and info["halo_hi"] == 0
    assert s.nx == info["nx_local"] == 72 + 8
    with pytest.raises(RuntimeError):
        s.forward_simulate(0)
    with pytest.raises(RuntimeError):
        s.get_gradient_vector()
    s._set_option("variant", 1)
    with pytest.raises(ValueError):
        s._march_begin(0, False, False)


def test_march_slabs_checks_what_it_is_given(case):
    """psv_march_slabs wants ALL slabs of one grid, in slab order, with one halo width and one arithmetic mode."""
    import psvwave_b200
    mk = lambda r, n, h: psvwave_b200.fdModel._slab(case["ini"], r, n, h)  # noqa: E731
    a, b, c = mk(0, 3, 8), mk(1, 3, 8), mk(2, 3, 8)
    with pytest.raises(ValueError):
        psvwave_b200.march_slabs([b, a, c], 0, False, False)  # wrong order
    with pytest.raises(ValueError):
        psvwave_b200.march_slabs([a, b], 0, False, False)  # one slab missing
    with pytest.raises(ValueError):
        psvwave_b200.march_slabs([a, mk(1, 3, 16), c], 0, False, False)  # another halo width
    with pytest.raises(RuntimeError):
        psvwave_b200.march_slabs([psvwave_b200.fdModel(case["ini"])], 0, False, False)  # not a slab
    b._set_option("fp32", 1)
    with pytest.raises(RuntimeError):
        psvwave_b200.march_slabs([a, b, c], 0, False, False)  # the slabs must agree on the FP32 mode
    b._set_option("fp32", 0)
    psvwave_b200.march_slabs([a, b, c], 0, False, False)  # and no failed attempt leaves a march open
    with pytest.raises(RuntimeError):
        psvwave_b200.march_slabs([a, b, c], 0, True, False)  # adjoint without a stored forward march
    psvwave_b200.march_slabs([a, b, c], 0, False, True)
    psvwave_b200.march_slabs([a, b, c], 0, True, False)
