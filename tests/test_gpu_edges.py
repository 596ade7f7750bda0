"""Edge cases of the path against the CPU oracle, every kernel variant, "exact" arithmetic (bit-identical):
the free window flush with the absorbing frame (no inner boundary), a 1- and 0-point Cerjan frame,
a snapshot at every step, a single receiver, rows wide enough for several CTA tiles of every variant,
and the smallest grid the 4th-order stencil accepts.
"""
import numpy as np
import pytest

from helpers import gpu_array, gradient_workflow, make_config, random_medium
from oracle.oracle import OracleModel

pytestmark = pytest.mark.gpu

ARRAYS = ("rtf_ux", "rtf_uz", "vx", "txz", "lambda_kernel", "mu_kernel", "density_l_kernel", "vp_kernel")

CASES = {
    "window_flush_with_frame": dict(nt=60, nx_inner=40, nz_inner=30, nx_inner_boundary=0, nz_inner_boundary=0, np_boundary=3,
                                    snapshot_interval=5),
    "one_point_frame": dict(nt=60, nx_inner=40, nz_inner=30, nx_inner_boundary=0, nz_inner_boundary=0, np_boundary=1,
                            snapshot_interval=4),
    "no_frame": dict(nt=50, nx_inner=40, nz_inner=30, nx_inner_boundary=3, nz_inner_boundary=3, np_boundary=0,
                     snapshot_interval=5),
    "snapshot_every_step": dict(nt=40, nx_inner=33, nz_inner=27, nx_inner_boundary=2, nz_inner_boundary=2, np_boundary=4,
                                snapshot_interval=1),
    "wide_rows_many_tiles": dict(nt=40, nx_inner=24, nz_inner=1400, nx_inner_boundary=2, nz_inner_boundary=5, np_boundary=6,
                                 snapshot_interval=8, ix_receivers=[20], iz_receivers=[703]),
    "smallest_grid": dict(nt=30, nx_inner=6, nz_inner=6, nx_inner_boundary=0, nz_inner_boundary=0, np_boundary=2,
                          snapshot_interval=3),
}


def config(path, **over):
    nxi, nzi = over["nx_inner"], over["nz_inner"]
    base = dict(n_sources=2, n_shots=2, ix_sources=[nxi // 3, nxi - 2], iz_sources=[2, nzi // 2], moment_angles=[20.0, 250.0],
                which_source_to_fire_in_which_shot=[[0], [1]], nr=1, ix_receivers=[nxi // 2], iz_receivers=[nzi // 3],
                npx=1, npz=1, np_factor=0.05, delay_cycles_per_shot=0.2, dt=0.00025)
    base.update(over)
    return make_config(path, **base)


@pytest.mark.parametrize("name", sorted(CASES))
def test_edge_case_every_variant_bit_identical(tmp_path, name):
    import psvwave_b200
    ini = str(tmp_path / f"{name}.ini")
    config(ini, **CASES[name])
    try:
        o = OracleModel(ini)
    except ValueError:
        with pytest.raises(ValueError):  # what the reference rejects, the library rejects
            psvwave_b200.fdModel(ini)
        return
    med = random_medium(o.nx, o.nz, seed=31)
    o.set_parameter_fields(*med)
    for s in range(o.n_shots):
        o.forward_simulate(s, True)
    o.set_observed_data(*[1.1 * np.array(a) for a in o.get_synthetic_data()])
    gradient_workflow(o)
    assert np.abs(o.get_gradient_vector()).max() > 0
    for variant in range(psvwave_b200.max_variant() + 1):
        m = psvwave_b200.fdModel(ini)
        m._set_option("variant", variant)
        m._set_option("exact", 1)
        m.set_parameter_fields(*med)
        for s in range(m.n_shots):
            m.forward_simulate(s, True)
        m.set_observed_data(*o.get_observed_data())
        gradient_workflow(m)
        for a in ARRAYS:
            assert np.array_equal(gpu_array(m, a), o.array(a)), f"{name}: variant {variant} {a}"
        assert np.array_equal(m.get_gradient_vector(), o.get_gradient_vector()), f"{name}: variant {variant} gradient"
