"""Out-of-bounds behaviour of every kernel family, checked with canaries (compute-sanitizer is closed on the GPU pool:
profiles/r2b_sanitizer_closed.md).

The device layout surrounds every grid array with guard bands and pads its rows to a multiple of 16 columns
 so that the +-2-row / +-1-pair neighbour loads and the bulk copies of whole row segments need no
bounds tests: what they fetch there only ever feeds frame points, whose results are discarded.  Option "poison_guards"
fills all of that -- guards and pad columns of the state, the medium, the kernels, the slack around snapshots and Lame
kernels -- with NaN canaries.  Then
  * every result must stay BIT-IDENTICAL to the run without canaries (nothing read out of bounds is ever used: one NaN
    reaching a grid point would spread over the whole grid within a few hundred steps), and
  * after the run every guard element must still hold the canary (nothing is ever written out of bounds).
"""
import numpy as np
import pytest

from conftest import SMALL_INI
from helpers import DYNAMIC, gpu_array, gradient_workflow, ragged_config, random_medium

pytestmark = pytest.mark.gpu

KERNELS = ("lambda_kernel", "mu_kernel", "density_l_kernel", "vp_kernel", "vs_kernel", "density_v_kernel")
FAMILIES = [dict(variant=3), dict(variant=3, pipe_warps=8, pipe_stages=3), dict(variant=3, pipe_warps=4, fused_strip=5),
            dict(variant=2), dict(variant=4), dict(variant=1), dict(variant=0), dict(resident=1),
            dict(resident=1, resident_rows=7), dict(variant=3, exact=1), dict(variant=3, fp32=1), dict(variant=2, fp32=1)]


def run(ini, medium, obs, poison, **options):
    import psvwave_b200
    m = psvwave_b200.fdModel(ini)
    for k, v in options.items():
        m._set_option(k, v)
    if poison:
        m._set_option("poison_guards", 1)
    m.set_parameter_fields(*medium)
    m.set_observed_data(*obs)
    gradient_workflow(m)
    out = {"ux": m.get_synthetic_data()[0], "uz": m.get_synthetic_data()[1], "gradient": m.get_gradient_vector(),
           "misfit": np.array([m.misfit])}
    for f in DYNAMIC + KERNELS:
        out[f] = gpu_array(m, f)
    for f, a in zip(DYNAMIC, m.get_snapshots()):
        out["snap_" + f] = a
    return out, (m._get_stat("guard_violations") if poison else 0)


@pytest.mark.parametrize("options", FAMILIES, ids=lambda o: "-".join(f"{k}{v}" for k, v in o.items()))
def test_nothing_out_of_bounds_is_used_or_written(tmp_path, options):
    ini = str(tmp_path / "ragged.ini")
    ragged_config(ini)  # odd sizes: nz = 47 -> 17 pad columns per row; sources / receivers next to the frame
    import psvwave_b200
    probe = psvwave_b200.fdModel(ini)
    medium = random_medium(probe.nx, probe.nz, seed=21)
    rng = np.random.default_rng(2)
    obs = 1e-3 * rng.standard_normal((2, probe.n_shots, probe.nr, probe.nt))
    plain, _ = run(ini, medium, obs, False, **options)
    canaries, violations = run(ini, medium, obs, True, **options)
    for name, a in plain.items():
        assert np.all(np.isfinite(a)), name
        assert np.array_equal(a, canaries[name]), f"{name}: a value read out of bounds reached the grid"
    assert violations == 0, f"{violations} guard elements were overwritten"


def test_canaries_on_a_batch_and_a_split_grid(golden):
    import psvwave_b200
    g = golden("small_multishot")
    medium = (g["vp"], g["vs"], g["rho"])
    # three models per launch (blockIdx.z)
    ms = [psvwave_b200.fdModel(SMALL_INI) for _ in range(3)]
    for k, m in enumerate(ms):
        m._set_option("poison_guards", 1)
        m.set_parameter_fields(medium[0] * (1 + 0.01 * k), medium[1], medium[2])
    for s in range(ms[0].n_shots):
        psvwave_b200.forward_batch(ms, s, True)
    for k, m in enumerate(ms):
        ref = psvwave_b200.fdModel(SMALL_INI)
        ref.set_parameter_fields(medium[0] * (1 + 0.01 * k), medium[1], medium[2])
        for s in range(ref.n_shots):
            ref.forward_simulate(s, True)
        assert np.array_equal(m.get_synthetic_data()[0], ref.get_synthetic_data()[0])
        assert m._get_stat("guard_violations") == 0
    # two slabs, peer-to-peer halo pulls: the halos are filled by the exchange, the guards beyond them never
    d = psvwave_b200.SlabDomain(SMALL_INI, n_slabs=2, halo_steps=1, options=dict(poison_guards=1))
    w = psvwave_b200.fdModel(SMALL_INI)
    for x in (d, w):
        x.set_parameter_fields(*medium)
        x.set_observed_data(g["obs_ux"], g["obs_uz"])
        gradient_workflow(x)
    assert np.array_equal(d.get_synthetic_data()[0], w.get_synthetic_data()[0])
    assert np.array_equal(d.get_gradient_vector(), w.get_gradient_vector())
    assert all(m._get_stat("guard_violations") == 0 for m in d.slabs)
