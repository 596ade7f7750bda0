"""Monte-Carlo ensembles (BASELINE.json configs[3], SURVEY 8e "replicas only"): n models of one
experiment advanced by ONE launch per time step (psv_forward_batch / psv_adjoint_batch).
The reference's route is n independent copies of the object (src/fdModel.cpp:64-97), so the bar
is: every model of the batch ends bit-identical to the same model simulated on its own by the
same kernels, and -- in "exact" arithmetic -- bit-identical to the CPU oracle.
"""
import numpy as np
import pytest

from helpers import DYNAMIC, gpu_array, ragged_config, random_medium
from oracle.oracle import OracleModel

pytestmark = pytest.mark.gpu

KERNELS = ("lambda_kernel", "mu_kernel", "density_l_kernel")
BATCH_VARIANTS = [dict(variant=2), dict(variant=3), dict(variant=3, pipe_warps=4), dict(variant=3, pipe_warps=8), dict(variant=4),
                  dict(variant=3, tune=1, graph=0)]  # launch geometry measured on the ensemble's own state arrays first


def fd(ini, **options):
    import psvwave_b200
    m = psvwave_b200.fdModel(ini)
    for k, v in options.items():
        m._set_option(k, v)
    return m


def gradient_of_shots(models, batched):
    import psvwave_b200
    n_shots = models[0].n_shots
    for s in range(n_shots):
        if batched:
            psvwave_b200.forward_batch(models, s, True)
        else:
            for m in models:
                m.forward_simulate(s, True)
    for m in models:
        m.calculate_l2_misfit()
        m.calculate_l2_adjoint_sources()
        m.reset_kernels()
    for s in range(n_shots):
        if batched:
            psvwave_b200.adjoint_batch(models, s)
        else:
            for m in models:
                m.adjoint_simulate(s)
    for m in models:
        m.map_kernels_to_velocity()


@pytest.mark.parametrize("opts", BATCH_VARIANTS, ids=lambda o: "-".join(f"{k}{v}" for k, v in o.items()))
def test_batch_equals_separate_runs_and_oracle(tmp_path, opts):
    ini = str(tmp_path / "ragged.ini")
    ragged_config(ini)
    n = 5
    media = [random_medium(61, 47, seed=100 + i) for i in range(n)]
    obs_model = OracleModel(ini)
    obs_model.set_parameter_fields(*random_medium(61, 47, seed=7))
    for s in range(obs_model.n_shots):
        obs_model.forward_simulate(s, False)
    obs = obs_model.get_synthetic_data()

    def make(exact):
        ms = [fd(ini, exact=exact, **opts) for _ in range(n)]
        for m, med in zip(ms, media):
            m.set_parameter_fields(*med)
            m.set_observed_data(*obs)
        return ms

    batch, single = make(1), make(1)
    gradient_of_shots(batch, True)
    gradient_of_shots(single, False)
    assert batch[0]._get_stat("launches") == single[0]._get_stat("launches")  # one launch per step for all n
    for i, (b, s) in enumerate(zip(batch, single)):
        for name in ("rtf_ux", "rtf_uz") + KERNELS + DYNAMIC:
            assert np.array_equal(gpu_array(b, name), gpu_array(s, name)), f"model {i} {name}"
        assert np.array_equal(b.get_gradient_vector(), s.get_gradient_vector())
        assert b.misfit == s.misfit
        for fb, fs in zip(b.get_snapshots(), s.get_snapshots()):
            assert np.array_equal(fb, fs)
    # against the oracle (first and last model of the ensemble)
    for i in (0, n - 1):
        o = OracleModel(ini)
        o.set_parameter_fields(*media[i])
        o.set_observed_data(*obs)
        gradient_of_shots([o], False)
        assert np.array_equal(batch[i].get_synthetic_data()[0], o.get_synthetic_data()[0])
        assert np.array_equal(batch[i].get_gradient_vector(), o.get_gradient_vector())
    # default arithmetic: the batch is still bit-identical to separate runs
    fb, fs = make(0), make(0)
    gradient_of_shots(fb, True)
    gradient_of_shots(fs, False)
    for b, s in zip(fb, fs):
        assert np.array_equal(b.get_gradient_vector(), s.get_gradient_vector())


def test_batch_rejects_mismatched_models(tmp_path):
    import psvwave_b200
    a_ini, b_ini = str(tmp_path / "a.ini"), str(tmp_path / "b.ini")
    ragged_config(a_ini)
    ragged_config(b_ini, nt=120)
    a, b = fd(a_ini), fd(b_ini)
    with pytest.raises(ValueError):
        psvwave_b200.forward_batch([a, b], 0, False)
    with pytest.raises(ValueError):
        psvwave_b200.forward_batch([a, a], 0, False)
    with pytest.raises(ValueError):
        psvwave_b200.forward_batch([], 0, False)
    c = fd(a_ini, variant=1)
    with pytest.raises(ValueError):
        psvwave_b200.forward_batch([c, a], 0, False)
    with pytest.raises((ValueError, RuntimeError)):
        psvwave_b200.adjoint_batch([a, fd(a_ini)], 0)  # no stored forward wavefield


def test_batch_of_one_is_forward_simulate(tmp_path):
    import psvwave_b200
    ini = str(tmp_path / "ragged.ini")
    ragged_config(ini)
    a, b = fd(ini, exact=1), fd(ini, exact=1)
    med = random_medium(61, 47, seed=3)
    a.set_parameter_fields(*med)
    b.set_parameter_fields(*med)
    psvwave_b200.forward_batch([a], 1, True)
    b.forward_simulate(1, True)
    assert np.array_equal(a.get_synthetic_data()[0], b.get_synthetic_data()[0])
