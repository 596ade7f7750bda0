"""Option "graph": the time loop of one (direction, shot, store) captured into a CUDA graph on first use and
replayed afterwards (the default for small grids, whose step kernel is shorter than the gap between two stream
launches; every other small-grid GPU test therefore also runs through it).
Capture and replay must give the arrays of the plain stream launches bit for bit, "exact" arithmetic the
CPU oracle's, and a replay must pick up new media, observed data and options.
"""
import numpy as np
import pytest

from helpers import gpu_array, gradient_workflow, make_config, random_medium
from oracle.oracle import OracleModel

pytestmark = pytest.mark.gpu

ARRAYS = ("rtf_ux", "rtf_uz", "vx", "vz", "txx", "tzz", "txz", "lambda_kernel", "mu_kernel", "density_l_kernel", "vp_kernel")


def config(path):
    return make_config(path, nt=84, nx_inner=40, nz_inner=70, nx_inner_boundary=3, nz_inner_boundary=2, np_boundary=5,
                       np_factor=0.05, snapshot_interval=7, n_sources=3, n_shots=2, ix_sources=[5, 20, 33], iz_sources=[4, 30, 9],
                       moment_angles=[10.0, 100.0, 215.0], which_source_to_fire_in_which_shot=[[0, 2], [1]], nr=3,
                       ix_receivers=[8, 21, 36], iz_receivers=[6, 6, 50], npx=2, npz=2, delay_cycles_per_shot=1, dt=0.00025)


def run(m, med, obs):
    m.set_parameter_fields(*med)
    m.set_observed_data(*obs)
    gradient_workflow_store(m)


def gradient_workflow_store(m):
    for s in range(m.n_shots):
        m.forward_simulate(s, True)
    m.calculate_l2_misfit()
    m.calculate_l2_adjoint_sources()
    m.reset_kernels()
    for s in range(m.n_shots):
        m.adjoint_simulate(s)
    m.map_kernels_to_velocity()


@pytest.mark.parametrize("options", [dict(), dict(exact=1), dict(variant=2), dict(pdl=1), dict(variant=4, exact=1)],
                         ids=lambda o: "-".join(f"{k}{v}" for k, v in o.items()) or "default")
def test_graph_replay_is_bit_identical_to_stream_launches(tmp_path, options):
    import psvwave_b200
    ini = str(tmp_path / "g.ini")
    config(ini)
    o = OracleModel(ini)
    meds = [random_medium(o.nx, o.nz, seed=s) for s in (3, 4)]
    o.set_parameter_fields(*random_medium(o.nx, o.nz, seed=5))
    for s in range(o.n_shots):
        o.forward_simulate(s, False)
    obs = [np.array(a) for a in o.get_synthetic_data()]
    plain, graph = psvwave_b200.fdModel(ini), psvwave_b200.fdModel(ini)
    for k, v in options.items():
        plain._set_option(k, v)
        graph._set_option(k, v)
    plain._set_option("graph", 0)
    graph._set_option("graph", 1)
    # round 0 captures, round 1 replays with another medium and other observed data, round 2 replays the first again
    for rnd, (med, scale) in enumerate([(meds[0], 1.0), (meds[1], 0.7), (meds[0], 1.0)]):
        ob = [scale * a for a in obs]
        run(plain, med, ob)
        run(graph, med, ob)
        for a in ARRAYS:
            assert np.array_equal(gpu_array(graph, a), gpu_array(plain, a)), f"round {rnd}: {a}"
        assert np.array_equal(graph.get_gradient_vector(), plain.get_gradient_vector())
        assert graph.misfit == plain.misfit
        assert graph._get_stat("launches") == plain._get_stat("launches")
    assert graph._get_option("graphs_cached") == 2 * graph.n_shots  # forward(store) and adjoint per shot
    if options.get("exact"):
        o.set_parameter_fields(*meds[0])
        o.set_observed_data(*obs)
        gradient_workflow_store(o)
        for a in ARRAYS:
            assert np.array_equal(gpu_array(graph, a), o.array(a)), a
    # an option that changes the launches gets its own graphs; switching the option off drops them all
    graph._set_option("fused_strip", 4)
    plain._set_option("fused_strip", 4)
    run(plain, meds[1], obs)
    run(graph, meds[1], obs)
    assert np.array_equal(graph.get_gradient_vector(), plain.get_gradient_vector())
    graph._set_option("graph", 0)
    assert graph._get_option("graphs_cached") == 0
    run(graph, meds[1], obs)
    assert np.array_equal(graph.get_gradient_vector(), plain.get_gradient_vector())


def test_graph_replay_of_a_model_ensemble(tmp_path):
    import psvwave_b200
    ini = str(tmp_path / "g.ini")
    config(ini)
    ms = [psvwave_b200.fdModel(ini) for _ in range(3)]
    ref = [psvwave_b200.fdModel(ini) for _ in range(3)]
    ms[0]._set_option("graph", 1)
    ref[0]._set_option("graph", 0)
    for rnd in range(2):
        for i, (a, b) in enumerate(zip(ms, ref)):
            med = random_medium(a.nx, a.nz, seed=10 * rnd + i)
            a.set_parameter_fields(*med)
            b.set_parameter_fields(*med)
        psvwave_b200.forward_batch(ms, 1, False)
        psvwave_b200.forward_batch(ref, 1, False)
        for a, b in zip(ms, ref):
            for x, y in zip(a.get_synthetic_data(), b.get_synthetic_data()):
                assert np.array_equal(x, y)
    assert ms[0]._get_option("graphs_cached") == 1


def test_auto_mode_captures_a_loop_when_it_comes_back(tmp_path):
    import psvwave_b200
    ini = str(tmp_path / "g.ini")
    config(ini)
    plain, auto = psvwave_b200.fdModel(ini), psvwave_b200.fdModel(ini)
    plain._set_option("graph", 0)
    assert auto._get_option("graph") == -1
    med = random_medium(auto.nx, auto.nz, seed=8)
    plain.set_parameter_fields(*random_medium(auto.nx, auto.nz, seed=9))
    for s in range(plain.n_shots):
        plain.forward_simulate(s, False)
    obs = plain.get_synthetic_data()
    cached = []
    for rnd in range(3):
        run(plain, med, obs)
        run(auto, med, obs)
        cached.append(auto._get_option("graphs_cached"))
        for a in ARRAYS:
            assert np.array_equal(gpu_array(auto, a), gpu_array(plain, a)), f"round {rnd}: {a}"
    assert cached == [0, 2 * auto.n_shots, 2 * auto.n_shots]
