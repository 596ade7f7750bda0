"""ONE fdModel object over several GPUs (psv_set_devices, include/psvwave_b200.h): the reference's own call order
(docs-source/source/typical-workflow.rst:5-13; fdModel::run_model, src/fdModel.cpp:1291-1308) on an object that spans
n devices must give what the same calls give on one device -- seismograms and misfit bit for bit (each shot runs the
same kernels on one GPU, the misfit is summed on the primary in the reference's order), kernels and gradient to
<= 1e-13 (the cross-GPU sum re-associates the additions of src/fdModel.cpp:832-858) -- and both must match the oracle.

Real multi-GPU cases need >= 2 devices (gpurun --gpus N; skipped otherwise).  On a single-GPU box the same machinery
(worker threads, queued simulate calls, gather of the synthetics, the "sum over replicas" invariant) is exercised with
several replicas on device 0 (PSVWAVE_ALLOW_DUPLICATE_DEVICES=1), where a device add stands in for the NCCL all-reduce.
"""
import os

import numpy as np
import pytest

from conftest import SMALL_INI, rel_l2
from helpers import DYNAMIC, gpu_array, gradient_workflow, ragged_config, random_medium
from oracle.oracle import OracleModel

pytestmark = pytest.mark.gpu

KERNELS = ("lambda_kernel", "mu_kernel", "density_l_kernel", "vp_kernel", "vs_kernel", "density_v_kernel")


def n_devices():
    import psvwave_b200
    return psvwave_b200.device_count()


def device_sets():
    n = n_devices()
    sets = [pytest.param([0, 0], id="2-replicas-on-gpu0"), pytest.param([0, 0, 0], id="3-replicas-on-gpu0")]
    for k in (2, 3, 4, 8):
        ids = list(range(k))
        sets.append(pytest.param(ids, id=f"{k}-gpus",
                                 marks=pytest.mark.skipif(n < k, reason=f"needs {k} CUDA devices")))
    return sets


@pytest.fixture(autouse=True)
def _allow_replicas_on_one_gpu(monkeypatch):
    monkeypatch.setenv("PSVWAVE_ALLOW_DUPLICATE_DEVICES", "1")


def prepare(ini, medium, obs, devices=None, **options):
    import psvwave_b200
    m = psvwave_b200.fdModel(ini)
    for k, v in options.items():
        m._set_option(k, v)
    if devices:
        m.set_devices(devices)
        assert m.devices() == devices
    m.set_parameter_fields(*medium)
    m.set_observed_data(*obs)
    return m


@pytest.mark.parametrize("devices", device_sets())
@pytest.mark.parametrize("exact", [1, 0])
def test_reference_call_order_on_a_device_group(tmp_path, devices, exact):
    ini = str(tmp_path / "ragged.ini")
    ragged_config(ini)
    o = OracleModel(ini)
    medium = random_medium(o.nx, o.nz, seed=11)
    rng = np.random.default_rng(5)
    obs = (1e-3 * rng.standard_normal((o.n_shots, o.nr, o.nt)), 1e-3 * rng.standard_normal((o.n_shots, o.nr, o.nt)))
    o.set_parameter_fields(*medium)
    o.set_observed_data(*obs)
    gradient_workflow(o)

    one = prepare(ini, medium, obs, exact=exact)
    gradient_workflow(one)
    grp = prepare(ini, medium, obs, devices=devices, exact=exact)
    gradient_workflow(grp)  # the unchanged user loop
    assert grp._get_stat("devices") == len(devices)
    assert grp._get_stat("allreduces") + grp._get_stat("local_sums") == 1

    for a, b in zip(grp.get_synthetic_data(), one.get_synthetic_data()):
        assert np.array_equal(a, b)
    assert grp.misfit == one.misfit
    for name in ("a_stf_ux", "a_stf_uz"):
        assert np.array_equal(gpu_array(grp, name), gpu_array(one, name))
    for f, a, b in zip(DYNAMIC, grp.get_snapshots(), one.get_snapshots()):
        assert np.array_equal(a, b), f
    for name in KERNELS:
        assert rel_l2(gpu_array(grp, name), gpu_array(one, name)) <= 1e-13, name
        assert rel_l2(gpu_array(grp, name), o.array(name)) <= 1e-10, name
    assert rel_l2(grp.get_gradient_vector(), one.get_gradient_vector()) <= 1e-13
    assert rel_l2(grp.get_gradient_vector(), o.get_gradient_vector()) <= 1e-10
    if exact:
        assert np.array_equal(grp.get_synthetic_data()[0], o.get_synthetic_data()[0])
        assert abs(grp.misfit - o.misfit) <= 1e-13 * o.misfit  # summation order (device partial sums)
    # the state the last simulate call left behind (adjoint of the last shot, on that shot's GPU)
    assert np.array_equal(gpu_array(grp, "vx"), gpu_array(one, "vx"))
    # the kernels keep accumulating across further adjoint runs until reset_kernels (src/fdModel.cpp:832-858)
    grp.adjoint_simulate(grp.n_shots - 1)
    one.adjoint_simulate(one.n_shots - 1)
    assert rel_l2(gpu_array(grp, "mu_kernel"), gpu_array(one, "mu_kernel")) <= 1e-13
    grp.map_kernels_to_velocity()
    grp.map_kernels_to_velocity()  # idempotent: the sum is taken once
    one.map_kernels_to_velocity()
    assert rel_l2(grp.get_gradient_vector(), one.get_gradient_vector()) <= 1e-13


@pytest.mark.parametrize("devices", device_sets()[:1] + device_sets()[3:4])
def test_group_run_gradient_descent_copy_and_return_to_one_device(golden, devices):
    g = golden("small_multishot")
    medium = (g["vp"], g["vs"], g["rho"])
    obs = (g["obs_ux"], g["obs_uz"])
    one = prepare(SMALL_INI, medium, obs, exact=1)
    grp = prepare(SMALL_INI, medium, obs, devices=devices, exact=1)
    # fused per-shot gradient (run_model semantics) and the device-resident descent loop
    assert abs(grp.run_gradient(True) - one.run_gradient(True)) <= 1e-13 * one.misfit
    assert rel_l2(grp.get_gradient_vector(), g["gradient"]) <= 1e-13
    m1 = one.gradient_descent(2, 0.25)
    m2 = grp.gradient_descent(2, 0.25)
    assert np.allclose(m1, m2, rtol=1e-12, atol=0)
    assert rel_l2(grp.get_model_vector(), one.get_model_vector()) <= 1e-14
    for a, b in zip(grp.get_parameter_fields(), one.get_parameter_fields()):
        assert rel_l2(a, b) <= 1e-14
    # deep copy of a group (src/fdModel.cpp:64-97): independent objects with the same state
    twin = grp.copy()
    assert twin.devices() == devices
    mv = twin.get_model_vector()
    twin.set_model_vector(mv * 1.01)
    assert np.array_equal(grp.get_model_vector(), mv)
    for s in range(twin.n_shots):
        twin.forward_simulate(s, False)
    assert not np.array_equal(twin.get_synthetic_data()[0], grp.get_synthetic_data()[0])
    # back to one device: everything the group computed stays with the object
    ux_before, grad_before = grp.get_synthetic_data()[0], grp.get_gradient_vector()
    lam_before = gpu_array(grp, "lambda_kernel")
    grp.set_devices([devices[0]])
    assert grp.devices() == [devices[0]] and grp._get_stat("launches") >= 0
    assert np.array_equal(grp.get_synthetic_data()[0], ux_before)
    assert np.array_equal(grp.get_gradient_vector(), grad_before)
    assert np.array_equal(gpu_array(grp, "lambda_kernel"), lam_before)


def test_group_reports_failures_of_queued_calls_and_bad_arguments(golden):
    import psvwave_b200
    g = golden("small_multishot")
    m = prepare(SMALL_INI, (g["vp"], g["vs"], g["rho"]), (g["obs_ux"], g["obs_uz"]), devices=[0, 0])
    with pytest.raises(ValueError):
        m.forward_simulate(m.n_shots)  # checked before anything is queued
    m.adjoint_simulate(1)  # no stored forward field: fails on the worker ...
    with pytest.raises(RuntimeError, match="forward"):
        m.synchronize()  # ... and is reported by the next call that waits
    m.map_kernels_to_velocity()
    assert m._get_stat("local_sums") == 0  # the failed run raised no "kernels are spread" flag: nothing to sum
    assert m._get_option("exact") == 0  # option reads wait for queued runs on a group handle too
    gradient_workflow(m)  # the object stays usable
    assert rel_l2(m.get_gradient_vector(), g["gradient"]) <= 1e-10
    with pytest.raises(ValueError):
        m.set_devices([])
    with pytest.raises(ValueError):
        m.set_devices([psvwave_b200.device_count()])
    with pytest.raises(RuntimeError, match="several GPUs"):
        psvwave_b200.forward_batch([m], 0, False)


@pytest.mark.skipif("n_devices() < 2", reason="needs 2 CUDA devices")
def test_object_moves_to_the_first_listed_device(golden):
    g = golden("small_multishot")
    m = prepare(SMALL_INI, (g["vp"], g["vs"], g["rho"]), (g["obs_ux"], g["obs_uz"]))
    m.forward_simulate(0)
    ux = m.get_synthetic_data()[0]
    m.set_devices([1, 0])
    assert m.devices() == [1, 0]
    assert np.array_equal(m.get_synthetic_data()[0], ux)
    gradient_workflow(m)
    assert rel_l2(m.get_gradient_vector(), g["gradient"]) <= 1e-10
    assert m._get_stat("allreduces") == 1


@pytest.mark.skipif("n_devices() < 2", reason="needs 2 CUDA devices")
def test_environment_variable_spreads_unchanged_scripts(golden, monkeypatch):
    import psvwave_b200
    g = golden("small_multishot")
    monkeypatch.setenv("PSVWAVE_DEVICES", "all")
    m = psvwave_b200.fdModel(SMALL_INI)
    assert m.devices() == list(range(psvwave_b200.device_count()))
    m.set_parameter_fields(g["vp"], g["vs"], g["rho"])
    m.set_observed_data(g["obs_ux"], g["obs_uz"])
    gradient_workflow(m)
    assert rel_l2(m.get_gradient_vector(), g["gradient"]) <= 1e-10


def test_group_run_model_and_receiver_files(golden, tmp_path, monkeypatch):
    """fdModel::run_model (src/fdModel.cpp:1291-1308) and the receiver files (:971-1006, :1053-1126) on an object that
    spans devices: same misfit / gradient / files as on one device; loaded observed data reach every replica."""
    import psvwave_b200
    from oracle.oracle import config_from_ini, write_ini
    g = golden("small_multishot")
    monkeypatch.chdir(tmp_path)
    files = {}
    results = []
    for tag, devices in (("one", None), ("grp", [0, 0, 0])):
        folder = tmp_path / tag
        folder.mkdir()
        ini = str(tmp_path / f"{tag}.ini")
        write_ini(ini, dict(config_from_ini(SMALL_INI), observed_data_folder=str(folder), stf_folder=str(folder)))
        m = prepare(ini, (g["vp"], g["vs"], g["rho"]), (g["obs_ux"], g["obs_uz"]), devices=devices, exact=1)
        m.run_model(False, True)
        results.append((m.misfit, m.get_gradient_vector()))
        m.write_receivers()
        files[tag] = {p.name: p.read_bytes() for p in sorted(folder.iterdir())}
        # observed := the synthetics just written; the misfit of the same synthetics must then vanish on every replica
        m.load_receivers(False)
        for s in range(m.n_shots):
            m.forward_simulate(s, False)
        assert m.calculate_l2_misfit_shot(m.n_shots - 1) <= 1e-20 * results[-1][0]
        m.calculate_l2_misfit()
        assert m.misfit <= 1e-20 * results[-1][0]
    assert files["one"].keys() == files["grp"].keys() and len(files["one"]) >= 2
    assert files["one"] == files["grp"]
    assert results[0][0] == results[1][0]
    assert rel_l2(results[1][1], results[0][1]) <= 1e-13
