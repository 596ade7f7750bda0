"""Multi-process path on CPU: psvwave_b200.parallel.ShotParallel at world_size 2 and 3 over gloo.

The shot map, the per-rank forward -> misfit -> adjoint loop, the ONE all-reduce of the
[lambda | mu | density_l] kernels and the scalar misfit all-reduce are host logic; here they
drive the CPU oracle in place of the CUDA model (tests/_gloo_worker.py) and must reproduce the
single-process gradient of all shots (the reference's run_model, src/fdModel.cpp:1291-1308)
up to the changed summation order across shots.
"""
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

from conftest import ROOT, SMALL_INI, rel_l2
from helpers import gradient_workflow
from oracle.oracle import OracleModel
from psvwave_b200.parallel import shots_of_rank


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_shot_map_is_a_partition():
    for n_shots in (1, 2, 5, 16):
        for world in (1, 2, 3, 8):
            dealt = sorted(s for r in range(world) for s in shots_of_rank(n_shots, r, world))
            assert dealt == list(range(n_shots))


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_gradient_equals_single_process(tmp_path, golden, world):
    g = golden("small_multishot")
    obs = str(tmp_path / "obs.npz")
    np.savez(obs, vp=g["vp"], vs=g["vs"], rho=g["rho"], obs_ux=g["obs_ux"], obs_uz=g["obs_uz"])
    port = _free_port()
    procs = []
    for rank in range(world):
        env = dict(os.environ, RANK=str(rank), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1",
                   MASTER_PORT=str(port), OMP_NUM_THREADS="2")
        procs.append(subprocess.Popen([sys.executable, os.path.join(ROOT, "tests", "_gloo_worker.py"), SMALL_INI,
                                       str(tmp_path), obs], env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT))
    for p in procs:
        out, _ = p.communicate(timeout=600)
        assert p.returncode == 0, out.decode()[-2000:]

    ref = OracleModel(SMALL_INI)
    ref.set_parameter_fields(g["vp"], g["vs"], g["rho"])
    for s in range(ref.n_shots):
        ref.forward_simulate(s, True)
    ref.set_observed_data(g["obs_ux"], g["obs_uz"])
    gradient_workflow(ref)
    ref_ux, _ = ref.get_synthetic_data()
    ranks = [np.load(tmp_path / f"rank{r}.npz") for r in range(world)]
    seen = []
    for r in ranks:
        # every rank ends with the full gradient and the global misfit
        assert rel_l2(r["gradient"], ref.get_gradient_vector()) <= 1e-13
        assert rel_l2(r["lambda_kernel"], ref.array("lambda_kernel")) <= 1e-13
        assert rel_l2(r["vs_kernel"], ref.array("vs_kernel")) <= 1e-13
        assert float(r["misfit"]) == pytest.approx(ref.misfit, rel=1e-13)
        assert np.array_equal(r["gradient"], ranks[0]["gradient"])
        # seismograms stay with the owning rank, bit-identical to the single-process run
        for s in r["shots"]:
            assert np.array_equal(r["ux"][s], ref_ux[s])
            seen.append(int(s))
    assert sorted(seen) == list(range(ref.n_shots))
