"""Host logic of the domain split on CPU: psvwave_b200.slab.SlabDomain at world_size 2 and 3 over
gloo (one slab per rank, halos as point-to-point messages) and with several slabs in one process,
driving the numpy test double of tests/_slab_fake.py.  Every decomposition must reproduce the
single-slab run BIT-IDENTICALLY: partition arithmetic, who sends which rows to whom, how many steps
fit between two exchanges (halo_rows / 4), the seismogram merge by owning slab, and the gathers.
The physics of the split (CUDA slabs against the whole-grid run and the oracle) is tests/test_gpu_slab.py.
"""
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

from conftest import ROOT
from _gloo_slab_worker import run
from _slab_fake import ToySlab, partition
from helpers import make_config
from psvwave_b200.slab import SlabDomain


def _ini(tmp_path):
    ini = str(tmp_path / "slab.ini")
    make_config(ini, nt=60, nx_inner=90, nz_inner=11, nx_inner_boundary=2, nz_inner_boundary=2, np_boundary=5,
                n_sources=3, n_shots=2, ix_sources=[4, 44, 45], iz_sources=[3, 4, 5], moment_angles=[0.0, 1.0, 2.0],
                which_source_to_fire_in_which_shot=[[0, 1], [2]], nr=6, ix_receivers=[0, 44, 45, 46, 20, 89],
                iz_receivers=[1, 2, 3, 4, 5, 6], snapshot_interval=5)
    return ini


def test_partition_covers_the_grid_once():
    for nx in (100, 101, 8100):
        for n in (1, 2, 3, 8):
            for halo in (4, 16):
                parts = [partition(nx, r, n, halo) for r in range(n)]
                owned = [(x0 + lo, x1 - hi) for x0, x1, lo, hi in parts]
                assert owned[0][0] == 0 and owned[-1][1] == nx
                assert all(a[1] == b[0] for a, b in zip(owned, owned[1:]))
                assert parts[0][2] == 0 and parts[-1][3] == 0


@pytest.mark.parametrize("n_slabs,halo_steps", [(2, 1), (3, 2), (4, 4)])
def test_in_process_slabs_equal_single_slab(tmp_path, n_slabs, halo_steps):
    ini = _ini(tmp_path)
    one = run(SlabDomain(ini, n_slabs=1, halo_steps=halo_steps, slab_factory=ToySlab, rank=0, world_size=1))
    split = SlabDomain(ini, n_slabs=n_slabs, halo_steps=halo_steps, slab_factory=ToySlab, rank=0, world_size=1)
    got = run(split)
    assert np.abs(one["field"]).max() > 0 and np.abs(one["ux"]).max() > 0
    for k in one:
        assert np.array_equal(got[k], one[k]), k
    assert split.exchanges == (split.nt - 1) // halo_steps


@pytest.mark.parametrize("world,halo_steps", [(2, 4), (3, 1)])
def test_one_slab_per_rank_over_gloo_equals_single_slab(tmp_path, world, halo_steps):
    ini = _ini(tmp_path)
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    procs = []
    for rank in range(world):
        env = dict(os.environ, RANK=str(rank), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port),
                   OMP_NUM_THREADS="1")
        procs.append(subprocess.Popen([sys.executable, os.path.join(ROOT, "tests", "_gloo_slab_worker.py"), ini,
                                       str(tmp_path), str(halo_steps)], env=env, stdout=subprocess.PIPE,
                                      stderr=subprocess.STDOUT))
    for p in procs:
        out, _ = p.communicate(timeout=600)
        assert p.returncode == 0, out.decode()[-2000:]
    one = run(SlabDomain(ini, n_slabs=1, slab_factory=ToySlab, rank=0, world_size=1))
    ranks = [np.load(tmp_path / f"rank{r}.npz") for r in range(world)]
    assert ranks[0]["owned"][0] == 0 and ranks[-1]["owned"][1] == 100
    for r in ranks:  # every rank ends with the whole-grid results
        for k in one:
            assert np.array_equal(r[k], one[k]), k
