"""Step-kernel geometry on the shape of ONE slab of the 8-way split of the 8100x4100 grid (1044 x 4100 points):
forward (no store) time per step for pipe_warps x fused_strip (run under gpurun, one GPU)."""
import json
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import psvwave_b200  # noqa: E402
from helpers import make_config  # noqa: E402

nx_inner = int(sys.argv[1]) if len(sys.argv) > 1 else 944
nz_inner = int(sys.argv[2]) if len(sys.argv) > 2 else 4000
ini = os.path.join(tempfile.mkdtemp(prefix="psv_sweep_"), "s.ini")
make_config(ini, nt=200, nx_inner=nx_inner, nz_inner=nz_inner, nx_inner_boundary=10, nz_inner_boundary=10, np_boundary=50,
            np_factor=0.0075, dx=2.5, dz=2.5, dt=2.5e-4, peak_frequency=8.0, n_sources=1, n_shots=1,
            ix_sources=[nx_inner // 2], iz_sources=[20], moment_angles=[90.0], which_source_to_fire_in_which_shot=[[0]],
            nr=8, ix_receivers=[int(x) for x in np.linspace(100, nx_inner - 100, 8)], iz_receivers=[10] * 8,
            snapshot_interval=50, npx=1, npz=1)
m = psvwave_b200.fdModel(ini)
pts = m.nx * m.nz
for warps in (0, 3, 4, 8):
    for strip in (0, 24, 32, 48, 64, 96, 131, 180, 261):
        m._set_option("pipe_warps", warps)
        m._set_option("fused_strip", strip)
        m.forward_simulate(0, False)
        m.forward_simulate(0, False)
        ms = m._get_stat("elapsed_ms")
        print(json.dumps({"grid": [m.nx, m.nz], "pipe_warps": warps, "fused_strip": strip,
                          "us_per_step": ms * 1e3 / m.nt, "G_updates_per_s": pts * m.nt / ms / 1e6}), flush=True)
