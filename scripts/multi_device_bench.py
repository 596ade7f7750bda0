"""ONE fdModel object over the GPUs of the box (psv_set_devices), driven by the reference's own user loop from a plain
single-process script -- no torchrun, no PyTorch: wall time of a whole gradient (all shots) on 1, 2, 4, 8 devices.

    python scripts/multi_device_bench.py [nt] [shots_per_gpu]      (run under gpurun --gpus N)

Workload: bench.py's C2 geometry (3100x1320, 2 stacked sources per shot, 300 receivers, si=20), n_shots = shots_per_gpu
x the largest device count (so every device count runs the SAME total work: strong scaling over shots).
Prints one JSON line per device count; `speedup` is against the 1-device line of the same run."""
import json
import os
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import psvwave_b200  # noqa: E402
from oracle.oracle import write_ini  # noqa: E402  (ini writer only)

nt = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
per_gpu = int(sys.argv[2]) if len(sys.argv) > 2 else 1
n_dev = psvwave_b200.device_count()
counts = [n for n in (1, 2, 4, 8) if n <= n_dev]
n_shots = per_gpu * counts[-1]
cfg = bench.workload_config(nt, n_shots)
ini = os.path.join(tempfile.mkdtemp(prefix="psv_md_"), "w.ini")
write_ini(ini, cfg)
vp, vs, rho = bench.marmousi_like(cfg["nx_inner"] + 2 * cfg["np_boundary"], cfg["nz_inner"] + 2 * cfg["np_boundary"])
base = None
ref_g = None
for n in counts:
    m = psvwave_b200.fdModel(ini)
    m.set_devices(list(range(n)))
    m.set_parameter_fields(vp, vs, rho)
    zero = np.zeros((m.n_shots, m.nr, m.nt))
    m.set_observed_data(zero, zero)
    walls = []
    for rep in range(3):
        t0 = time.perf_counter()
        # the reference's documented call order (docs-source/source/typical-workflow.rst:5-13), unchanged
        for s in range(m.n_shots):
            m.forward_simulate(s)
        m.calculate_l2_misfit()
        m.calculate_l2_adjoint_sources()
        m.reset_kernels()
        for s in range(m.n_shots):
            m.adjoint_simulate(s)
        m.map_kernels_to_velocity()
        g = m.get_gradient_vector()
        walls.append(time.perf_counter() - t0)
    wall = min(walls)
    if base is None:
        base, ref_g = wall, g
    out = dict(devices=n, n_shots=m.n_shots, nt=nt, grid=[m.nx, m.nz], wall_s_per_gradient=wall,
               updates_per_s=2.0 * m.n_shots * m.nx * m.nz * m.nt / wall, speedup=base / wall,
               allreduce_ms=m._get_stat("allreduce_ms") if n > 1 else 0.0,
               rel_l2_gradient_vs_1_device=float(np.linalg.norm(g - ref_g) / np.linalg.norm(ref_g)))
    print(json.dumps(out), flush=True)
    del m
