"""What ncu looks at (round 2): ONE forward(store) + adjoint of the bench workload's shot 0 with nt = 60, after two
untimed passes (launch-geometry tuning, module loading), bracketed by cudaProfilerStart/Stop:

    ncu --profile-from-start off ... python scripts/ncu_target.py

Launches inside the bracket: 3 snapshot + 57 plain step launches, the receiver prefix, 3 correlation + 57 plain."""
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402  (cudaProfilerStart/Stop only)

import bench  # noqa: E402
import psvwave_b200  # noqa: E402

nt = int(os.environ.get("PSV_NCU_NT", "60"))
cfg = bench.workload_config(nt, 2)
ini = os.path.join(tempfile.mkdtemp(prefix="psv_ncu_"), "w.ini")
bench.write_ini(ini, cfg)
m = psvwave_b200.fdModel(ini)
for kv in sys.argv[1:]:
    k, v = kv.split("=")
    m._set_option(k, int(v))
m.set_parameter_fields(*bench.marmousi_like(m.nx, m.nz))
rng = np.random.default_rng(0)
m.set_observed_data(1e-6 * rng.standard_normal((m.n_shots, m.nr, m.nt)), 1e-6 * rng.standard_normal((m.n_shots, m.nr, m.nt)))


def gradient():
    m.reset_kernels()
    m.forward_simulate(0, True)
    m.calculate_l2_adjoint_sources()
    m.adjoint_simulate(0)


gradient()
gradient()
m.synchronize()
torch.cuda.profiler.start()
gradient()
m.synchronize()
torch.cuda.profiler.stop()
print("tuned_warps", m._get_stat("tuned_warps"), "tuned_strip", m._get_stat("tuned_strip"), "elapsed_ms", m._get_stat("elapsed_ms"))
