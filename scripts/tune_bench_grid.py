"""The tuner on the bench workload (3100x1320, 300 receivers, 2 stacked sources): forward(store) + adjoint per-step times
with the formula geometry (tune=0) and the measured one (default)."""
import json, os, sys, tempfile
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import bench, psvwave_b200
cfg = bench.workload_config(1000, 2)
ini = os.path.join(tempfile.mkdtemp(), "w.ini"); bench.write_ini(ini, cfg)
med = bench.marmousi_like(cfg["nx_inner"] + 100, cfg["nz_inner"] + 100)
for opts in ({"tune": 0}, {}, {"tune": 0, "pipe_warps": 3, "fused_strip": 84}):
    m = psvwave_b200.fdModel(ini)
    for k, v in opts.items(): m._set_option(k, v)
    m.set_parameter_fields(*med)
    z = np.zeros((m.n_shots, m.nr, m.nt)); m.set_observed_data(z, z)
    out = {"options": opts}
    for rep in range(2):
        m.forward_simulate(0, True); out["fwd_us"] = m._get_stat("elapsed_ms") * 1e3 / m.nt
        m.calculate_l2_adjoint_sources(); m.reset_kernels()
        m.adjoint_simulate(0); out["adj_us"] = m._get_stat("elapsed_ms") * 1e3 / m.nt
    out.update(warps=m._get_stat("tuned_warps"), strip=m._get_stat("tuned_strip"), tune_ms=m._get_stat("tune_ms"))
    print(json.dumps(out), flush=True)
