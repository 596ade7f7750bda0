"""C1 (the reference's default 250x150 grid, nt 8000) with the time loop launched kernel by kernel, with programmatic
dependent launch, and replayed from a CUDA graph (option "graph"): device time per march and wall time per gradient.
Run under gpurun; prints one JSON line per option set.
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import psvwave_b200  # noqa: E402
from helpers import C1_INI, notebook2_models  # noqa: E402


def measure(options):
    m = psvwave_b200.fdModel(C1_INI)
    for k, v in options.items():
        m._set_option(k, v)
    start, target = notebook2_models(m)
    m.set_parameter_fields(*target)
    m.forward_simulate(0, False)
    obs = m.get_synthetic_data()
    m.set_parameter_fields(*start)
    m.set_observed_data(*obs)
    pts = m.nx * m.nz * m.nt
    out = {"options": options}
    t0 = time.perf_counter()
    m.forward_simulate(0, True)
    out["first_forward_store_wall_s"] = time.perf_counter() - t0  # includes capture + instantiation with "graph"
    for name, call in (("forward_nostore", lambda: m.forward_simulate(0, False)), ("forward_store", lambda: m.forward_simulate(0, True))):
        call()
        call()
        out[name + "_updates_per_s"] = pts / (m._get_stat("elapsed_ms") * 1e-3)
        out[name + "_us_per_step"] = m._get_stat("elapsed_ms") * 1e3 / m.nt
    m.calculate_l2_misfit()
    m.calculate_l2_adjoint_sources()
    m.reset_kernels()
    m.adjoint_simulate(0)
    m.reset_kernels()
    m.adjoint_simulate(0)
    out["adjoint_updates_per_s"] = pts / (m._get_stat("elapsed_ms") * 1e-3)
    out["adjoint_us_per_step"] = m._get_stat("elapsed_ms") * 1e3 / m.nt
    walls = []
    for _ in range(3):
        t0 = time.perf_counter()
        m.reset_kernels()
        m.forward_simulate(0, True)
        m.calculate_l2_misfit()
        m.calculate_l2_adjoint_sources()
        m.adjoint_simulate(0)
        m.map_kernels_to_velocity()
        g = m.get_gradient_vector()
        walls.append(time.perf_counter() - t0)
    out["gradient_wall_s_per_shot"] = min(walls)
    out["gradient_norm"] = float(np.linalg.norm(g))
    return out


if __name__ == "__main__":
    if len(sys.argv) > 1:  # option sets as JSON objects on the command line
        for a in sys.argv[1:]:
            print(json.dumps(measure(json.loads(a))), flush=True)
        sys.exit(0)
    for opts in ({"graph": 0}, {"graph": 0, "pdl": 1}, {"graph": 1}, {"graph": 1, "pdl": 1}, {}):
        print(json.dumps(measure(opts)), flush=True)
