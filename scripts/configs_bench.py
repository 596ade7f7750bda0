"""Rows of BASELINE.md's table that bench.py does not print (run under gpurun, one GPU):
  C1  configs/default_testing_configuration.ini (250x150, nt=8000): forward (no store), forward (store),
      adjoint, gradient per shot -- on the B200 and, beside it, the unmodified reference core
      (oracle/_ref, OpenMP on all host threads) on the same box in the same run;
  C2  Marmousi-scale forward simulation (bench.py's grid and medium, 2 stacked sources), no store / store.
One JSON line per config.  GPU times are CUDA-event times of the whole time loop ("elapsed_ms").
"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402

import bench  # noqa: E402
import psvwave_b200  # noqa: E402
from helpers import C1_INI, notebook2_models  # noqa: E402
from oracle import oracle  # noqa: E402  (cpu baseline only)


def gpu_times(m, obs):
    pts = m.nx * m.nz * m.nt
    out = {}
    m.forward_simulate(0, False)  # warm-up
    m.forward_simulate(0, False)
    out["forward_nostore_updates_per_s"] = pts / (m._get_stat("elapsed_ms") * 1e-3)
    m.forward_simulate(0, True)
    out["forward_store_updates_per_s"] = pts / (m._get_stat("elapsed_ms") * 1e-3)
    if obs is not None:
        m.set_observed_data(*obs)
    m.calculate_l2_misfit()
    m.calculate_l2_adjoint_sources()
    m.reset_kernels()
    m.adjoint_simulate(0)
    out["adjoint_updates_per_s"] = pts / (m._get_stat("elapsed_ms") * 1e-3)
    t0 = time.perf_counter()
    m.reset_kernels()
    m.forward_simulate(0, True)
    m.calculate_l2_misfit()
    m.calculate_l2_adjoint_sources()
    m.adjoint_simulate(0)
    m.map_kernels_to_velocity()
    g = m.get_gradient_vector()
    out["gradient_s_per_shot"] = time.perf_counter() - t0
    out["gradient_norm"] = float(np.linalg.norm(g))
    return out


def cpu_times(m, obs):
    pts = m.nx * m.nz * m.nt
    out = {}

    def timed(f):
        best = 1e30
        for _ in range(2):
            t0 = time.perf_counter()
            f()
            best = min(best, time.perf_counter() - t0)
        return best

    out["forward_nostore_updates_per_s"] = pts / timed(lambda: m.forward_simulate(0, False))
    out["forward_store_updates_per_s"] = pts / timed(lambda: m.forward_simulate(0, True))
    m.set_observed_data(*obs)
    m.calculate_l2_misfit()
    m.calculate_l2_adjoint_sources()
    m.reset_kernels()
    out["adjoint_updates_per_s"] = pts / timed(lambda: m.adjoint_simulate(0))
    return out


def main():
    # ---- C1 ----
    g = psvwave_b200.fdModel(C1_INI)
    start, target = notebook2_models(g)
    g.set_parameter_fields(*target)
    g.forward_simulate(0, False)
    obs = g.get_synthetic_data()
    g.set_parameter_fields(*start)
    c1 = {"config": "C1 default ini", "grid": [g.nx, g.nz], "nt": g.nt, "b200": gpu_times(g, obs)}
    if oracle.ref_available():
        r = oracle.RefModel(C1_INI)
        r.set_parameter_fields(*start)
        c1["cpu_reference"] = dict(cpu_times(r, obs), cores=bench.omp_threads(), build="g++ -O3 -fopenmp (reference Release flags)")
        r.map_kernels_to_velocity()
        # 1-thread row: a fresh process so that OMP_NUM_THREADS takes effect
        import subprocess
        code = ("import sys,time; sys.path.insert(0,%r); from oracle import oracle; m=oracle.RefModel(%r); "
                "t=time.perf_counter(); m.forward_simulate(0, False); print(m.nx*m.nz*m.nt/(time.perf_counter()-t))" % (ROOT, C1_INI))
        out = subprocess.run([sys.executable, "-c", code], env=dict(os.environ, OMP_NUM_THREADS="1"), capture_output=True, text=True)
        try:
            c1["cpu_reference_1_thread"] = {"forward_nostore_updates_per_s": float(out.stdout.strip().splitlines()[-1])}
        except Exception:
            c1["cpu_reference_1_thread"] = {"error": out.stderr[-200:]}
    print(json.dumps(c1), flush=True)
    del g

    # ---- C2 forward ----
    import tempfile
    geom = dict(bench.GEOM)
    cfg = bench.workload_config(geom["nt"], 1, geom)
    ini = os.path.join(tempfile.mkdtemp(prefix="psv_c2_"), "c2.ini")
    bench.write_ini(ini, cfg)
    m = psvwave_b200.fdModel(ini)
    m.set_parameter_fields(*bench.marmousi_like(m.nx, m.nz))
    pts = m.nx * m.nz * m.nt
    m.forward_simulate(0, False)
    m.forward_simulate(0, False)
    a = pts / (m._get_stat("elapsed_ms") * 1e-3)
    m.forward_simulate(0, True)
    b = pts / (m._get_stat("elapsed_ms") * 1e-3)
    print(json.dumps({"config": "C2 Marmousi-scale forward, 2 stacked sources, 300 receivers", "grid": [m.nx, m.nz],
                      "nt": m.nt, "snapshot_interval": m.snapshot_interval,
                      "b200": {"forward_nostore_updates_per_s": a, "forward_store_updates_per_s": b,
                               "seconds_per_shot_store": pts / b}}), flush=True)


if __name__ == "__main__":
    main()
