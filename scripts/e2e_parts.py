"""Host-side parts of bench.py's end-to-end step on the bench grid: set_parameter_fields from pinned memory,
get_gradient_vector to the host (run under gpurun)."""
import os, sys, tempfile, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import bench, psvwave_b200
cfg = bench.workload_config(100, 16)
ini = os.path.join(tempfile.mkdtemp(), "w.ini"); bench.write_ini(ini, cfg)
m = psvwave_b200.fdModel(ini)
med = bench.marmousi_like(m.nx, m.nz)
pinned = [bench.pinned_copy(a) for a in med]
host = [t.numpy() for t in pinned]
def t(f, n=6):
    out = []
    for _ in range(n):
        t0 = time.perf_counter(); r = f(); out.append((time.perf_counter() - t0) * 1e3)
    return [round(x, 2) for x in out]
print("set_parameter_fields (pinned) ms", t(lambda: m.set_parameter_fields(*host)))
print("set_parameter_fields (pageable) ms", t(lambda: m.set_parameter_fields(*med)))
print("get_gradient_vector, result dropped each time ms", t(lambda: m.get_gradient_vector()))
keep = []
print("get_gradient_vector, results kept ms", t(lambda: keep.append(m.get_gradient_vector())))
print("reset_kernels ms", t(lambda: m.reset_kernels()))
print("map_kernels ms", t(lambda: m.map_kernels_to_velocity()))
print("adjoint sources ms", t(lambda: m.calculate_l2_adjoint_sources()))
print("misfit shot ms", t(lambda: m.calculate_l2_misfit_shot(0)))
print("release ms", t(lambda: m.release_snapshots(0)))
