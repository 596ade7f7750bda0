import sys, json, os, tempfile
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import numpy as np, psvwave_b200
from helpers import make_config
for (nxi, nzi) in ((944, 4000), (1950, 4000), (8000, 4000), (3000, 1220)):
    ini = os.path.join(tempfile.mkdtemp(), "s.ini")
    make_config(ini, nt=200, nx_inner=nxi, nz_inner=nzi, nx_inner_boundary=10, nz_inner_boundary=10, np_boundary=50, np_factor=0.0075, dx=2.5, dz=2.5, dt=2.5e-4, peak_frequency=8.0, n_sources=1, n_shots=1, ix_sources=[nxi // 2], iz_sources=[20], moment_angles=[90.0], which_source_to_fire_in_which_shot=[[0]], nr=8, ix_receivers=[int(x) for x in np.linspace(100, nxi - 100, 8)], iz_receivers=[10] * 8, snapshot_interval=50, npx=1, npz=1)
    for tune in (0, -1):
        m = psvwave_b200.fdModel(ini); m._set_option("tune", tune)
        m.forward_simulate(0, False); m.forward_simulate(0, False)
        ms = m._get_stat("elapsed_ms")
        print(json.dumps(dict(grid=[m.nx, m.nz], tune=tune, us_per_step=ms * 1e3 / m.nt, G=m.nx * m.nz * m.nt / ms / 1e6, warps=m._get_stat("tuned_warps"), strip=m._get_stat("tuned_strip"), tune_ms=m._get_stat("tune_ms"))), flush=True)
        del m
