"""Domain split inside ONE process (BASELINE.json configs[4]): the large grid cut into one x-slab per GPU,
the march native (psv_march_slabs: peer-to-peer halo pulls over NVLink, CUDA events between the slabs' streams; no NCCL,
no PyTorch).  Strong scaling of forward(store) + adjoint (with kernel accumulation) of one shot:

    python scripts/slab_inprocess_bench.py [--nt 1000] [--halo-steps 4] [--counts 1,2,4,8] [--nx-inner 8000] [--nz-inner 4000]

One JSON line per GPU count (wall clock around the marches with every stream synchronised on both sides; best of 2
after a warm-up); `speedup` is against the first line of the run.  --check compares a small split problem bit for bit
with its whole-grid run first."""
import argparse
import json
import os
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import psvwave_b200  # noqa: E402
from helpers import make_config  # noqa: E402


def check(n, tmp):
    ini = os.path.join(tmp, "check.ini")
    make_config(ini, nt=300, nx_inner=46 * n + 40, nz_inner=61, nx_inner_boundary=4, nz_inner_boundary=3,
                np_boundary=9, np_factor=0.03, n_sources=2, n_shots=1, ix_sources=[20, 23 * n + 20],
                iz_sources=[30, 12], moment_angles=[30.0, 100.0], which_source_to_fire_in_which_shot=[[0, 1]],
                nr=4, ix_receivers=[5, 23 * n + 19, 23 * n + 20, 46 * n + 30], iz_receivers=[8, 9, 40, 41],
                snapshot_interval=7, npx=1, npz=1, dt=0.00025)
    d = psvwave_b200.SlabDomain(ini, n_slabs=n, halo_steps=2, options=dict(exact=1), devices=list(range(n)))
    rng = np.random.default_rng(5)
    med = [base * (1 + 0.05 * rng.standard_normal((d.nx, d.nz))) for base in (2000.0, 800.0, 1500.0)]
    obs = [rng.standard_normal((1, d.nr, d.nt)) * 1e-3 for _ in range(2)]

    def workflow(m):
        m.set_parameter_fields(*med)
        m.forward_simulate(0, True)
        m.set_observed_data(*obs)
        m.calculate_l2_misfit()
        m.calculate_l2_adjoint_sources()
        m.reset_kernels()
        m.adjoint_simulate(0)
        m.map_kernels_to_velocity()
        return m.get_synthetic_data()[0], m.get_kernels(), m.get_gradient_vector()

    ux, kernels, grad = workflow(d)
    w = psvwave_b200.fdModel(ini)
    w._set_option("exact", 1)
    wux, wk, wg = workflow(w)
    ok = bool(np.array_equal(ux, wux) and all(np.array_equal(a, b) for a, b in zip(kernels, wk))
              and np.array_equal(grad, wg) and np.abs(wg).max() > 0)
    print(json.dumps({"check": "split grid, one slab per GPU, peer-to-peer halos vs whole grid, exact arithmetic",
                      "slabs": n, "devices": [int(m.device) for m in d.slabs], "bit_identical": ok,
                      "exchanges": d.exchanges}), flush=True)
    return ok


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--nt", type=int, default=1000)
    ap.add_argument("--halo-steps", type=int, default=4)
    ap.add_argument("--nx-inner", type=int, default=8000)
    ap.add_argument("--nz-inner", type=int, default=4000)
    ap.add_argument("--counts", default="")
    ap.add_argument("--check", action="store_true")
    args = ap.parse_args()
    n_dev = psvwave_b200.device_count()
    counts = [int(c) for c in args.counts.split(",")] if args.counts else [n for n in (1, 2, 4, 8) if n <= n_dev]
    tmp = tempfile.mkdtemp(prefix="psv_slab_ip_")
    if args.check:
        for n in counts:
            if n > 1 and not check(n, tmp):
                sys.exit(1)
    ini = os.path.join(tmp, "c5.ini")
    nxi, nzi = args.nx_inner, args.nz_inner
    make_config(ini, nt=args.nt, nx_inner=nxi, nz_inner=nzi, nx_inner_boundary=10, nz_inner_boundary=10, np_boundary=50,
                np_factor=0.0075, dx=2.5, dz=2.5, dt=2.5e-4, peak_frequency=8.0, n_sources=1, n_shots=1,
                ix_sources=[nxi // 2], iz_sources=[20], moment_angles=[90.0], which_source_to_fire_in_which_shot=[[0]],
                nr=64, ix_receivers=[int(x) for x in np.linspace(100, nxi - 100, 64)], iz_receivers=[10] * 64,
                snapshot_interval=50, npx=1, npz=1)
    base = None
    for n in counts:
        d = psvwave_b200.SlabDomain(ini, n_slabs=n, halo_steps=args.halo_steps, devices=list(range(n)))
        for k, m in enumerate(d.slabs):  # homogeneous + one seeded anomaly, slab by slab (no whole-grid host array)
            a, b = d._rows(k)
            x = np.arange(a, b)[:, None]
            z = np.arange(d.nz)[None, :]
            blob = 1 + 0.1 * np.exp(-(((x - d.nx * 0.4) / 300.0) ** 2 + ((z - d.nz * 0.5) / 300.0) ** 2))
            m.set_parameter_fields(2000.0 * blob, 800.0 * blob, 1500.0 * np.ones_like(blob))
        d.set_observed_data(np.zeros((1, d.nr, d.nt)), np.zeros((1, d.nr, d.nt)))

        def step():
            d.reset_kernels()
            d.forward_simulate(0, True)
            d.calculate_l2_misfit()
            d.calculate_l2_adjoint_sources()
            d.adjoint_simulate(0)

        step()  # warm-up
        walls, marches = [], []
        for _ in range(2):
            d.synchronize()
            t0 = time.perf_counter()
            step()
            d.synchronize()
            walls.append(time.perf_counter() - t0)
            # the marches alone: forward without the host-side seismogram merge
            d.synchronize()
            t0 = time.perf_counter()
            d._march(0, False, True)
            d._march(0, True, False)
            d.synchronize()
            marches.append(time.perf_counter() - t0)
        dt, dm = min(walls), min(marches)
        if base is None:
            base = dm
        updates = 2.0 * d.nx * d.nz * d.nt
        print(json.dumps({"workload": "C5-large-grid-forward+adjoint", "grid": [d.nx, d.nz], "nt": d.nt, "n_gpus": n,
                          "transport": "in-process, peer-to-peer pulls (psv_march_slabs)", "halo_rows": d.halo_rows,
                          "exchanges_per_march": d.exchanges if n > 1 else 0, "scaling": "strong",
                          "seconds_marches": dm, "updates_per_s": updates / dm, "speedup": base / dm,
                          "seconds_with_host_merge": dt, "misfit": d.misfit,
                          "halo_bytes_per_exchange_per_side": 5 * d.halo_rows * int(d.info[0]["pitch"]) * 8}), flush=True)
        del d


if __name__ == "__main__":
    main()
