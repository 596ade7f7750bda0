"""Domain split on N GPUs (BASELINE.json configs[4]), one process per GPU:
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        scripts/slab_bench.py [--nt 1000] [--halo-steps 4] [--nx-inner 8000] [--nz-inner 4000] [--check]
--check  first runs a small split problem in "exact" arithmetic and compares it BIT FOR BIT with the
         whole-grid run of the same problem on rank 0's GPU (halos over NCCL point-to-point).
Then times forward(store) + adjoint (with kernel accumulation) of one shot on the large grid: wall
clock between barriers around the marches (max over ranks), one JSON line from rank 0.
With N = 1 (plain `python scripts/slab_bench.py`) the same problem runs as one whole-grid slab.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import psvwave_b200  # noqa: E402
from helpers import make_config  # noqa: E402


def barrier(world):
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
        torch.cuda.synchronize()


def check(rank, world, tmp):
    ini = os.path.join(tmp, "check.ini")
    make_config(ini, nt=300, nx_inner=46 * world + 40, nz_inner=61, nx_inner_boundary=4, nz_inner_boundary=3,
                np_boundary=9, np_factor=0.03, n_sources=2, n_shots=1, ix_sources=[20, 23 * world + 20],
                iz_sources=[30, 12], moment_angles=[30.0, 100.0], which_source_to_fire_in_which_shot=[[0, 1]],
                nr=4, ix_receivers=[5, 23 * world + 19, 23 * world + 20, 46 * world + 30], iz_receivers=[8, 9, 40, 41],
                snapshot_interval=7, npx=1, npz=1, dt=0.00025)
    d = psvwave_b200.SlabDomain(ini, halo_steps=2, options=dict(exact=1))
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
    ok = True
    if rank == 0:
        w = psvwave_b200.fdModel(ini)
        w._set_option("exact", 1)
        wux, wk, wg = workflow(w)
        ok = bool(np.array_equal(ux, wux) and all(np.array_equal(a, b) for a, b in zip(kernels, wk))
                  and np.array_equal(grad, wg) and np.abs(wg).max() > 0)
        print(json.dumps({"check": "split grid over NCCL vs whole grid, exact arithmetic", "slabs": world,
                          "bit_identical": ok, "exchanges": d.exchanges}), flush=True)
    return ok


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--nt", type=int, default=1000)
    ap.add_argument("--halo-steps", type=int, default=4)
    ap.add_argument("--nx-inner", type=int, default=8000)
    ap.add_argument("--nz-inner", type=int, default=4000)
    ap.add_argument("--check", action="store_true")
    ap.add_argument("--sync-exchange", action="store_true", help="host-synchronised halo exchange (debug)")
    args = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    psvwave_b200.set_default_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    tmp = f"/tmp/psv_slab_{rank}"
    os.makedirs(tmp, exist_ok=True)
    if args.check and not check(rank, world, tmp):
        sys.exit(1)

    ini = os.path.join(tmp, "c5.ini")
    nxi, nzi = args.nx_inner, args.nz_inner
    make_config(ini, nt=args.nt, nx_inner=nxi, nz_inner=nzi, nx_inner_boundary=10, nz_inner_boundary=10, np_boundary=50,
                np_factor=0.0075, dx=2.5, dz=2.5, dt=2.5e-4, peak_frequency=8.0, n_sources=1, n_shots=1,
                ix_sources=[nxi // 2], iz_sources=[20], moment_angles=[90.0], which_source_to_fire_in_which_shot=[[0]],
                nr=64, ix_receivers=[int(x) for x in np.linspace(100, nxi - 100, 64)], iz_receivers=[10] * 64,
                snapshot_interval=50, npx=1, npz=1)
    d = psvwave_b200.SlabDomain(ini, halo_steps=args.halo_steps, async_exchange=not args.sync_exchange)
    for k, m in enumerate(d.slabs):  # homogeneous + one seeded anomaly, set slab by slab (no whole-grid host array)
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
    barrier(world)
    t0 = time.perf_counter()
    step()
    barrier(world)
    dt = time.perf_counter() - t0
    t = torch.tensor([dt], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dt = float(t.item())
    dev_ms = d.slabs[0]._get_stat("elapsed_ms")
    if rank == 0:
        updates = 2.0 * d.nx * d.nz * d.nt
        print(json.dumps({"workload": "C5-large-grid-forward+adjoint", "grid": [d.nx, d.nz], "nt": d.nt, "n_gpus": world,
                          "halo_rows": d.halo_rows, "stream_ordered_exchange": bool(world > 1 and d._stream_ordered()), "exchanges_per_march": d.exchanges, "scaling": "strong",
                          "seconds": dt, "updates_per_s": updates / dt, "adjoint_device_ms_rank0": dev_ms,
                          "misfit": d.misfit,
                          "halo_bytes_per_exchange_per_side": 5 * d.halo_rows * int(d.info[0]["pitch"]) * 8}), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
