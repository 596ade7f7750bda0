#!/usr/bin/env python
"""Benchmark of the P-SV time loop (BASELINE.json metric: grid-point updates/s, fwd+adj, and
FWI gradient time per shot; % of the HBM roofline).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...

Workload (config.workload = "C2-marmousi-scale-gradient"): synthetic
Marmousi-scale elastic model, 3000 x 1220 inner grid + 50-point Cerjan frame = 3100 x 1320
points, nt = 4000, snapshot every 20 steps, 300 receivers, shots of 2 stacked sources.
One STEP = the FWI gradient contribution of ONE shot per GPU:
    forward_simulate(shot, store_fields=True)  -> per-shot L2 misfit -> adjoint sources
    -> adjoint_simulate(shot) with in-loop accumulation of the rho/lambda/mu kernels
    (N > 1: + one all-reduce of the three Lame kernels over NCCL, every step)
i.e. 2 * nx * nz * nt grid-point updates per GPU per step.  Shots are dealt round-robin to
the ranks (weak scaling: one shot per GPU per step).

`value` is timed with CUDA events on the solver's stream with all inputs resident in HBM.
`e2e` is the same step through the public Python API with HOST buffers: set_parameter_fields
from pinned host memory before, get_gradient_vector to the host after, inside the timed region.
`roofline` is the dominant sweep kernel: algorithmic bytes per launch over its
mean CUDA-event duration sampled inside the timed steps.
`cpu_baseline` / `--impl reference`: the UNMODIFIED reference core (oracle/_ref, OpenMP, all
host threads -- set explicitly, torchrun's OMP_NUM_THREADS=1 is overridden) on a bounded sample
of the same workload: same grid and medium, NT_SAMPLE time steps (fixed, stated in both arms'
`config`).  `parity` compares the CUDA path with that very reference run (seismograms, the
three Lame kernels, the gradient) on the sample.  Nothing here reads /root/reference at run time.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

WORKLOAD = "C2-marmousi-scale-gradient"
NT_SAMPLE = 400  # time steps of the reference arm / cpu_baseline sample (same grid, medium, sources, receivers)
GEOM = dict(nx_inner=3000, nz_inner=1220, np_boundary=50, nt=4000, snapshot_interval=20, n_shots=16, nr=300)


# ---------------------------------------------------------------------------------------------
# synthetic Marmousi-scale configuration + medium
# ---------------------------------------------------------------------------------------------
def workload_config(nt: int, n_shots: int, geom=GEOM) -> dict:
    nxi, nzi = geom["nx_inner"], geom["nz_inner"]
    n_sources = 2 * n_shots
    xs = np.linspace(150, nxi - 150, n_sources).round().astype(int)
    return dict(
        nt=nt, nx_inner=nxi, nz_inner=nzi, nx_inner_boundary=10, nz_inner_boundary=10, dx=2.5, dz=2.5, dt=2.5e-4,
        np_boundary=geom["np_boundary"], np_factor=0.0075, scalar_rho=2200.0, scalar_vp=3000.0, scalar_vs=1600.0,
        npx=1, npz=1, peak_frequency=8.0, source_timeshift=0.0, delay_cycles_per_shot=8.0,
        n_sources=n_sources, n_shots=n_shots, ix_sources=[int(v) for v in xs], iz_sources=[8] * n_sources,
        moment_angles=[float(a) for a in np.linspace(0, 170, n_sources)],
        which_source_to_fire_in_which_shot=[[2 * s, 2 * s + 1] for s in range(n_shots)],
        nr=geom["nr"], ix_receivers=[int(v) for v in np.linspace(20, nxi - 20, geom["nr"]).round()],
        iz_receivers=[10] * geom["nr"], snapshot_interval=geom["snapshot_interval"],
        observed_data_folder=".", stf_folder=".")


def write_ini(path: str, cfg: dict) -> None:
    def lst(x):
        if isinstance(x, (list, tuple)):
            return "{" + ", ".join(lst(v) for v in x) + "}"
        return repr(x)
    sections = {
        "domain": ["nt", "nx_inner", "nz_inner", "nx_inner_boundary", "nz_inner_boundary", "dx", "dz", "dt"],
        "boundary": ["np_boundary", "np_factor"],
        "medium": ["scalar_rho", "scalar_vp", "scalar_vs"],
        "sources": ["peak_frequency", "n_sources", "n_shots", "source_timeshift", "delay_cycles_per_shot",
                    "moment_angles", "ix_sources", "iz_sources", "which_source_to_fire_in_which_shot"],
        "receivers": ["nr", "ix_receivers", "iz_receivers"],
        "inversion": ["snapshot_interval"],
        "basis": ["npx", "npz"],
        "output": ["observed_data_folder", "stf_folder"],
    }
    with open(path, "w") as fh:
        for sec, keys in sections.items():
            fh.write(f"[{sec}]\n")
            for k in keys:
                v = cfg[k]
                fh.write(f"{k} = {v if isinstance(v, str) else lst(v)}\n")


def marmousi_like(nx: int, nz: int, seed: int = 1234, blob_seed: int | None = None):
    """40 dipping layers, vp 1500..4700 m/s increasing with depth + 5 % smoothed noise,
    vs = vp/1.9 floored at 400 m/s, rho = 310 vp^0.25 (Gardner).  blob_seed adds a +5 %
    Gaussian-blob perturbation (the 'true' model the observed data come from)."""
    rng = np.random.default_rng(seed)
    x = np.arange(nx)[:, None] / nx
    z = np.arange(nz)[None, :] / nz
    dip = 0.15 * x + 0.03 * np.sin(6.0 * x + 1.0)
    layer = np.clip(np.floor((z + dip) / 1.2 * 40), 0, 39).astype(int)
    steps = np.sort(rng.uniform(0, 1, 40)) * 0.6 + np.linspace(0, 0.4, 40)
    vp = 1500.0 + 3200.0 * steps[layer]
    noise = rng.standard_normal((nx // 8 + 2, nz // 8 + 2))
    noise = np.kron(noise, np.ones((8, 8)))[:nx, :nz]
    for _ in range(2):  # cheap smoothing
        noise = (noise + np.roll(noise, 1, 0) + np.roll(noise, -1, 0) + np.roll(noise, 1, 1) + np.roll(noise, -1, 1)) / 5
    vp = np.clip(vp * (1 + 0.05 * noise / max(noise.std(), 1e-12) * 0.5), 1500.0, 4900.0)
    if blob_seed is not None:
        r2 = np.random.default_rng(blob_seed)
        for _ in range(6):
            cx, cz, w = r2.uniform(0.1, 0.9), r2.uniform(0.2, 0.9), r2.uniform(0.03, 0.08)
            vp = vp * (1 + 0.05 * np.exp(-((x - cx) ** 2 + (z - cz) ** 2) / (2 * w * w)))
    vs = np.maximum(vp / 1.9, 400.0)
    rho = 310.0 * vp ** 0.25
    return np.ascontiguousarray(vp), np.ascontiguousarray(vs), np.ascontiguousarray(rho)


# ---------------------------------------------------------------------------------------------
# clocks
# ---------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap,power.draw")

    def __init__(self, index: int):
        self.index, self.rows, self._stop, self._t = index, [], threading.Event(), None

    def _run(self):
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5)
                if out.returncode == 0 and out.stdout.strip():
                    self.rows.append([c.strip() for c in out.stdout.strip().split(",")])
            except Exception:
                pass
            self._stop.wait(0.2)

    def __enter__(self):
        self._t = threading.Thread(target=self._run, daemon=True)
        self._t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        self._t.join(timeout=6)

    def summary(self) -> dict:
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unsampled"]}
        sm = sorted(float(r[0]) for r in self.rows)
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        reasons = [n for i, n in enumerate(names) if any(r[2 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(self.rows[0][1]), "reasons": reasons,
                "samples": len(sm), "power_w_max": max(float(r[6]) for r in self.rows)}


# ---------------------------------------------------------------------------------------------
# the reference arm / cpu baseline: unmodified reference core on a bounded sample
# ---------------------------------------------------------------------------------------------
def host_threads() -> int:
    """All the host threads this process may run on; also makes libgomp use them (torchrun exports
    OMP_NUM_THREADS=1 to its workers, which would silently time a one-thread reference)."""
    import ctypes
    try:
        n = len(os.sched_getaffinity(0))
    except AttributeError:
        n = os.cpu_count() or 1
    n = max(1, n)
    os.environ["OMP_NUM_THREADS"] = str(n)
    try:
        gomp = ctypes.CDLL("libgomp.so.1", mode=ctypes.RTLD_GLOBAL)
        gomp.omp_set_dynamic(0)
        gomp.omp_set_num_threads(n)
        return int(gomp.omp_get_max_threads())
    except Exception:
        return n


def cpu_sample_run(nt: int = NT_SAMPLE, geom=GEOM, steps: int = 1, warmup: int = 0, keep: bool = False):
    """Time forward(store)+adjoint of shot 0 on the CPU checker for `nt` time steps of the same
    grid/medium.  Returns a dict: rate (updates/s, best step), seconds per step (mean), sample text,
    kind, nt, threads and -- with keep=True -- the arrays the parity check compares."""
    threads = host_threads()
    from oracle import oracle
    kind = "reference" if oracle.ref_available() else "port"
    cfg = workload_config(nt, 1, geom)
    d = tempfile.mkdtemp(prefix="psv_cpu_")
    ini = os.path.join(d, "sample.ini")
    write_ini(ini, cfg)
    m = oracle.RefModel(ini) if kind == "reference" else oracle.OracleModel(ini)
    m.set_parameter_fields(*marmousi_like(m.nx, m.nz))
    zero = np.zeros((1, m.nr, nt))  # observed data: explicitly zero (the reference leaves the arrays uninitialised)
    m.set_observed_data(zero, zero)

    def gradient():
        t0 = time.perf_counter()
        m.forward_simulate(0, True)
        m.calculate_l2_misfit()
        m.calculate_l2_adjoint_sources()
        m.reset_kernels()
        m.adjoint_simulate(0)
        m.map_kernels_to_velocity()
        return time.perf_counter() - t0

    for _ in range(warmup):
        gradient()
    times = [gradient() for _ in range(max(1, steps))]
    updates = 2.0 * m.nx * m.nz * nt
    best = min(times)
    sample = (f"{kind} core (OpenMP, -O3, {threads} threads), shot 0 of {WORKLOAD}: same {m.nx}x{m.nz} grid and "
              f"medium, nt={nt} of {geom['nt']} time steps, forward(store)+adjoint, best of {len(times)}")
    out = dict(rate=updates / best, sec=sum(times) / len(times), sample=sample, kind=kind, nt=nt, threads=threads,
               ini=ini)
    if keep:
        out["arrays"] = dict(ux=m.array("rtf_ux").copy(), uz=m.array("rtf_uz").copy(),
                             kernels=np.stack([m.array(k).copy() for k in
                                               ("lambda_kernel", "mu_kernel", "density_l_kernel")]),
                             gradient=m.get_gradient_vector(), misfit=float(m.misfit))
    return out


def gpu_parity(ref: dict, options=()) -> dict:
    """The CUDA path (with the solver options of this run) on the sample the reference just ran (same .ini, same
    medium, observed data zero): relative L2 differences of what the hot path produces."""
    import psvwave_b200
    from psvwave_b200 import _array_ids  # noqa: F401  (name -> psv_array id)
    m = psvwave_b200.fdModel(ref["ini"])
    for kv in options:
        k, v = kv.split("=")
        m._set_option(k, int(v))
    m.set_parameter_fields(*marmousi_like(m.nx, m.nz))
    zero = np.zeros((1, m.nr, m.nt))
    m.set_observed_data(zero, zero)
    m.forward_simulate(0, True)
    m.calculate_l2_misfit()
    m.calculate_l2_adjoint_sources()
    m.reset_kernels()
    m.adjoint_simulate(0)
    m.map_kernels_to_velocity()
    a = ref["arrays"]

    def rel(x, y):
        n = float(np.linalg.norm(y))
        return float(np.linalg.norm(np.asarray(x) - y) / (n if n > 0 else 1.0))

    ux, uz = m.get_synthetic_data()
    ids = psvwave_b200._array_ids()
    kern = np.stack([m._get_array(ids[k]).reshape(m.nx, m.nz) for k in
                     ("lambda_kernel", "mu_kernel", "density_l_kernel")])
    return {"against": ref["kind"], "nt": ref["nt"], "rel_l2_ux": rel(ux, a["ux"]), "rel_l2_uz": rel(uz, a["uz"]),
            "rel_l2_kernels": rel(kern, a["kernels"]), "rel_l2_gradient": rel(m.get_gradient_vector(), a["gradient"]),
            "rel_misfit": abs(m.misfit - a["misfit"]) / abs(a["misfit"]),
            "tolerance": 1e-3 if m._get_option("fp32") else 1e-10}


def sample_config() -> dict:
    return {"workload": WORKLOAD, "grid": [GEOM["nx_inner"] + 2 * GEOM["np_boundary"],
                                           GEOM["nz_inner"] + 2 * GEOM["np_boundary"]],
            "nt": GEOM["nt"], "snapshot_interval": GEOM["snapshot_interval"], "sources_per_shot": 2,
            "receivers": GEOM["nr"], "reference_sample_nt": NT_SAMPLE, "shots_per_gpu_per_step": 1,
            "step": "forward(store)+misfit+adjoint(+kernel accumulation) of one shot; the GPU arm runs all nt time "
                    "steps per step, the reference arm reference_sample_nt of them (a rate metric)",
            "l2": f"working set {(8 * (GEOM['nx_inner'] + 2 * GEOM['np_boundary']) * (GEOM['nz_inner'] + 2 * GEOM['np_boundary']) * 8) >> 20}"
                  " MiB per sweep > 126 MB L2 (no flush needed)"}


def run_reference_arm(args) -> None:
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    r = cpu_sample_run(NT_SAMPLE, steps=args.steps, warmup=args.warmup)
    cfg = sample_config()  # the same dict as the GPU arm's `config`, key for key
    line = {
        "impl": "reference", "metric": "grid-point updates/s (fwd+adj)", "value": r["rate"], "unit": "updates/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["sec"] * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": cfg,
        "arm": {"parallelism": f"one shot on {r['threads']} host threads (OpenMP)", "kind": r["kind"]},
        "cpu_baseline": {"value": r["rate"], "unit": "updates/s", "cores": r["threads"], "kind": r["kind"],
                         "sample": r["sample"]},
        "e2e": {"value": r["rate"], "unit": "updates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


# ---------------------------------------------------------------------------------------------
# extras (reported next to the headline, never part of `value`): the other two multi-GPU layouts, driven by rank 0
# alone from ONE process over all `n_gpus` devices while the other ranks wait on the host
# ---------------------------------------------------------------------------------------------
def extra_slab_split(n_gpus: int, nt: int = 200) -> dict:
    """BASELINE.json configs[4]: ONE 8100 x 4100 grid cut into one x-slab per GPU (psv_march_slabs:
    peer-to-peer halo pulls over NVLink), forward(store) + adjoint of one shot -- strong scaling over n_gpus."""
    import psvwave_b200
    cfg = workload_config(nt, 1, dict(nx_inner=8000, nz_inner=4000, np_boundary=50, nr=64, snapshot_interval=50))
    cfg.update(n_sources=1, n_shots=1, ix_sources=[4000], iz_sources=[20], moment_angles=[90.0],
               which_source_to_fire_in_which_shot=[[0]], snapshot_interval=50)
    ini = os.path.join(tempfile.mkdtemp(prefix="psv_bench_slab_"), "c5.ini")
    write_ini(ini, cfg)
    d = psvwave_b200.SlabDomain(ini, n_slabs=n_gpus, halo_steps=4, devices=list(range(n_gpus)), rank=0, world_size=1)
    for k, m in enumerate(d.slabs):  # homogeneous + one anomaly, slab by slab (no whole-grid host array)
        a, b = d._rows(k)
        x, z = np.arange(a, b)[:, None], np.arange(d.nz)[None, :]
        blob = 1 + 0.1 * np.exp(-(((x - d.nx * 0.4) / 300.0) ** 2 + ((z - d.nz * 0.5) / 300.0) ** 2))
        m.set_parameter_fields(2000.0 * blob, 800.0 * blob, 1500.0 * np.ones_like(blob))
    zero = np.zeros((1, d.nr, d.nt))
    d.set_observed_data(zero, zero)
    d.reset_kernels()
    d.forward_simulate(0, True)
    d.calculate_l2_adjoint_sources()
    d.adjoint_simulate(0)  # warm-up of both marches
    best = None
    for _ in range(2):
        d.synchronize()
        t0 = time.perf_counter()
        d._march(0, False, True)
        d._march(0, True, False)
        d.synchronize()
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    # parity of the split (outside the timed region): the same shot on the WHOLE grid on one GPU.  Every slab point is
    # computed by the same kernel on the same operands, so seismograms and sensitivity kernels must agree bit for bit.
    def rel(x, y):
        n = float(np.linalg.norm(y))
        return float(np.linalg.norm(np.asarray(x) - y) / (n if n > 0 else 1.0))

    def clean_pass(obj):
        obj.reset_kernels()
        obj.forward_simulate(0, True)
        obj.calculate_l2_adjoint_sources()
        obj.adjoint_simulate(0)
        return obj.get_synthetic_data(), obj.get_kernels()

    (sux, suz), sk = clean_pass(d)
    psvwave_b200.set_default_device(0)
    w = psvwave_b200.fdModel(ini)
    x, z = np.arange(d.nx)[:, None], np.arange(d.nz)[None, :]
    blob = 1 + 0.1 * np.exp(-(((x - d.nx * 0.4) / 300.0) ** 2 + ((z - d.nz * 0.5) / 300.0) ** 2))
    w.set_parameter_fields(2000.0 * blob, 800.0 * blob, 1500.0 * np.ones_like(blob))
    w.set_observed_data(zero, zero)
    (wux, wuz), wk = clean_pass(w)
    parity = {"against": "the whole grid on one GPU (same library, one launch per step)", "n_slabs": n_gpus,
              "rel_l2_ux": rel(sux, wux), "rel_l2_uz": rel(suz, wuz),
              "rel_l2_kernels": rel(np.stack([np.asarray(a) for a in sk]), np.stack([np.asarray(a) for a in wk])),
              "bit_identical": bool(np.array_equal(sux, wux) and np.array_equal(suz, wuz) and
                                    all(np.array_equal(a, b) for a, b in zip(sk, wk))),
              "tolerance": 0.0}
    del w, sk, wk
    return {"workload": "C5-large-grid-forward+adjoint, one x-slab per GPU, peer-to-peer halos in one process",
            "parity": parity,
            "grid": [d.nx, d.nz], "nt": d.nt, "n_gpus": n_gpus, "halo_rows": d.halo_rows, "scaling": "strong",
            "exchanges_per_march": d.exchanges if n_gpus > 1 else 0, "seconds": best,
            "value": 2.0 * d.nx * d.nz * d.nt / best, "unit": "updates/s",
            "timing": "host clock around both marches, every stream synchronised on both sides, best of 2"}


def extra_one_object(n_gpus: int, nt: int = 2000) -> dict:
    """psv_set_devices: ONE fdModel object over n_gpus devices, the reference's own user loop from a plain
    single-process script (no torchrun, no PyTorch): one shot per GPU of the headline workload's geometry."""
    import psvwave_b200
    cfg = workload_config(nt, n_gpus)
    ini = os.path.join(tempfile.mkdtemp(prefix="psv_bench_obj_"), "c2.ini")
    write_ini(ini, cfg)
    m = psvwave_b200.fdModel(ini)
    m.set_devices(list(range(n_gpus)))
    m.set_parameter_fields(*marmousi_like(m.nx, m.nz))
    zero = np.zeros((m.n_shots, m.nr, m.nt))
    m.set_observed_data(zero, zero)
    best = None
    for _ in range(3):
        m.synchronize()
        t0 = time.perf_counter()
        for s in range(m.n_shots):
            m.forward_simulate(s)
        m.calculate_l2_misfit()
        m.calculate_l2_adjoint_sources()
        m.reset_kernels()
        for s in range(m.n_shots):
            m.adjoint_simulate(s)
        m.map_kernels_to_velocity()
        g = m.get_gradient_vector()
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    # parity (outside the timed region): the same shots one after the other on ONE device.  The group adds the per-GPU
    # kernels with one all-reduce, the single device adds shot by shot: same terms, another association (<= 1e-13).
    from psvwave_b200.parallel import ShotParallel
    group_misfit = float(m.misfit)
    psvwave_b200.set_default_device(0)
    one = psvwave_b200.fdModel(ini)
    one.set_parameter_fields(*marmousi_like(one.nx, one.nz))
    one.set_observed_data(zero, zero)
    one_misfit = ShotParallel(one, rank=0, world_size=1).gradient(release_snapshots=True)
    g1 = one.get_gradient_vector()
    n1 = float(np.linalg.norm(g1))
    parity = {"against": "one device running the same shots one after the other", "n_devices": n_gpus,
              "rel_l2_gradient": float(np.linalg.norm(g - g1) / (n1 if n1 > 0 else 1.0)),
              "rel_misfit": abs(group_misfit - one_misfit) / abs(one_misfit) if one_misfit else 0.0,
              "tolerance": 1e-13}
    parity["ok"] = bool(parity["rel_l2_gradient"] <= 1e-13 and parity["rel_misfit"] <= 1e-13)
    del one
    return {"workload": "C2 geometry, one shot per GPU, ONE object (psv_set_devices), the reference's call order",
            "parity": parity,
            "grid": [m.nx, m.nz], "nt": m.nt, "n_gpus": n_gpus, "n_shots": m.n_shots, "scaling": "weak", "seconds": best,
            "value": 2.0 * m.n_shots * m.nx * m.nz * m.nt / best, "unit": "updates/s",
            "allreduce_ms": m._get_stat("allreduce_ms") if n_gpus > 1 else 0.0,
            "gradient_norm": float(np.linalg.norm(g)),
            "timing": "host clock around the whole gradient incl. the gradient vector on the host, best of 3"}


def extra_default_config() -> dict:
    """BASELINE.json configs[0]: the reference's own default configuration (250 x 150 grid, nt 8000, 4 stacked sources,
    19 receivers) -- one forward_simulate and one whole gradient on one GPU.  Grids this small run the whole time loop
    in ONE cooperative launch with the strips resident in shared memory (resident.cuh)."""
    import psvwave_b200
    m = psvwave_b200.fdModel(os.path.join(ROOT, "configs", "default_testing_configuration.ini"))
    zero = np.zeros((m.n_shots, m.nr, m.nt))
    m.set_observed_data(zero, zero)
    best = None
    for _ in range(3):
        t0 = time.perf_counter()
        m.reset_kernels()
        m.forward_simulate(0, True)
        fwd_ms = m._get_stat("elapsed_ms")
        ctas = m._get_stat("resident_ctas")
        m.calculate_l2_misfit()
        m.calculate_l2_adjoint_sources()
        m.adjoint_simulate(0)
        adj_ms = m._get_stat("elapsed_ms")
        m.map_kernels_to_velocity()
        m.get_gradient_vector()
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    pts = float(m.nx) * m.nz * m.nt
    return {"workload": "C1 default_testing_configuration.ini, single shot", "grid": [m.nx, m.nz], "nt": m.nt,
            "forward_us_per_step": fwd_ms * 1e3 / m.nt, "adjoint_us_per_step": adj_ms * 1e3 / m.nt,
            "value": 2.0 * pts / ((fwd_ms + adj_ms) * 1e-3), "unit": "updates/s",
            "gradient_wall_s_per_shot": best, "resident_ctas": ctas,
            "timing": "CUDA events around each time loop; host clock around the whole gradient, best of 3"}


def extra_monte_carlo(n_models: int = 64) -> dict:
    """BASELINE.json configs[3]: an ensemble of media on the reference's default configuration, all
    forward-simulated together by psv_forward_batch (the reference's route: n copies of the object run one after the
    other, src/fdModel.cpp:64-97).  Device time of the whole ensemble from CUDA events on the leader's stream."""
    import psvwave_b200
    ini = os.path.join(ROOT, "configs", "default_testing_configuration.ini")
    rng = np.random.default_rng(2024)

    def smooth(a, n=6):
        for _ in range(n):
            a = (a + np.roll(a, 1, 0) + np.roll(a, -1, 0) + np.roll(a, 1, 1) + np.roll(a, -1, 1)) / 5.0
        return a

    models = []
    for _ in range(n_models):
        x = psvwave_b200.fdModel(ini)
        f = [1 + 0.05 * smooth(rng.standard_normal((x.nx, x.nz))) / 0.2 for _ in range(3)]
        x.set_parameter_fields(2000.0 * f[0], 800.0 * f[1], 1500.0 * f[2])
        models.append(x)
    lead = models[0]
    psvwave_b200.forward_batch(models, 0, False)  # warm-up
    best = None
    for _ in range(2):
        psvwave_b200.forward_batch(models, 0, False)
        dev = lead._get_stat("elapsed_ms") * 1e-3
        best = dev if best is None else min(best, dev)
    # parity: two members of the ensemble against the same media simulated on their own (same kernels: bit for bit)
    same = True
    for k in (1, n_models - 1):
        ux, uz = (np.array(a) for a in models[k].get_synthetic_data())
        solo = psvwave_b200.fdModel(ini)
        solo._set_option("resident", 0)
        solo.set_parameter_fields(*models[k].get_parameter_fields())
        solo.forward_simulate(0, False)
        sx, sz = solo.get_synthetic_data()
        same = same and bool(np.array_equal(ux, sx) and np.array_equal(uz, sz))
        del solo
    pts = float(lead.nx) * lead.nz * lead.nt
    return {"workload": "C4-monte-carlo-forward: models of default_testing_configuration.ini in one psv_forward_batch",
            "models": n_models, "grid": [lead.nx, lead.nz], "nt": lead.nt, "seconds": best,
            "models_per_s": n_models / best, "value": n_models * pts / best, "unit": "updates/s",
            "parity": {"against": "members 1 and n-1 simulated on their own (per-step kernels)", "bit_identical": same},
            "timing": "CUDA events around the ensemble's time loop, best of 2"}


# ---------------------------------------------------------------------------------------------
# our arm
# ---------------------------------------------------------------------------------------------
def pinned_copy(a: np.ndarray):
    import torch
    t = torch.empty(a.shape, dtype=torch.float64, pin_memory=True)
    t.numpy()[...] = a
    return t  # keep the tensor alive; use .numpy() views


def run_ours(args) -> None:
    import torch
    import torch.distributed as dist

    import psvwave_b200
    from psvwave_b200.parallel import ShotParallel

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if psvwave_b200.device_count() < 1:
        raise RuntimeError("bench.py needs a CUDA device (no CPU fallback)")
    torch.cuda.set_device(local)
    host_group = None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        host_group = dist.new_group(backend="gloo")  # host-side waits (no kernel spinning on the waiting GPUs)
    psvwave_b200.set_default_device(local)

    geom = dict(GEOM)
    if args.nt:
        geom["nt"] = args.nt
    n_shots = max(geom["n_shots"], world)
    cfg = workload_config(geom["nt"], n_shots, geom)
    tmp = tempfile.mkdtemp(prefix=f"psv_bench_{rank}_")
    ini = os.path.join(tmp, "c2.ini")
    write_ini(ini, cfg)
    m = psvwave_b200.fdModel(ini)
    for kv in args.option:
        k, v = kv.split("=")
        m._set_option(k, int(v))
    nx, nz, nt = m.nx, m.nz, m.nt
    start = marmousi_like(nx, nz)
    true = marmousi_like(nx, nz, blob_seed=4321)
    sp = ShotParallel(m, rank=rank, world_size=world)
    shots = sp.my_shots()
    need = args.steps + args.warmup
    my = [shots[i % len(shots)] for i in range(need)]

    # observed data = synthetics of the perturbed model for the shots this rank will touch
    m.set_parameter_fields(*true)
    for s in sorted(set(my)):
        m.forward_simulate(s, False)
    m.set_observed_data(*m.get_synthetic_data())
    m.set_parameter_fields(*start)
    pinned = [pinned_copy(a) for a in start]
    host_model = [t.numpy() for t in pinned]

    def step(shot, e2e):
        if e2e:
            m.set_parameter_fields(*host_model)
        m.reset_kernels()
        m.forward_simulate(shot, True)
        launches = m._get_stat("launches")
        prof = [m._get_stat("stress_ms"), m._get_stat("velocity_ms")]
        mis = m.calculate_l2_misfit_shot(shot)
        m.calculate_l2_adjoint_sources()
        m.adjoint_simulate(shot)
        launches += m._get_stat("launches") + 3
        prof += [m._get_stat("stress_ms"), m._get_stat("velocity_ms")]
        m.release_snapshots(shot)
        sp.allreduce_kernels()
        m.map_kernels_to_velocity()
        g = m.get_gradient_vector() if e2e else None
        return mis, launches + 1, prof, g

    def barrier():
        m.synchronize()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    m._set_option("profile", 1)  # CUDA events around ~32 sampled plain launches per time loop (< 0.1 % of a step)
    for i in range(args.warmup):
        step(my[i], False)
    barrier()
    launches_total, profs = 0, []
    with ClockSampler(local) as clocks:
        t_wall = time.perf_counter()
        m._timer_start()
        for i in range(args.steps):
            _, l, p, _ = step(my[args.warmup + i], False)
            launches_total += l
            profs.append(p)
        dev_ms = m._timer_stop()
        barrier()
        wall_ms = (time.perf_counter() - t_wall) * 1e3
    m._set_option("profile", 0)
    # one more step without the sampling events: what the default configuration does (reported, not the headline)
    barrier()
    m._timer_start()
    step(my[args.warmup], False)
    unprofiled_ms = m._timer_stop()

    # end to end through the public API with host buffers (two untimed steps first: the binding hands out its large
    # arrays in recycled page-locked buffers, and the first ones have to be allocated)
    for i in range(2):
        _, _, _, g = step(my[args.warmup + i % max(args.steps, 1)], True)
    barrier()
    t0 = time.perf_counter()
    for i in range(args.steps):
        _, _, _, g = step(my[args.warmup + i], True)
    barrier()
    e2e_ms = (time.perf_counter() - t0) * 1e3
    grad_bytes = int(g.nbytes)

    def rank_max(v):
        if world == 1:
            return v
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    dev_ms, wall_ms, e2e_ms = rank_max(dev_ms), rank_max(wall_ms), rank_max(e2e_ms)

    # N > 1: parity of the sharded path itself (outside the timed region).  The only cross-shot coupling of the
    # reference is the additive kernel accumulation (src/fdModel.cpp:832-858) and the misfit sum (:1128-1145): every
    # rank runs ITS shot of a short sample (1500 steps), the kernels are all-reduced over NCCL exactly as
    # in the timed steps, and rank 0 compares with ONE GPU running all N shots one after the other.
    parity_multi = None
    if world > 1:
        barrier()
        # 1500 steps, not the CPU arm's 400: the wavefields of neighbouring shots have to overlap for the ORDER of the
        # sum to matter (after 400 steps the delayed sources have barely fired and every rank's kernels are disjoint)
        nt_parity = min(geom["nt"], 1500)
        pcfg = workload_config(nt_parity, world, geom)
        pini = os.path.join(tmp, "parity.ini")
        write_ini(pini, pcfg)

        def sample_model():
            x = psvwave_b200.fdModel(pini)
            for kv in args.option:
                k, v = kv.split("=")
                x._set_option(k, int(v))
            x.set_parameter_fields(*start)
            zero = np.zeros((x.n_shots, x.nr, x.nt))
            x.set_observed_data(zero, zero)
            return x

        ids = psvwave_b200._array_ids()
        names = ("lambda_kernel", "mu_kernel", "density_l_kernel")
        ms = sample_model()
        mis_n = ShotParallel(ms, rank=rank, world_size=world).gradient(release_snapshots=True)
        if rank == 0:
            kern_n = np.stack([ms._get_array(ids[k]) for k in names])
            grad_n = np.array(ms.get_gradient_vector())
        del ms
        if rank == 0:
            one = sample_model()
            mis_1 = ShotParallel(one, rank=0, world_size=1).gradient(release_snapshots=True)
            kern_1 = np.stack([one._get_array(ids[k]) for k in names])
            grad_1 = np.array(one.get_gradient_vector())
            del one

            def rel(x, y):
                n = float(np.linalg.norm(y))
                return float(np.linalg.norm(x - y) / (n if n > 0 else 1.0))

            tol = 1e-4 if any(o.startswith("fp32=1") for o in args.option) else 1e-12
            parity_multi = {"against": f"one GPU running the same {world} shots one after the other (same library)",
                            "n_ranks": world, "nt": nt_parity, "rel_l2_kernels": rel(kern_n, kern_1),
                            "rel_l2_gradient": rel(grad_n, grad_1),
                            "rel_misfit": abs(mis_n - mis_1) / abs(mis_1) if mis_1 else 0.0, "tolerance": tol}
            parity_multi["ok"] = bool(max(parity_multi["rel_l2_kernels"], parity_multi["rel_l2_gradient"],
                                          parity_multi["rel_misfit"]) <= tol)
        barrier()

    # the other two multi-GPU layouts, from rank 0's process alone over all GPUs; the other ranks wait on the host
    extra = {}
    if not args.no_extra:
        barrier()
        if rank == 0:
            for name, fn in (("slab_split", extra_slab_split), ("one_object_all_gpus", extra_one_object),
                             ("default_config", lambda n: extra_default_config()),
                             ("monte_carlo", lambda n: extra_monte_carlo())):
                try:
                    extra[name] = fn(world)
                except Exception as e:  # an extra never takes the headline down
                    extra[name] = {"error": f"{type(e).__name__}: {e}"}
            psvwave_b200.set_default_device(local)
        if world > 1:
            dist.barrier(group=host_group)
    updates_per_step = 2.0 * nx * nz * nt * world
    value = updates_per_step * args.steps / (dev_ms * 1e-3)
    e2e_value = updates_per_step * args.steps / (e2e_ms * 1e-3)

    if rank == 0:
        peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
        if os.path.exists(peaks_path):
            peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        else:
            peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
        P = np.array(profs)  # columns: fwd stress, fwd velocity, adj stress, adj velocity (ms / launch)
        pts = float(nx) * nz
        bytes_per_launch = {"stress": 12 * 8 * pts, "velocity": 10 * 8 * pts, "fused": 16 * 8 * pts}
        moved_per_launch = {"stress": 10 * 8 * pts, "velocity": 8 * 8 * pts, "fused": 13 * 8 * pts}
        fused = m._get_option("variant") >= 2
        fp32 = bool(m._get_option("fp32"))  # optional mode, never the default: words are 4 bytes there
        if fp32:
            bytes_per_launch = {k: v / 2 for k, v in bytes_per_launch.items()}
            moved_per_launch = {k: v / 2 for k, v in moved_per_launch.items()}
        if fused:
            ms = float(P[:, [0, 2]].mean())
            kern, nbytes, moved = "fused stress+velocity sweep", bytes_per_launch["fused"], moved_per_launch["fused"]
            if m._get_option("variant") == 4 and not fp32:
                # the sampled launches of variant 4 are its two-step launches: 2 grid-point updates
                # per point, each credited with the single-pass 128 B
                kern, nbytes = "fused two-step sweep", 2 * bytes_per_launch["fused"]
        else:
            s_ms, v_ms = float(P[:, [0, 2]].mean()), float(P[:, [1, 3]].mean())
            if s_ms >= v_ms:
                kern, ms, nbytes, moved = "stress sweep", s_ms, bytes_per_launch["stress"], moved_per_launch["stress"]
            else:
                kern, ms, nbytes, moved = ("velocity sweep", v_ms, bytes_per_launch["velocity"],
                                           moved_per_launch["velocity"])
        achieved = nbytes / (ms * 1e-3) / 1e9 if ms > 0 else 0.0
        traffic, traffic_source = None, None
        tpath = os.path.join(ROOT, "profiles", "traffic.json")
        traffic_stale = None
        if os.path.exists(tpath):  # an ncu capture of the step kernel of ONE library build, not of this run
            import hashlib
            tj = json.load(open(tpath))
            traffic = tj.get(kern + (" (fp32=1)" if fp32 else ""))
            traffic_source = "profiles/traffic.json (" + str(tj.get("_source", "ncu --set full")) + ")"
            sha = hashlib.sha256(open(psvwave_b200.LIBRARY_PATH, "rb").read()).hexdigest()[:16]
            traffic_stale = bool(tj.get("_library_sha16") != sha)  # true: the capture is of another build of the library
        real = traffic if traffic else moved
        roofline = {"bound": "hbm", "kernel": kern, "achieved": achieved, "peak": peak, "unit": "GB/s",
                    "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_source, "traffic_stale": traffic_stale,
                    "peak_source": peak_src, "algorithmic_bytes_per_launch": nbytes, "ms_per_launch": ms,
                    "note": "achieved/frac (= algorithmic_*) credit the single-pass ALGORITHMIC 128 B per update (SURVEY "
                            "8d) and can exceed 1 because the kernel moves fewer bytes (la, b_vz, taper are not "
                            "streamed); moved_gbs / moved_frac are the physical bandwidth on the bus -- read THOSE as "
                            "the distance to the bandwidth bound",
                    "algorithmic_gbs": achieved, "algorithmic_frac": achieved / peak,
                    "moved_bytes_per_launch": real, "moved_gbs": real / (ms * 1e-3) / 1e9 if ms > 0 else 0.0,
                    "moved_frac": (real / (ms * 1e-3) / 1e9) / peak if ms > 0 else 0.0,
                    "all_ms_per_launch": {"fwd_stress": float(P[:, 0].mean()), "fwd_velocity": float(P[:, 1].mean()),
                                          "adj_stress": float(P[:, 2].mean()), "adj_velocity": float(P[:, 3].mean())}}
        cpu = parity = None
        if world == 1 and not args.no_cpu_baseline:
            r = cpu_sample_run(NT_SAMPLE, keep=True)
            cpu = {"value": r["rate"], "unit": "updates/s", "cores": r["threads"], "kind": r["kind"],
                   "sample": r["sample"]}
            parity = gpu_parity(r, args.option)
        cfg_out = sample_config()  # identical in both arms (the driver compares them); what is specific to this arm
        if args.nt:                 # goes to `arm`
            cfg_out.update({"grid": [nx, nz], "nt": nt})
        arm = {"parallelism": f"shots x{world}" + (" + one all-reduce of the kernels per step" if world > 1 else ""),
               "variant": m._get_option("variant"), "exact": m._get_option("exact"), "fp32": int(fp32)}
        line = {
            "metric": "grid-point updates/s (fwd+adj)", "value": value, "unit": "updates/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": dev_ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32 wavefield, f64 seismograms/kernels (optional mode)" if fp32 else "f64", "data": "synthetic",
            "config": cfg_out, "arm": arm,
            "gradient_ms_per_shot": dev_ms / args.steps, "wall_ms_per_step": wall_ms / args.steps,
            "unprofiled_ms_per_step": unprofiled_ms,
            "clocks": clocks.summary(),
            "e2e": {"value": e2e_value, "unit": "updates/s", "ms_per_step": e2e_ms / args.steps,
                    "h2d_bytes_per_step": int(sum(a.nbytes for a in host_model)), "d2h_bytes_per_step": grad_bytes},
            "gpu_launches": int(launches_total),
            "roofline": roofline,
        }
        if cpu:
            line["cpu_baseline"] = cpu
            line["parity"] = parity
        if parity_multi:
            line["parity"] = parity_multi
        if extra:
            line["extra"] = extra
        emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


_JSON_OUT = None


def emit(line: dict) -> None:
    out = _JSON_OUT or sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--nt", type=int, default=0, help="override the number of time steps (debug only)")
    ap.add_argument("--option", action="append", default=[], help="solver option name=value (debug only)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extra", action="store_true", help="skip the slab-split / one-object extras")
    args = ap.parse_args()
    # The solver and the reference print progress lines on stdout (src/fdModel.cpp:396,472);
    # keep stdout for the ONE JSON line: everything else goes to stderr.
    global _JSON_OUT
    sys.stdout.flush()
    _JSON_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
