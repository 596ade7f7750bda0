"""Parity AT THE SCALE THAT IS BENCHMARKED: the CUDA path (default kernel variant, through the C ABI) against the
compiled UNMODIFIED reference core (oracle/_ref, /root/reference/src/fdModel.cpp:526-969 forward + adjoint loops) on

  * the bench workload's grid (3100 x 1320, `bench.workload_config`: `marmousi_like` medium, 300 receivers,
    snapshot every 20 steps) -- with the launch geometry bench.py times (chosen by the tuner at the first time
    loop: 3, 4 or 8 consumer warps per CTA, multi-row strips on every SM; the geometry that ran is recorded) and
    with each compiled CTA width pinned;
  * the C5 grid (8100 x 4100, BASELINE.json configs[4]), whole and cut into 2 and 4 in-process x-slabs with halo
    exchange (psvwave_b200.SlabDomain).

Time axes are short (the reference does ~0.3 G updates/s) and the sources are a lattice of simultaneous high-frequency
Ricker sources, so that after a few hundred steps the wavefield is non-zero on the whole grid and every CTA tile,
strip boundary and slab cut carries signal.  Compared: seismograms, three snapshots of all five fields, the six
sensitivity kernels, the gradient vector, the misfit.  Tolerances: default arithmetic <= 1e-10 relative L2 (north_star),
"exact" arithmetic bit-identical (np.array_equal).
"""
import gc
import os

import numpy as np
import pytest

import bench
from conftest import rel_l2
from helpers import DYNAMIC, array_ids, gpu_array
from oracle import oracle

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not oracle.ref_available(), reason="oracle/_ref not built")]

TOL = 1e-10
LAME = ("lambda_kernel", "mu_kernel", "density_l_kernel")
VELK = ("vp_kernel", "vs_kernel", "density_v_kernel")


def lattice_config(geom, nt, f_peak, step_x, step_z, extra_rows=()):
    """bench.workload_config geometry with ONE shot of lattice sources firing (almost) together; `extra_rows`
    (inner-grid rows) get a line of sources of their own (rows next to slab cuts)."""
    cfg = bench.workload_config(nt, 1, geom)
    xs = np.arange(step_x // 2, geom["nx_inner"] - 8, step_x)
    zs = np.arange(step_z // 2, geom["nz_inner"] - 8, step_z)
    ix, iz = (a.ravel() for a in np.meshgrid(xs, zs, indexing="ij"))
    for r in extra_rows:
        ez = np.arange(step_z // 3, geom["nz_inner"] - 8, 3 * step_z)
        ix, iz = np.concatenate([ix, np.full(ez.size, r)]), np.concatenate([iz, ez])
    n = ix.size
    # receivers spread over the whole grid, each 2-3 cells from a source (the workload's own line of receivers near the
    # surface would see nothing of most sources within a few hundred steps); the cut-adjacent sources come last
    pick = np.unique(np.concatenate([np.linspace(0, n - 1, cfg["nr"] - 24).round().astype(int),
                                     np.arange(n - 24, n)]))[: cfg["nr"]]
    cfg.update(nr=int(pick.size), ix_receivers=[int(ix[q]) + 2 for q in pick], iz_receivers=[int(iz[q]) + 3 for q in pick])
    cfg.update(n_sources=n, ix_sources=[int(v) for v in ix], iz_sources=[int(v) for v in iz],
               moment_angles=[float(round(a, 1)) for a in np.linspace(0.0, 340.0, n)],
               which_source_to_fire_in_which_shot=[list(range(n))], peak_frequency=float(f_peak),
               delay_cycles_per_shot=0.0)
    return cfg


def run_workflow(m):
    """forward(store) -> misfit -> adjoint sources -> adjoint -> map with zero observed data (a_stf = synthetics)."""
    m.forward_simulate(0, True)
    zero = np.zeros((1, m.nr, m.nt))  # explicitly: the reference leaves rtf_u*_true uninitialised (src/fdModel.cpp:139-191)
    m.set_observed_data(zero, zero)
    m.calculate_l2_misfit()
    m.calculate_l2_adjoint_sources()
    m.reset_kernels()
    m.adjoint_simulate(0)
    m.map_kernels_to_velocity()


def reference_results(ini, medium, snaps):
    bench.host_threads()  # all host cores for the reference's OpenMP loops
    r = oracle.RefModel(ini)
    r.set_parameter_fields(*medium)
    run_workflow(r)
    out = dict(ux=r.array("rtf_ux").copy(), uz=r.array("rtf_uz").copy(), misfit=float(r.misfit),
               gradient=r.get_gradient_vector())
    for k in LAME + VELK:
        out[k] = r.array(k).copy()
    for f in DYNAMIC:
        out["accu_" + f] = r.array("accu_" + f)[0, snaps].copy()
        out["end_" + f] = r.array(f).copy()
    assert np.abs(out["ux"]).max() > 0 and np.abs(out["gradient"]).max() > 0
    del r
    gc.collect()
    return out


def compare_whole(m, ref, snaps, exact):
    def close(a, b, what):
        if exact:
            assert np.array_equal(a, b), f"{what}: not bit-identical (rel L2 {rel_l2(a, b):.3e})"
        else:
            e = rel_l2(a, b)
            assert e <= TOL, f"{what}: rel L2 {e:.3e} > {TOL}"

    ux, uz = m.get_synthetic_data()
    close(ux, ref["ux"], "ux")
    close(uz, ref["uz"], "uz")
    assert abs(m.misfit - ref["misfit"]) <= (1e-13 if exact else TOL) * abs(ref["misfit"])
    ids = array_ids()
    for f in DYNAMIC:
        a = m._get_array(ids["accu_" + f]).reshape(m.n_shots, m.snapshots, m.nx, m.nz)[0, snaps]
        close(a, ref["accu_" + f], "snapshot " + f)
        del a
        close(gpu_array(m, f), ref["end_" + f], "final adjoint " + f)
    for k in LAME + VELK:
        close(gpu_array(m, k), ref[k], k)
    close(m.get_gradient_vector(), ref["gradient"], "gradient")


# ---- the bench workload's grid ---------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def bench_grid(tmp_path_factory):
    nt, snaps = 300, [1, 7, 14]
    cfg = lattice_config(bench.GEOM, nt, f_peak=40.0, step_x=120, step_z=110)
    ini = str(tmp_path_factory.mktemp("scale") / "bench_grid.ini")
    bench.write_ini(ini, cfg)
    medium = bench.marmousi_like(bench.GEOM["nx_inner"] + 2 * bench.GEOM["np_boundary"],
                                 bench.GEOM["nz_inner"] + 2 * bench.GEOM["np_boundary"])
    return dict(ini=ini, medium=medium, snaps=snaps, ref=reference_results(ini, medium, snaps))


@pytest.mark.parametrize("exact", [0, 1], ids=["default-arithmetic", "exact"])
def test_bench_grid_vs_compiled_reference(bench_grid, exact, record_property):
    """The default configuration: the launch geometry is whatever the tuner measures on this box at the first time
    loop (option "tune" = -1 on a grid this size); the geometry that actually ran is recorded with the result."""
    import psvwave_b200
    m = psvwave_b200.fdModel(bench_grid["ini"])
    assert (m.nx, m.nz) == (3100, 1320) and m._get_option("variant") == 3 and m._get_option("tune") == -1
    m._set_option("exact", exact)
    m.set_parameter_fields(*bench_grid["medium"])
    run_workflow(m)
    warps, strip = int(m._get_stat("tuned_warps")), int(m._get_stat("tuned_strip"))
    record_property("tuned_warps", warps)
    record_property("tuned_strip", strip)
    print(f"geometry that ran: {warps} consumer warps per CTA, strips of {strip} rows")
    assert warps in (3, 4, 8) and strip >= 1  # one of the compiled widths, chosen by measurement
    compare_whole(m, bench_grid["ref"], bench_grid["snaps"], bool(exact))


@pytest.mark.parametrize("warps", [3, 4, 8])
def test_bench_grid_every_compiled_width_vs_compiled_reference(bench_grid, warps):
    """Each compiled CTA width pinned (tuner off), so that the geometry the tuner did NOT pick on this box is compared
    with the reference on the bench grid as well."""
    import psvwave_b200
    m = psvwave_b200.fdModel(bench_grid["ini"])
    m._set_option("tune", 0)
    m._set_option("pipe_warps", warps)
    m.set_parameter_fields(*bench_grid["medium"])
    run_workflow(m)
    assert m._get_option("pipe_warps") == warps
    compare_whole(m, bench_grid["ref"], bench_grid["snaps"], False)


# ---- the C5 grid, whole and split ----------------------------------------------------------------------------------
C5_GEOM = dict(nx_inner=8000, nz_inner=4000, np_boundary=50, nt=40, snapshot_interval=10, n_shots=1, nr=300)


@pytest.fixture(scope="module")
def c5_grid(tmp_path_factory):
    nt, snaps = 40, [1, 2, 3]
    # extra source lines 1-2 rows either side of the slab cuts (global rows 2025, 4050, 6075 = inner rows - 50)
    cuts = [g - 50 + o for g in (2025, 4050, 6075) for o in (-2, 1)]
    cfg = lattice_config(C5_GEOM, nt, f_peak=400.0, step_x=90, step_z=80, extra_rows=cuts)
    ini = str(tmp_path_factory.mktemp("scale") / "c5_grid.ini")
    bench.write_ini(ini, cfg)
    nx, nz = 8100, 4100
    rng = np.random.default_rng(5)
    x = np.arange(nx)[:, None] / nx
    z = np.arange(nz)[None, :] / nz
    vp = 2000.0 + 1500.0 * z + 200.0 * np.sin(9.0 * x) + 300.0 * np.exp(-((x - 0.52) ** 2 + (z - 0.4) ** 2) / 0.005)
    vp = vp * (1.0 + 0.01 * rng.standard_normal((nx, nz)))
    medium = (np.ascontiguousarray(vp), np.ascontiguousarray(vp / 1.8), np.ascontiguousarray(310.0 * vp ** 0.25))
    return dict(ini=ini, medium=medium, snaps=snaps, ref=reference_results(ini, medium, snaps))


@pytest.mark.parametrize("exact", [0, 1], ids=["default-arithmetic", "exact"])
def test_c5_grid_vs_compiled_reference(c5_grid, exact):
    import psvwave_b200
    m = psvwave_b200.fdModel(c5_grid["ini"])
    assert (m.nx, m.nz) == (8100, 4100)
    m._set_option("exact", exact)
    m.set_parameter_fields(*c5_grid["medium"])
    run_workflow(m)
    compare_whole(m, c5_grid["ref"], c5_grid["snaps"], bool(exact))


@pytest.mark.parametrize("n_slabs", [2, 4])
def test_c5_grid_split_in_slabs_vs_compiled_reference(c5_grid, n_slabs):
    """x-slabs of ~4000 / ~2000 rows with 16-row halos (4 steps between exchanges), exact arithmetic: bit-identical
    to the reference although every slab recomputes its halo."""
    import psvwave_b200
    ref = c5_grid["ref"]
    d = psvwave_b200.SlabDomain(c5_grid["ini"], n_slabs=n_slabs, halo_steps=4, options=dict(exact=1))
    d.set_parameter_fields(*c5_grid["medium"])
    run_workflow(d)
    ux, uz = d.get_synthetic_data()
    assert np.array_equal(ux, ref["ux"]) and np.array_equal(uz, ref["uz"])
    assert abs(d.misfit - ref["misfit"]) <= 1e-13 * abs(ref["misfit"])
    for a, k in zip(d.get_kernels(), VELK):
        assert np.array_equal(a, ref[k]), k
    assert np.array_equal(d.get_gradient_vector(), ref["gradient"])
    ids = array_ids()
    for f in ("vx", "txz"):
        got = d.gather_grid(lambda m, f=f: m._get_array(ids[f]))
        assert np.array_equal(got, ref["end_" + f]), f
    assert d.exchanges == (d.nt - 1) // 4
