"""Shared builders for the parity tests (inputs only -- no checking logic lives here)."""
import os

import numpy as np

from oracle.oracle import config_from_ini, write_ini

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
C1_INI = os.path.join(ROOT, "configs", "default_testing_configuration.ini")

DYNAMIC = ("vx", "vz", "txx", "tzz", "txz")


def notebook2_models(m):
    """Start and three-circle target of the reference's notebook 2, cell 3."""
    IX, IZ = m.get_coordinates(True)
    shape = IX.shape
    vp0, vs0, rho0 = np.full(shape, 2000.0), np.full(shape, 800.0), np.full(shape, 1500.0)

    def circle(fx):
        xm = (IX.max() + IX.min()) * fx
        zm = (IZ.max() + IZ.min()) / 2
        return ((IX - xm) ** 2 + (IZ - zm) ** 2) ** 0.5 < 15

    c = circle(1 / 2)
    vs = vs0 * (1 - 0.1 * c)
    vp = vp0 * (1 - 0.1 * c)
    rho = rho0.copy()
    c = circle(1 / 6)
    vs = vs * (1 + 0.1 * c)
    rho = rho * (1 + 0.1 * c)
    c = circle(5 / 6)
    vp = vp * (1 + 0.1 * c)
    rho = rho * (1 + 0.1 * c)
    return (vp0, vs0, rho0), (vp, vs, rho)


def gradient_workflow(m):
    """The reference's documented call order (docs-source/source/typical-workflow.rst:5-13)."""
    for s in range(m.n_shots):
        m.forward_simulate(s)
    m.calculate_l2_misfit()
    m.calculate_l2_adjoint_sources()
    m.reset_kernels()
    for s in range(m.n_shots):
        m.adjoint_simulate(s)
    m.map_kernels_to_velocity()


def random_medium(nx, nz, seed, amp=0.05):
    rng = np.random.default_rng(seed)
    vp = 2000.0 * (1 + amp * rng.standard_normal((nx, nz)))
    vs = 800.0 * (1 + amp * rng.standard_normal((nx, nz)))
    rho = 1500.0 * (1 + amp * rng.standard_normal((nx, nz)))
    return vp, vs, rho


def make_config(path, **overrides):
    """Write a variation of the default test configuration and return its dict."""
    cfg = config_from_ini(C1_INI)
    cfg.update(overrides)
    write_ini(path, cfg)
    return cfg


def ragged_config(path, **extra):
    """Odd grid sizes (nz not a multiple of 2/16), 3 shots of different source counts,
    receivers sharing a point and sitting next to a source, basis blocks."""
    base = dict(nt=240, nx_inner=47, nz_inner=33, nx_inner_boundary=4, nz_inner_boundary=3,
                np_boundary=7, np_factor=0.04, n_sources=4, n_shots=3,
                ix_sources=[8, 9, 30, 40], iz_sources=[6, 7, 20, 5],
                moment_angles=[0.0, 45.0, 90.0, 135.0],
                which_source_to_fire_in_which_shot=[[1, 0], [2], [3]],
                nr=5, ix_receivers=[3, 9, 9, 44, 20], iz_receivers=[3, 7, 7, 30, 16],
                snapshot_interval=6, npx=3, npz=9, delay_cycles_per_shot=2, dt=0.00025)
    base.update(extra)
    return make_config(path, **base)


def array_ids():
    """psv_array enumerators of include/psvwave_b200.h -> {reference field name: id}."""
    import re
    text = open(os.path.join(ROOT, "include", "psvwave_b200.h")).read()
    body = re.search(r"typedef enum psv_array \{(.*?)\} psv_array;", text, re.S).group(1)
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    names = [n.strip().split("=")[0].strip() for n in body.split(",") if n.strip()]
    out = {}
    for i, n in enumerate(names):
        key = n[len("PSV_"):].lower()
        key = {"rtf_ux_true": "rtf_ux_true", "rtf_uz_true": "rtf_uz_true"}.get(key, key)
        out[key] = i
    return out


def gpu_array(m, name):
    """Any internal array of the CUDA model, shaped like the reference's (src/fdModel.h:209-252)."""
    ids = array_ids()
    flat = m._get_array(ids[name])
    if name in ("t",):
        return flat
    if name == "stf":
        return flat.reshape(m.n_sources, m.nt)
    if name == "moment":
        return flat.reshape(m.n_sources, 2, 2)
    if name.startswith(("rtf_", "a_stf_")):
        return flat.reshape(m.n_shots, m.nr, m.nt)
    if name.startswith("accu_"):
        return flat.reshape(m.n_shots, m.snapshots, m.nx, m.nz)
    return flat.reshape(m.nx, m.nz)
