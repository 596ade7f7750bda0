"""Generate tests/golden/*.npz from the UNMODIFIED reference core.

Run in the dev container (where /root/reference exists):

    make -C oracle && python tests/golden/make_golden.py

It drives oracle/_ref/libpsv_ref.so (reference sources compiled where they lie, see
oracle/Makefile) through oracle/ref_shim.cpp and stores the reference's outputs for
three cases.  The .npz files are committed; the GPU box (no /root/reference) checks the
CUDA path and the C restatement against them.

Cases
  c1_homogeneous   configs/default_testing_configuration.ini, homogeneous medium,
                   forward_simulate(0, store_fields=True)                     
  c1_notebook2     same ini, three-circle target of notebook 2 cell 3 as observed data,
                   homogeneous start: misfit + full gradient workflow.  The reference's
                   own notebook prints misfit 1920401.82 and |proj g| 7.71590D+01 for this
                   (notebooks/Notebook 2 ...ipynb:240,443).
  small_multishot  tests/golden/small_multishot.ini: 2 shots, stacked sources, duplicate
                   receiver position, 2x4 basis blocks, random heterogeneous medium.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle.oracle import RefModel, config_from_ini, write_ini  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
C1 = os.path.join(ROOT, "configs", "default_testing_configuration.ini")
SMALL = os.path.join(HERE, "small_multishot.ini")


def small_config():
    cfg = config_from_ini(C1)
    cfg.update(nt=400, nx_inner=60, nz_inner=40, nx_inner_boundary=4, nz_inner_boundary=6,
               np_boundary=10, np_factor=0.03, n_sources=3, n_shots=2,
               ix_sources=[10, 30, 50], iz_sources=[5, 5, 8], moment_angles=[90, 33.5, 180],
               which_source_to_fire_in_which_shot=[[0, 2], [1]],
               nr=4, ix_receivers=[5, 20, 20, 55], iz_receivers=[30, 30, 30, 5],
               snapshot_interval=5, npx=2, npz=4, delay_cycles_per_shot=3)
    return cfg


def small_medium(nx, nz):
    rng = np.random.default_rng(20261104)
    vp = 2000.0 * (1 + 0.05 * rng.standard_normal((nx, nz)))
    vs = 800.0 * (1 + 0.05 * rng.standard_normal((nx, nz)))
    rho = 1500.0 * (1 + 0.05 * rng.standard_normal((nx, nz)))
    return vp, vs, rho


def notebook2_target(m):
    """Notebook 2 cell 3 (three circles) restated."""
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
    for s in range(m.n_shots):
        m.forward_simulate(s)
    m.calculate_l2_misfit()
    m.calculate_l2_adjoint_sources()
    m.reset_kernels()
    for s in range(m.n_shots):
        m.adjoint_simulate(s)
    m.map_kernels_to_velocity()


def main():
    # ---- c1_homogeneous -------------------------------------------------------------
    m = RefModel(C1)
    m.forward_simulate(0, True)
    ux, uz = m.get_synthetic_data()
    snaps = {}
    for k in (100, 400, 799):
        for f in ("vx", "vz", "txx", "tzz", "txz"):
            a = m.array("accu_" + f)[0, k]
            snaps[f"snapnorm_{f}_{k}"] = np.linalg.norm(a)
    np.savez_compressed(
        os.path.join(HERE, "c1_homogeneous.npz"),
        ux_dec8=ux[:, :, ::8], uz_dec8=uz[:, :, ::8],
        sum_ux=ux.sum(), sum_uz=uz.sum(), norm_ux=np.linalg.norm(ux), norm_uz=np.linalg.norm(uz),
        snap_vx_400_dec=m.array("accu_vx")[0, 400, ::3, ::3],
        snap_txz_400_dec=m.array("accu_txz")[0, 400, ::3, ::3],
        final_vx_dec=m.array("vx")[::3, ::3], final_tzz_dec=m.array("tzz")[::3, ::3],
        taper_row=m.array("taper")[:, 75], stf_dec=m.array("stf")[:, ::16],
        moment=m.array("moment"), **snaps)
    print("c1_homogeneous: sum_ux", ux.sum(), "sum_uz", uz.sum(),
          "norms", np.linalg.norm(ux), np.linalg.norm(uz))

    # ---- c1_notebook2 -----------------------------------------------------------------
    start, target = notebook2_target(m)
    m.set_parameter_fields(*target)
    m.forward_simulate(0, True)
    ux_obs, uz_obs = m.get_synthetic_data()
    m.set_observed_data(ux_obs, uz_obs)
    m.set_parameter_fields(*start)
    gradient_workflow(m)
    g = m.get_gradient_vector()
    kv = m.get_kernels()
    np.savez_compressed(
        os.path.join(HERE, "c1_notebook2.npz"),
        misfit=m.misfit, g_norm=np.linalg.norm(g), g_maxabs=np.abs(g).max(),
        g_dec7=g[::7], kernel_absmax=np.array([np.abs(k).max() for k in kv]),
        obs_norm_ux=np.linalg.norm(ux_obs), obs_norm_uz=np.linalg.norm(uz_obs),
        obs_ux_dec8=ux_obs[:, :, ::8], obs_uz_dec8=uz_obs[:, :, ::8],
        lambda_kernel_dec=m.array("lambda_kernel")[::3, ::3],
        mu_kernel_dec=m.array("mu_kernel")[::3, ::3],
        density_l_kernel_dec=m.array("density_l_kernel")[::3, ::3],
        vs_kernel_dec=kv[1][::3, ::3])
    print("c1_notebook2: misfit", m.misfit, "|g|2", np.linalg.norm(g), "max|g|", np.abs(g).max(),
          "kernel absmax", [float(np.abs(k).max()) for k in kv])
    del m

    # ---- small_multishot -----------------------------------------------------------------
    write_ini(SMALL, small_config())
    m = RefModel(SMALL)
    vp, vs, rho = small_medium(m.nx, m.nz)
    m.set_parameter_fields(vp, vs, rho)
    for s in range(m.n_shots):
        m.forward_simulate(s, True)
    ux, uz = m.get_synthetic_data()
    obs_ux, obs_uz = 0.9 * ux, 1.1 * uz
    m.set_observed_data(obs_ux, obs_uz)
    model_vec = m.get_model_vector()
    snap = {f"snap_{f}": m.array("accu_" + f)[:, [0, 17, 60]].copy()
            for f in ("vx", "vz", "txx", "tzz", "txz")}
    fwd_end = {f"fwd_end_{f}": m.array(f).copy() for f in ("vx", "vz", "txx", "tzz", "txz")}
    gradient_workflow(m)
    np.savez_compressed(
        os.path.join(HERE, "small_multishot.npz"),
        vp=vp, vs=vs, rho=rho, ux=ux, uz=uz, obs_ux=obs_ux, obs_uz=obs_uz,
        misfit=m.misfit, a_stf_ux=m.array("a_stf_ux"), a_stf_uz=m.array("a_stf_uz"),
        lambda_kernel=m.array("lambda_kernel"), mu_kernel=m.array("mu_kernel"),
        density_l_kernel=m.array("density_l_kernel"), vp_kernel=m.array("vp_kernel"),
        vs_kernel=m.array("vs_kernel"), density_v_kernel=m.array("density_v_kernel"),
        gradient=m.get_gradient_vector(), model_vector=model_vec,
        taper=m.array("taper"), stf=m.array("stf"), moment=m.array("moment"),
        lm=m.array("lm"), la=m.array("la"), mu=m.array("mu"), b_vx=m.array("b_vx"),
        adj_end_vx=m.array("vx"), adj_end_txx=m.array("txx"), **snap, **fwd_end)
    print("small_multishot: misfit", m.misfit)


if __name__ == "__main__":
    main()
