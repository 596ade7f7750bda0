/* TEST INFRASTRUCTURE ONLY -- the product path (psvwave_b200/) never links, imports or
 * executes anything in oracle/.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may use it, and only as the checker / baseline.
 *
 * Plain-C restatement of the psvWave P-SV velocity-stress time loop, written from the
 * reference's algorithm (file:line citations are into /root/reference/).  Every
 * floating-point expression keeps the reference's evaluation order, so that when this
 * file is compiled the way the reference is (x86-64 SSE2, no FMA contraction, no
 * fast-math) the results are BIT-IDENTICAL to the reference build; tests/test_oracle.py
 * asserts exactly that against oracle/_ref/libpsv_ref.so (the unmodified reference core)
 * whenever that library is present, and against the tests/golden fixtures (generated from the
 * reference by tests/golden/make_golden.py) everywhere.  Parity status: PINNED.
 *
 * Layout contract (src/contiguous_arrays.cpp:14-21): grid arrays are row-major [ix][iz]
 * with iz fastest; receiver arrays [shot][receiver][it]; snapshots [shot][snap][ix][iz].
 * Unlike the reference (int indices, src/contiguous_arrays.h:9-14) all indices here are
 * 64-bit.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "psv_oracle.h"

#define ORC_PI_TRUNC 3.14159265 /* src/fdModel.cpp:14 -- used for the moment tensor only */

typedef int64_t i64;

struct orc_model {
  orc_config c;
  /* derived geometry (src/fdModel.cpp:398-407,460-461) */
  int nx, nz, nx_free, nz_free, free_parameters, snapshots;
  double alpha;
  int *ix_src, *iz_src, *ix_rec, *iz_rec, *shot_off, *shot_src;
  double *angles;
  /* grid arrays */
  double *vx, *vz, *txx, *tzz, *txz;
  double *lm, *la, *mu, *b_vx, *b_vz, *rho, *vp, *vs;
  double *lambda_kernel, *mu_kernel, *density_l_kernel;
  double *vp_kernel, *vs_kernel, *density_v_kernel;
  double *taper;
  /* signals */
  double *t, *stf, *moment;
  double *rtf_ux, *rtf_uz, *rtf_ux_true, *rtf_uz_true, *a_stf_ux, *a_stf_uz;
  double *accu_vx, *accu_vz, *accu_txx, *accu_tzz, *accu_txz;
  double misfit;
};

static double *dalloc(i64 n) { return (double *)calloc((size_t)(n > 0 ? n : 1), sizeof(double)); }
static int *ialloc(i64 n) { return (int *)calloc((size_t)(n > 0 ? n : 1), sizeof(int)); }

#define G(m, ix, iz) ((i64)(ix) * (m)->nz + (iz))

/* ---- setup ---------------------------------------------------------------------- */

/* src/fdModel.cpp:1034-1051 */
void orc_update_from_velocity(orc_model *m) {
  const i64 n = (i64)m->nx * m->nz;
  for (i64 i = 0; i < n; ++i) {
    m->mu[i] = pow(m->vs[i], 2) * m->rho[i];
    m->lm[i] = pow(m->vp[i], 2) * m->rho[i];
    m->la[i] = m->lm[i] - 2 * m->mu[i];
    m->b_vx[i] = 1.0 / m->rho[i];
    m->b_vz[i] = m->b_vx[i];
  }
}

/* src/fdModel.cpp:193-299 (initialize_arrays) */
static void orc_initialize(orc_model *m) {
  const orc_config *c = &m->c;
  const int nx = m->nx, nz = m->nz, nt = c->nt;
  /* :197-212 indices in the ini are relative to the physical domain */
  for (int r = 0; r < c->nr; ++r) {
    m->ix_rec[r] += c->np_boundary;
    m->iz_rec[r] += c->np_boundary;
  }
  for (int s = 0; s < c->n_sources; ++s) {
    m->ix_src[s] += c->np_boundary;
    m->iz_src[s] += c->np_boundary;
  }
  /* :215-231 Ricker wavelet; the delay counts the source's position inside its shot */
  for (int shot = 0; shot < c->n_shots; ++shot) {
    for (int k = 0; k < m->shot_off[shot + 1] - m->shot_off[shot]; ++k) {
      const int src = m->shot_src[m->shot_off[shot] + k];
      for (unsigned int it = 0; it < (unsigned int)nt; ++it) {
        m->t[it] = it * c->dt;
        const double f = 1.0 / m->alpha;
        const double tau = m->t[it] - 1.4 / f - c->delay_cycles_per_shot * k / f;
        m->stf[(i64)src * nt + it] =
            (1 - 2 * pow(M_PI * f * tau, 2)) * exp(-pow(M_PI * f * tau, 2));
      }
    }
  }
  /* :234-246 rotated double-couple, truncated pi */
  for (int s = 0; s < c->n_sources; ++s) {
    m->moment[s * 4 + 0] = cos(m->angles[s] * ORC_PI_TRUNC / 180.0) * 1e15;
    m->moment[s * 4 + 1] = -sin(m->angles[s] * ORC_PI_TRUNC / 180.0) * 1e15;
    m->moment[s * 4 + 2] = -sin(m->angles[s] * ORC_PI_TRUNC / 180.0) * 1e15;
    m->moment[s * 4 + 3] = -cos(m->angles[s] * ORC_PI_TRUNC / 180.0) * 1e15;
  }
  /* :249-261 homogeneous medium */
  for (i64 i = 0; i < (i64)nx * nz; ++i) {
    m->vp[i] = c->scalar_vp;
    m->vs[i] = c->scalar_vs;
    m->rho[i] = c->scalar_rho;
  }
  orc_update_from_velocity(m);
  /* :263-298 the taper counts, for every point, how many of the np_boundary nested
   * frames contain it (= min(np, 1 + distance to the nearest edge)), then maps the
   * count through a Gaussian. */
  for (int ix = 0; ix < nx; ++ix) {
    for (int iz = 0; iz < nz; ++iz) {
      int d = ix;
      if (nx - 1 - ix < d) d = nx - 1 - ix;
      if (iz < d) d = iz;
      if (nz - 1 - iz < d) d = nz - 1 - iz;
      int layers = d + 1;
      if (layers > c->np_boundary) layers = c->np_boundary;
      if (layers < 0) layers = 0;
      const double cnt = (double)layers;
      m->taper[G(m, ix, iz)] = exp(-pow(c->np_factor * (c->np_boundary - cnt), 2));
    }
  }
}

const char *orc_check_config(const orc_config *c) {
  /* src/fdModel.cpp:414-452,461-467 */
  if (c->nt <= 0 || c->snapshot_interval <= 0) return "nt and snapshot_interval must be positive";
  if (c->n_shots < 0 || c->n_sources < 0 || c->nr < 0) return "negative counts";
  if (c->shot_offsets[c->n_shots] != c->n_sources)
    return "Mismatch between n_sources and sources.which_source_to_fire_in_which_shot.";
  if (floor((double)c->nt / c->snapshot_interval) != (double)(c->nt / c->snapshot_interval))
    return "Snapshot interval and size of accumulator don't match!";
  return NULL;
}

orc_model *orc_create(const orc_config *c) {
  if (orc_check_config(c)) return NULL;
  orc_model *m = (orc_model *)calloc(1, sizeof(orc_model));
  m->c = *c;
  m->nx = c->nx_inner + 2 * c->np_boundary;
  m->nz = c->nz_inner + 2 * c->np_boundary;
  m->nx_free = c->nx_inner - 2 * c->nx_inner_boundary;
  m->nz_free = c->nz_inner - 2 * c->nz_inner_boundary;
  m->free_parameters = 3 * m->nx_free * m->nz_free / (c->npx * c->npz);
  m->alpha = 1.0 / c->peak_frequency;
  m->snapshots = c->nt / c->snapshot_interval;
  const i64 grid = (i64)m->nx * m->nz;
  const i64 rec = (i64)c->n_shots * c->nr * c->nt;
  const i64 accu = (i64)c->n_shots * m->snapshots * grid;

  m->ix_src = ialloc(c->n_sources);
  m->iz_src = ialloc(c->n_sources);
  m->angles = dalloc(c->n_sources);
  m->shot_off = ialloc(c->n_shots + 1);
  m->shot_src = ialloc(c->n_sources);
  m->ix_rec = ialloc(c->nr);
  m->iz_rec = ialloc(c->nr);
  memcpy(m->ix_src, c->ix_sources, sizeof(int) * c->n_sources);
  memcpy(m->iz_src, c->iz_sources, sizeof(int) * c->n_sources);
  memcpy(m->angles, c->moment_angles, sizeof(double) * c->n_sources);
  memcpy(m->shot_off, c->shot_offsets, sizeof(int) * (c->n_shots + 1));
  memcpy(m->shot_src, c->shot_sources, sizeof(int) * c->n_sources);
  memcpy(m->ix_rec, c->ix_receivers, sizeof(int) * c->nr);
  memcpy(m->iz_rec, c->iz_receivers, sizeof(int) * c->nr);
  /* the copies above own the index lists from here on */
  m->c.ix_sources = m->ix_src;
  m->c.iz_sources = m->iz_src;
  m->c.moment_angles = m->angles;
  m->c.shot_offsets = m->shot_off;
  m->c.shot_sources = m->shot_src;
  m->c.ix_receivers = m->ix_rec;
  m->c.iz_receivers = m->iz_rec;

  double **grids[] = {&m->vx, &m->vz, &m->txx, &m->tzz, &m->txz, &m->lm, &m->la, &m->mu,
                      &m->b_vx, &m->b_vz, &m->rho, &m->vp, &m->vs, &m->lambda_kernel,
                      &m->mu_kernel, &m->density_l_kernel, &m->vp_kernel, &m->vs_kernel,
                      &m->density_v_kernel, &m->taper};
  for (size_t i = 0; i < sizeof(grids) / sizeof(grids[0]); ++i) *grids[i] = dalloc(grid);
  m->t = dalloc(c->nt);
  m->stf = dalloc((i64)c->n_sources * c->nt);
  m->moment = dalloc((i64)c->n_sources * 4);
  double **recs[] = {&m->rtf_ux, &m->rtf_uz, &m->rtf_ux_true, &m->rtf_uz_true, &m->a_stf_ux,
                     &m->a_stf_uz};
  for (size_t i = 0; i < 6; ++i) *recs[i] = dalloc(rec);
  double **accus[] = {&m->accu_vx, &m->accu_vz, &m->accu_txx, &m->accu_tzz, &m->accu_txz};
  for (size_t i = 0; i < 5; ++i) *accus[i] = dalloc(accu);
  orc_initialize(m);
  return m;
}

void orc_destroy(orc_model *m) {
  if (!m) return;
  double *d[] = {m->vx, m->vz, m->txx, m->tzz, m->txz, m->lm, m->la, m->mu, m->b_vx, m->b_vz,
                 m->rho, m->vp, m->vs, m->lambda_kernel, m->mu_kernel, m->density_l_kernel,
                 m->vp_kernel, m->vs_kernel, m->density_v_kernel, m->taper, m->t, m->stf,
                 m->moment, m->rtf_ux, m->rtf_uz, m->rtf_ux_true, m->rtf_uz_true, m->a_stf_ux,
                 m->a_stf_uz, m->accu_vx, m->accu_vz, m->accu_txx, m->accu_tzz, m->accu_txz,
                 m->angles};
  for (size_t i = 0; i < sizeof(d) / sizeof(d[0]); ++i) free(d[i]);
  int *ip[] = {m->ix_src, m->iz_src, m->ix_rec, m->iz_rec, m->shot_off, m->shot_src};
  for (size_t i = 0; i < 6; ++i) free(ip[i]);
  free(m);
}

/* ---- the two stencil sweeps (shared by forward and adjoint; sdt = +dt or -dt) ------ */
/* Written as  f + sdt*(...)  for the forward run and  f - dt*(...)  for the adjoint run
 * exactly as the reference spells them (:615-642 vs :882-909); IEEE-754 makes
 * f - dt*X and f + (-dt)*X identical bit for bit, so one body with a signed step
 * serves both. */

/* src/fdModel.cpp:599-644 and :864-911 */
static void orc_stress_sweep(orc_model *m, double sdt) {
  const int nx = m->nx, nz = m->nz;
  const double dx = m->c.dx, dz = m->c.dz;
  const double c1 = 9.0 / 8.0, c2 = 1.0 / 24.0; /* src/fdModel.h:201-202 */
  const double *vx = m->vx, *vz = m->vz, *lm = m->lm, *la = m->la, *mu = m->mu, *tp = m->taper;
  double *txx = m->txx, *tzz = m->tzz, *txz = m->txz;
#pragma omp parallel for
  for (int ix = 2; ix < nx - 2; ++ix) {
    for (int iz = 2; iz < nz - 2; ++iz) {
      const i64 i = G(m, ix, iz);
      const i64 xp1 = i + nz, xp2 = i + 2 * (i64)nz, xm1 = i - nz, xm2 = i - 2 * (i64)nz;
      const i64 zp1 = i + 1, zp2 = i + 2, zm1 = i - 1, zm2 = i - 2;
      const double dxp_vx = c1 * (vx[xp1] - vx[i]) + c2 * (vx[xm1] - vx[xp2]);
      const double dzm_vz = c1 * (vz[i] - vz[zm1]) + c2 * (vz[zm2] - vz[zp1]);
      const double dzp_vx = c1 * (vx[zp1] - vx[i]) + c2 * (vx[zm1] - vx[zp2]);
      const double dxm_vz = c1 * (vz[i] - vz[xm1]) + c2 * (vz[xm2] - vz[xp1]);
      txx[i] = tp[i] * (txx[i] + sdt * (lm[i] * dxp_vx / dx + la[i] * dzm_vz / dz));
      tzz[i] = tp[i] * (tzz[i] + sdt * (la[i] * dxp_vx / dx + lm[i] * dzm_vz / dz));
      txz[i] = tp[i] * (txz[i] + sdt * mu[i] * (dzp_vx / dz + dxm_vz / dx));
    }
  }
}

/* src/fdModel.cpp:646-678 and :913-948.  Forward: b*dt*(..); adjoint: vx - b*dt*(..).
 * (b*dt)*X with a negative dt equals -(b*dt*X) exactly, so the signed form is exact. */
static void orc_velocity_sweep(orc_model *m, double sdt) {
  const int nx = m->nx, nz = m->nz;
  const double dx = m->c.dx, dz = m->c.dz;
  const double c1 = 9.0 / 8.0, c2 = 1.0 / 24.0;
  const double *txx = m->txx, *tzz = m->tzz, *txz = m->txz, *bx = m->b_vx, *bz = m->b_vz,
               *tp = m->taper;
  double *vx = m->vx, *vz = m->vz;
#pragma omp parallel for
  for (int ix = 2; ix < nx - 2; ++ix) {
    for (int iz = 2; iz < nz - 2; ++iz) {
      const i64 i = G(m, ix, iz);
      const i64 xp1 = i + nz, xp2 = i + 2 * (i64)nz, xm1 = i - nz, xm2 = i - 2 * (i64)nz;
      const i64 zp1 = i + 1, zp2 = i + 2, zm1 = i - 1, zm2 = i - 2;
      const double dxm_txx = c1 * (txx[i] - txx[xm1]) + c2 * (txx[xm2] - txx[xp1]);
      const double dzm_txz = c1 * (txz[i] - txz[zm1]) + c2 * (txz[zm2] - txz[zp1]);
      const double dxp_txz = c1 * (txz[xp1] - txz[i]) + c2 * (txz[xm1] - txz[xp2]);
      const double dzp_tzz = c1 * (tzz[zp1] - tzz[i]) + c2 * (tzz[zm1] - tzz[zp2]);
      vx[i] = tp[i] * (vx[i] + bx[i] * sdt * (dxm_txx / dx + dzm_txz / dz));
      vz[i] = tp[i] * (vz[i] + bz[i] * sdt * (dxp_txz / dx + dzp_tzz / dz));
    }
  }
}

static void orc_zero_fields(orc_model *m) {
  const size_t bytes = sizeof(double) * (size_t)m->nx * m->nz;
  memset(m->vx, 0, bytes);
  memset(m->vz, 0, bytes);
  memset(m->txx, 0, bytes);
  memset(m->tzz, 0, bytes);
  memset(m->txz, 0, bytes);
}

/* ---- forward (src/fdModel.cpp:526-787) --------------------------------------------- */
int orc_forward(orc_model *m, int i_shot, int store_fields) {
  const orc_config *c = &m->c;
  if (i_shot < 0 || i_shot >= c->n_shots) return 1;
  const int nt = c->nt, si = c->snapshot_interval, nz = m->nz;
  const double dx = c->dx, dz = c->dz, dt = c->dt;
  const i64 grid = (i64)m->nx * m->nz;
  orc_zero_fields(m);
  for (int it = 0; it < nt; ++it) {
    /* :555-572 snapshot = state before step `it` */
    if (it % si == 0 && store_fields) {
      const i64 off = ((i64)i_shot * m->snapshots + it / si) * grid;
      memcpy(m->accu_vx + off, m->vx, sizeof(double) * grid);
      memcpy(m->accu_vz + off, m->vz, sizeof(double) * grid);
      memcpy(m->accu_txx + off, m->txx, sizeof(double) * grid);
      memcpy(m->accu_tzz + off, m->tzz, sizeof(double) * grid);
      memcpy(m->accu_txz + off, m->txz, sizeof(double) * grid);
    }
    /* :576-596 displacement = running time integral of velocity */
    for (int r = 0; r < c->nr; ++r) {
      const i64 row = ((i64)i_shot * c->nr + r) * nt;
      const i64 p = G(m, m->ix_rec[r], m->iz_rec[r]);
      if (it == 0) {
        m->rtf_ux[row] = dt * m->vx[p] / (dx * dz);
        m->rtf_uz[row] = dt * m->vz[p] / (dx * dz);
      } else {
        m->rtf_ux[row + it] = m->rtf_ux[row + it - 1] + dt * m->vx[p] / (dx * dz);
        m->rtf_uz[row + it] = m->rtf_uz[row + it - 1] + dt * m->vz[p] / (dx * dz);
      }
    }
    orc_stress_sweep(m, dt);
    orc_velocity_sweep(m, dt);
    /* :681-751 moment-tensor injection, sources of the shot in list order; note the
     * reference scales every component with b_vz at the target point. */
    for (int k = m->shot_off[i_shot]; k < m->shot_off[i_shot + 1]; ++k) {
      const int s = m->shot_src[k];
      const double w = m->stf[(i64)s * nt + it];
      const double *M = m->moment + s * 4;
      const int sx = m->ix_src[s], sz = m->iz_src[s];
      const i64 p = G(m, sx, sz);
      const i64 xm1 = p - nz, xp1 = p + nz, zm1 = p - 1, zp1 = p + 1;
      const i64 xp1zm1 = p + nz - 1, xm1zp1 = p - nz + 1, xm1zm1 = p - nz - 1;
      const double dx4 = dx * dx * dx * dx, dz4 = dz * dz * dz * dz;
      const double *b = m->b_vz;
      /* (x,x) */
      m->vx[xm1] -= M[0] * w * dt * b[xm1] / dx4;
      m->vx[p] += M[0] * w * dt * b[p] / dx4;
      /* (z,z) */
      m->vz[zm1] -= M[3] * w * dt * b[zm1] / dz4;
      m->vz[p] += M[3] * w * dt * b[p] / dz4;
      /* (x,z) */
      m->vx[xm1zp1] += 0.25 * M[1] * w * dt * b[xm1zp1] / dx4;
      m->vx[zp1] += 0.25 * M[1] * w * dt * b[zp1] / dx4;
      m->vx[xm1zm1] -= 0.25 * M[1] * w * dt * b[xm1zm1] / dx4;
      m->vx[zm1] -= 0.25 * M[1] * w * dt * b[zm1] / dx4;
      /* (z,x) */
      m->vz[xp1zm1] += 0.25 * M[2] * w * dt * b[xp1zm1] / dz4;
      m->vz[xp1] += 0.25 * M[2] * w * dt * b[xp1] / dz4;
      m->vz[xm1zm1] -= 0.25 * M[2] * w * dt * b[xm1zm1] / dz4;
      m->vz[xm1] -= 0.25 * M[2] * w * dt * b[xm1] / dz4;
    }
  }
  return 0;
}

/* ---- adjoint (src/fdModel.cpp:789-969) --------------------------------------------- */
int orc_adjoint(orc_model *m, int i_shot) {
  const orc_config *c = &m->c;
  if (i_shot < 0 || i_shot >= c->n_shots) return 1;
  const int nt = c->nt, si = c->snapshot_interval;
  const double dx = c->dx, dz = c->dz, dt = c->dt;
  const i64 grid = (i64)m->nx * m->nz;
  const int x0 = c->np_boundary + c->nx_inner_boundary;
  const int x1 = c->np_boundary + c->nx_inner - c->nx_inner_boundary;
  const int z0 = c->np_boundary + c->nz_inner_boundary;
  const int z1 = c->np_boundary + c->nz_inner - c->nz_inner_boundary;
  orc_zero_fields(m);
  for (int it = nt - 1; it >= 0; --it) {
    /* :817-861 correlate the stored forward state with the current adjoint state */
    if (it % si == 0) {
      const i64 off = ((i64)i_shot * m->snapshots + it / si) * grid;
      const double *fvx = m->accu_vx + off, *fvz = m->accu_vz + off, *fxx = m->accu_txx + off,
                   *fzz = m->accu_tzz + off, *fxz = m->accu_txz + off;
#pragma omp parallel for
      for (int ix = x0; ix < x1; ++ix) {
        for (int iz = z0; iz < z1; ++iz) {
          const i64 i = G(m, ix, iz);
          const double la = m->la[i], lm = m->lm[i], mu = m->mu[i];
          const double txx = m->txx[i], tzz = m->tzz[i], txz = m->txz[i];
          m->density_l_kernel[i] -= si * dt * (fvx[i] * m->vx[i] + fvz[i] * m->vz[i]);
          m->lambda_kernel[i] +=
              si * dt *
              (((fxx[i] - (fzz[i] * la) / lm) + (fzz[i] - (fxx[i] * la) / lm)) *
               ((txx - (tzz * la) / lm) + (tzz - (txx * la) / lm))) /
              ((lm - ((la * la) / (lm))) * (lm - ((la * la) / (lm))));
          m->mu_kernel[i] +=
              si * dt * 2 *
              ((((txx - (tzz * la) / lm) * (fxx[i] - (fzz[i] * la) / lm)) +
                ((tzz - (txx * la) / lm) * (fzz[i] - (fxx[i] * la) / lm))) /
                   ((lm - ((la * la) / (lm))) * (lm - ((la * la) / (lm)))) +
               2 * (txz * fxz[i] / (4 * mu * mu)));
        }
      }
    }
    orc_stress_sweep(m, -dt);
    orc_velocity_sweep(m, -dt);
    /* :951-958 adjoint sources at the receivers, in receiver order */
    for (int r = 0; r < c->nr; ++r) {
      const i64 p = G(m, m->ix_rec[r], m->iz_rec[r]);
      const i64 q = ((i64)i_shot * c->nr + r) * nt + it;
      m->vx[p] += dt * m->b_vx[p] * m->a_stf_ux[q] / (dx * dz);
      m->vz[p] += dt * m->b_vz[p] * m->a_stf_uz[q] / (dx * dz);
    }
  }
  return 0;
}

/* ---- misfit / adjoint sources / kernels ---------------------------------------------- */

/* src/fdModel.cpp:1128-1145 (serial, shot -> receiver -> time, ux term then uz term) */
double orc_calculate_l2_misfit(orc_model *m) {
  const i64 n = (i64)m->c.n_shots * m->c.nr * m->c.nt;
  const double dt = m->c.dt;
  double acc = 0;
  for (i64 i = 0; i < n; ++i) {
    acc += 0.5 * dt * pow(m->rtf_ux_true[i] - m->rtf_ux[i], 2);
    acc += 0.5 * dt * pow(m->rtf_uz_true[i] - m->rtf_uz[i], 2);
  }
  m->misfit = acc;
  return acc;
}

/* src/fdModel.cpp:1147-1162 */
void orc_calculate_l2_adjoint_sources(orc_model *m) {
  const i64 n = (i64)m->c.n_shots * m->c.nr * m->c.nt;
  for (i64 i = 0; i < n; ++i) {
    m->a_stf_ux[i] = m->rtf_ux[i] - m->rtf_ux_true[i];
    m->a_stf_uz[i] = m->rtf_uz[i] - m->rtf_uz_true[i];
  }
}

/* src/fdModel.cpp:1310-1324 (only the three Lame-basis kernels) */
void orc_reset_kernels(orc_model *m) {
  const size_t bytes = sizeof(double) * (size_t)m->nx * m->nz;
  memset(m->lambda_kernel, 0, bytes);
  memset(m->mu_kernel, 0, bytes);
  memset(m->density_l_kernel, 0, bytes);
}

/* src/fdModel.cpp:1164-1183 */
void orc_map_kernels_to_velocity(orc_model *m) {
  const i64 n = (i64)m->nx * m->nz;
  for (i64 i = 0; i < n; ++i) {
    const double vp = m->vp[i], vs = m->vs[i];
    m->vp_kernel[i] = 2 * vp * m->lambda_kernel[i] / m->b_vx[i];
    m->vs_kernel[i] = (2 * vs * m->mu_kernel[i] - 4 * vs * m->lambda_kernel[i]) / m->b_vx[i];
    m->density_v_kernel[i] = m->density_l_kernel[i] + (vp * vp - 2 * vs * vs) * m->lambda_kernel[i] +
                             vs * vs * m->mu_kernel[i];
  }
}

/* src/fdModel.cpp:1291-1308 */
void orc_run_model(orc_model *m, int simulate_adjoint) {
  for (int s = 0; s < m->c.n_shots; ++s) orc_forward(m, s, 1);
  orc_calculate_l2_misfit(m);
  if (simulate_adjoint) {
    orc_calculate_l2_adjoint_sources(m);
    orc_reset_kernels(m);
    for (int s = 0; s < m->c.n_shots; ++s) orc_adjoint(m, s);
    orc_map_kernels_to_velocity(m);
  }
}

/* ---- model / gradient vectors (src/fdModel.cpp:1358-1515) ------------------------------
 * The free window is cut into npx x npz blocks; block (bx,bz) is vector entry
 * bx + bz*(nx_free/npx) -- x fastest, i.e. transposed w.r.t. the grid layout -- and the
 * vector is [vp blocks | vs blocks | rho blocks]. */
typedef enum { VEC_GET_MEAN, VEC_SET_FILL, VEC_GET_SUM } vec_mode;

static int orc_vector_walk(orc_model *m, double *v, i64 n, vec_mode mode) {
  const orc_config *c = &m->c;
  const int npx = c->npx, npz = c->npz;
  const int nbx = m->nx_free / npx, nbz = m->nz_free / npz;
  const int per = m->nx_free * m->nz_free / (npx * npz);
  if (n != 3 * (i64)per) return 1;
  double *a0 = mode == VEC_GET_SUM ? m->vp_kernel : m->vp;
  double *a1 = mode == VEC_GET_SUM ? m->vs_kernel : m->vs;
  double *a2 = mode == VEC_GET_SUM ? m->density_v_kernel : m->rho;
  for (int bx = 0; bx < nbx; ++bx) {
    for (int bz = 0; bz < nbz; ++bz) {
      const int k = bx + bz * m->nx_free / npx;
      const int gx = bx * npx + c->nx_inner_boundary + c->np_boundary;
      const int gz = bz * npz + c->nz_inner_boundary + c->np_boundary;
      if (mode != VEC_SET_FILL) v[k] = v[k + per] = v[k + 2 * per] = 0;
      for (int sx = 0; sx < npx; ++sx) {
        for (int sz = 0; sz < npz; ++sz) {
          const i64 i = G(m, gx + sx, gz + sz);
          if (mode == VEC_GET_MEAN) { /* :1400-1406 */
            v[k] += a0[i] / (npx * npz);
            v[k + per] += a1[i] / (npx * npz);
            v[k + 2 * per] += a2[i] / (npx * npz);
          } else if (mode == VEC_GET_SUM) { /* :1501-1509 plain block sum */
            v[k] += a0[i];
            v[k + per] += a1[i];
            v[k + 2 * per] += a2[i];
          } else { /* :1449-1451 */
            a0[i] = v[k];
            a1[i] = v[k + per];
            a2[i] = v[k + 2 * per];
          }
        }
      }
    }
  }
  if (mode == VEC_SET_FILL) orc_update_from_velocity(m); /* :1456 */
  return 0;
}

int orc_get_model_vector(orc_model *m, double *out, long long n) {
  return orc_vector_walk(m, out, n, VEC_GET_MEAN);
}
int orc_set_model_vector(orc_model *m, const double *in, long long n) {
  return orc_vector_walk(m, (double *)in, n, VEC_SET_FILL);
}
int orc_get_gradient_vector(orc_model *m, double *out, long long n) {
  return orc_vector_walk(m, out, n, VEC_GET_SUM);
}

/* ---- introspection used by oracle/oracle.py -------------------------------------------- */
long orc_get_int(orc_model *m, const char *name) {
#define I(n, v) \
  if (!strcmp(name, n)) return (v);
  I("nt", m->c.nt) I("nx", m->nx) I("nz", m->nz) I("nx_inner", m->c.nx_inner)
  I("nz_inner", m->c.nz_inner) I("nx_inner_boundary", m->c.nx_inner_boundary)
  I("nz_inner_boundary", m->c.nz_inner_boundary) I("np_boundary", m->c.np_boundary)
  I("n_sources", m->c.n_sources) I("n_shots", m->c.n_shots) I("nr", m->c.nr)
  I("snapshot_interval", m->c.snapshot_interval) I("snapshots", m->snapshots)
  I("nx_free_parameters", m->nx_free) I("nz_free_parameters", m->nz_free)
  I("free_parameters", m->free_parameters) I("basis_gridpoints_x", m->c.npx)
  I("basis_gridpoints_z", m->c.npz)
#undef I
  return -999999;
}

double orc_get_real(orc_model *m, const char *name) {
#define R(n, v) \
  if (!strcmp(name, n)) return (v);
  R("dx", m->c.dx) R("dz", m->c.dz) R("dt", m->c.dt) R("np_factor", m->c.np_factor)
  R("scalar_rho", m->c.scalar_rho) R("scalar_vp", m->c.scalar_vp) R("scalar_vs", m->c.scalar_vs)
  R("peak_frequency", m->c.peak_frequency) R("alpha", m->alpha)
  R("t0", m->c.source_timeshift) R("delay_cycles_per_shot", m->c.delay_cycles_per_shot)
  R("misfit", m->misfit)
#undef R
  return -999999.0;
}

double *orc_array(orc_model *m, const char *name, long long *count) {
  const i64 grid = (i64)m->nx * m->nz;
  const i64 rec = (i64)m->c.n_shots * m->c.nr * m->c.nt;
  const i64 accu = (i64)m->c.n_shots * m->snapshots * grid;
#define A(field, cnt)         \
  if (!strcmp(name, #field)) { \
    *count = (cnt);           \
    return m->field;          \
  }
  A(vx, grid) A(vz, grid) A(txx, grid) A(tzz, grid) A(txz, grid)
  A(lm, grid) A(la, grid) A(mu, grid) A(b_vx, grid) A(b_vz, grid)
  A(rho, grid) A(vp, grid) A(vs, grid)
  A(lambda_kernel, grid) A(mu_kernel, grid) A(density_l_kernel, grid)
  A(vp_kernel, grid) A(vs_kernel, grid) A(density_v_kernel, grid) A(taper, grid)
  A(t, (i64)m->c.nt) A(stf, (i64)m->c.n_sources * m->c.nt) A(moment, (i64)m->c.n_sources * 4)
  A(rtf_ux, rec) A(rtf_uz, rec) A(rtf_ux_true, rec) A(rtf_uz_true, rec)
  A(a_stf_ux, rec) A(a_stf_uz, rec)
  A(accu_vx, accu) A(accu_vz, accu) A(accu_txx, accu) A(accu_tzz, accu) A(accu_txz, accu)
#undef A
  *count = 0;
  return NULL;
}

int *orc_int_array(orc_model *m, const char *name, long long *count) {
#define A(n, field, cnt) \
  if (!strcmp(name, n)) { \
    *count = (cnt);      \
    return m->field;     \
  }
  A("ix_sources", ix_src, m->c.n_sources) A("iz_sources", iz_src, m->c.n_sources)
  A("ix_receivers", ix_rec, m->c.nr) A("iz_receivers", iz_rec, m->c.nr)
#undef A
  *count = 0;
  return NULL;
}
