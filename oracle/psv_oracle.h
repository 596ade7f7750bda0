/* TEST INFRASTRUCTURE ONLY (see psv_oracle.c).  C interface of the CPU restatement. */
#ifndef PSV_ORACLE_H
#define PSV_ORACLE_H

#ifdef __cplusplus
extern "C" {
#endif

/* The reference's 30-argument constructor (src/fdModel.h:50-64) as a plain struct.
 * which_source_to_fire_in_which_shot is given in CSR form: the sources of shot s are
 * shot_sources[shot_offsets[s] .. shot_offsets[s+1]). */
typedef struct orc_config {
  int nt, nx_inner, nz_inner, nx_inner_boundary, nz_inner_boundary;
  double dx, dz, dt;
  int np_boundary;
  double np_factor;
  double scalar_rho, scalar_vp, scalar_vs;
  int npx, npz;
  double peak_frequency, source_timeshift, delay_cycles_per_shot;
  int n_sources, n_shots;
  const int *ix_sources, *iz_sources;
  const double *moment_angles;
  const int *shot_offsets, *shot_sources;
  int nr;
  const int *ix_receivers, *iz_receivers;
  int snapshot_interval;
} orc_config;

typedef struct orc_model orc_model;

const char *orc_check_config(const orc_config *c);
orc_model *orc_create(const orc_config *c);
void orc_destroy(orc_model *m);

int orc_forward(orc_model *m, int i_shot, int store_fields);
int orc_adjoint(orc_model *m, int i_shot);
void orc_update_from_velocity(orc_model *m);
double orc_calculate_l2_misfit(orc_model *m);
void orc_calculate_l2_adjoint_sources(orc_model *m);
void orc_reset_kernels(orc_model *m);
void orc_map_kernels_to_velocity(orc_model *m);
void orc_run_model(orc_model *m, int simulate_adjoint);
int orc_get_model_vector(orc_model *m, double *out, long long n);
int orc_set_model_vector(orc_model *m, const double *in, long long n);
int orc_get_gradient_vector(orc_model *m, double *out, long long n);

long orc_get_int(orc_model *m, const char *name);
double orc_get_real(orc_model *m, const char *name);
double *orc_array(orc_model *m, const char *name, long long *count);
int *orc_int_array(orc_model *m, const char *name, long long *count);

#ifdef __cplusplus
}
#endif
#endif
