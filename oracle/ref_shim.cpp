// TEST INFRASTRUCTURE ONLY -- never linked, imported or executed by the product path.
//
// extern "C" shim over the UNMODIFIED reference solver core.  It is compiled together
// with the reference's own sources where they lie (/root/reference/src/fdModel.cpp,
// /root/reference/src/contiguous_arrays.cpp; headers from ext/eigen, ext/inih) by
// oracle/Makefile into oracle/_ref/libpsv_ref.so.  No reference source is copied into
// this repository: this file only *calls* the reference's public class
// (src/fdModel.h:35-318) and hands out pointers to its public array fields
// (src/fdModel.h:209-252), so that
//   * oracle/psv_oracle.c (our CPU restatement) can be pinned against the real thing,
//   * tests/golden/*.npz can be generated (tests/golden/make_golden.py),
//   * bench.py --impl reference can time the reference's own OpenMP loops.
#include <cstring>
#include <exception>
#include <string>
#include <vector>

#include "fdModel.h"

namespace {
thread_local std::string g_err;

template <class F>
int guarded(F &&f) {
  try {
    f();
    return 0;
  } catch (const std::invalid_argument &e) {
    g_err = e.what();
    return 1;
  } catch (const std::length_error &e) {
    g_err = e.what();
    return 2;
  } catch (const std::exception &e) {
    g_err = e.what();
    return 3;
  }
}
} // namespace

extern "C" {

const char *ref_last_error() { return g_err.c_str(); }

void *ref_create(const char *ini_path) {
  fdModel *m = nullptr;
  if (guarded([&] { m = new fdModel(ini_path); }) != 0) return nullptr;
  return m;
}

void *ref_clone(void *h) {
  fdModel *m = nullptr;
  if (guarded([&] { m = new fdModel(*static_cast<fdModel *>(h)); }) != 0) return nullptr;
  return m;
}

void ref_destroy(void *h) { delete static_cast<fdModel *>(h); }

long ref_get_int(void *h, const char *name) {
  fdModel *m = static_cast<fdModel *>(h);
  const std::string n(name);
#define I(field) \
  if (n == #field) return m->field;
  I(nt) I(nx) I(nz) I(nx_inner) I(nz_inner) I(nx_inner_boundary) I(nz_inner_boundary)
  I(np_boundary) I(n_sources) I(n_shots) I(nr) I(snapshot_interval) I(snapshots)
  I(nx_free_parameters) I(nz_free_parameters) I(free_parameters)
  I(basis_gridpoints_x) I(basis_gridpoints_z)
#undef I
  return -999999;
}

double ref_get_real(void *h, const char *name) {
  fdModel *m = static_cast<fdModel *>(h);
  const std::string n(name);
#define R(field) \
  if (n == #field) return m->field;
  R(dx) R(dz) R(dt) R(np_factor) R(scalar_rho) R(scalar_vp) R(scalar_vs)
  R(peak_frequency) R(alpha) R(t0) R(delay_cycles_per_shot) R(misfit)
#undef R
  return -999999.0;
}

// Pointer to one of the reference's public arrays plus its element count.
double *ref_array(void *h, const char *name, long long *count) {
  fdModel *m = static_cast<fdModel *>(h);
  const std::string n(name);
  const long long grid = (long long)m->nx * m->nz;
  const long long rec = (long long)m->n_shots * m->nr * m->nt;
  const long long accu = (long long)m->n_shots * m->snapshots * grid;
#define A(field, cnt) \
  if (n == #field) {  \
    *count = (cnt);   \
    return m->field;  \
  }
  A(vx, grid) A(vz, grid) A(txx, grid) A(tzz, grid) A(txz, grid)
  A(lm, grid) A(la, grid) A(mu, grid) A(b_vx, grid) A(b_vz, grid)
  A(rho, grid) A(vp, grid) A(vs, grid)
  A(lambda_kernel, grid) A(mu_kernel, grid) A(density_l_kernel, grid)
  A(vp_kernel, grid) A(vs_kernel, grid) A(density_v_kernel, grid)
  A(taper, grid)
  A(t, (long long)m->nt) A(stf, (long long)m->n_sources * m->nt)
  A(moment, (long long)m->n_sources * 4)
  A(rtf_ux, rec) A(rtf_uz, rec) A(rtf_ux_true, rec) A(rtf_uz_true, rec)
  A(a_stf_ux, rec) A(a_stf_uz, rec)
  A(accu_vx, accu) A(accu_vz, accu) A(accu_txx, accu) A(accu_tzz, accu) A(accu_txz, accu)
#undef A
  *count = 0;
  return nullptr;
}

int *ref_int_array(void *h, const char *name, long long *count) {
  fdModel *m = static_cast<fdModel *>(h);
  const std::string n(name);
#define A(field, cnt) \
  if (n == #field) {  \
    *count = (cnt);   \
    return m->field;  \
  }
  A(ix_sources, m->n_sources) A(iz_sources, m->n_sources)
  A(ix_receivers, m->nr) A(iz_receivers, m->nr)
#undef A
  *count = 0;
  return nullptr;
}

int ref_forward(void *h, int i_shot, int store_fields, int verbose) {
  return guarded([&] {
    static_cast<fdModel *>(h)->forward_simulate(i_shot, store_fields != 0, verbose != 0);
  });
}

int ref_adjoint(void *h, int i_shot, int verbose) {
  return guarded([&] { static_cast<fdModel *>(h)->adjoint_simulate(i_shot, verbose != 0); });
}

int ref_update_from_velocity(void *h) {
  return guarded([&] { static_cast<fdModel *>(h)->update_from_velocity(); });
}
int ref_calculate_l2_misfit(void *h) {
  return guarded([&] { static_cast<fdModel *>(h)->calculate_l2_misfit(); });
}
int ref_calculate_l2_adjoint_sources(void *h) {
  return guarded([&] { static_cast<fdModel *>(h)->calculate_l2_adjoint_sources(); });
}
int ref_reset_kernels(void *h) {
  return guarded([&] { static_cast<fdModel *>(h)->reset_kernels(); });
}
int ref_map_kernels_to_velocity(void *h) {
  return guarded([&] { static_cast<fdModel *>(h)->map_kernels_to_velocity(); });
}
int ref_run_model(void *h, int verbose, int simulate_adjoint) {
  return guarded(
      [&] { static_cast<fdModel *>(h)->run_model(verbose != 0, simulate_adjoint != 0); });
}

int ref_get_model_vector(void *h, double *out, long long n) {
  return guarded([&] {
    dynamic_vector v = static_cast<fdModel *>(h)->get_model_vector();
    if (v.size() != n) throw std::invalid_argument("model vector size mismatch");
    std::memcpy(out, v.data(), sizeof(double) * n);
  });
}
int ref_set_model_vector(void *h, const double *in, long long n) {
  return guarded([&] {
    dynamic_vector v(n);
    std::memcpy(v.data(), in, sizeof(double) * n);
    static_cast<fdModel *>(h)->set_model_vector(v);
  });
}
int ref_get_gradient_vector(void *h, double *out, long long n) {
  return guarded([&] {
    dynamic_vector v = static_cast<fdModel *>(h)->get_gradient_vector();
    if (v.size() != n) throw std::invalid_argument("gradient vector size mismatch");
    std::memcpy(out, v.data(), sizeof(double) * n);
  });
}

int ref_write_receivers(void *h) {
  return guarded([&] { static_cast<fdModel *>(h)->write_receivers(); });
}
int ref_load_receivers(void *h, int verbose) {
  return guarded([&] { static_cast<fdModel *>(h)->load_receivers(verbose != 0); });
}
int ref_write_sources(void *h) {
  return guarded([&] { static_cast<fdModel *>(h)->write_sources(); });
}
// files in the current directory: kernel_vp.txt kernel_vs.txt kernel_density.txt (src/fdModel.cpp:1326-1356)
int ref_write_kernels(void *h) {
  return guarded([&] { static_cast<fdModel *>(h)->write_kernels(); });
}
int ref_load_model(void *h, const char *de_path, const char *vp_path, const char *vs_path, int verbose) {
  return guarded([&] { static_cast<fdModel *>(h)->load_model(de_path, vp_path, vs_path, verbose != 0); });
}
// forward run with the wavefield dumps into ./snapshots/ (src/fdModel.cpp:753-776)
int ref_forward_dump(void *h, int i_shot, int store_fields, int output_wavefields) {
  return guarded([&] {
    static_cast<fdModel *>(h)->forward_simulate(i_shot, store_fields != 0, false, output_wavefields != 0);
  });
}

// The explicit 30-argument constructor (src/fdModel.h:50-64), fed from the same flat record the
// C ABI's psv_create takes (include/psvwave_b200.h: psv_config).
struct ref_config {
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
  const int *shot_offsets;
  const int *shot_sources;
  int nr;
  const int *ix_receivers, *iz_receivers;
  int snapshot_interval;
  const char *observed_data_folder, *stf_folder;
};

void *ref_create_explicit(const ref_config *c) {
  fdModel *m = nullptr;
  const int rc = guarded([&] {
    std::vector<std::vector<int>> which;
    for (int s = 0; s < c->n_shots; ++s)
      which.emplace_back(c->shot_sources + c->shot_offsets[s], c->shot_sources + c->shot_offsets[s + 1]);
    m = new fdModel(c->nt, c->nx_inner, c->nz_inner, c->nx_inner_boundary, c->nz_inner_boundary, c->dx, c->dz, c->dt,
                    c->np_boundary, c->np_factor, c->scalar_rho, c->scalar_vp, c->scalar_vs, c->npx, c->npz,
                    c->peak_frequency, c->source_timeshift, c->delay_cycles_per_shot, c->n_sources, c->n_shots,
                    std::vector<int>(c->ix_sources, c->ix_sources + c->n_sources),
                    std::vector<int>(c->iz_sources, c->iz_sources + c->n_sources),
                    std::vector<double>(c->moment_angles, c->moment_angles + c->n_sources), which, c->nr,
                    std::vector<int>(c->ix_receivers, c->ix_receivers + c->nr),
                    std::vector<int>(c->iz_receivers, c->iz_receivers + c->nr), c->snapshot_interval,
                    c->observed_data_folder, c->stf_folder);
  });
  return rc == 0 ? m : nullptr;
}

} // extern "C"
