// extern "C" boundary (include/psvwave_b200.h) over psv::Model.
#include <algorithm>
#include <cstring>
#include <sstream>
#include <string>
#include <vector>

#include "model.hpp"
#include "multi.hpp"

struct psv_model {
  psv::Model *m;
  psv::DeviceGroup *g = nullptr;  // set by psv_set_devices with more than one device: `m` is then replica 0
};

namespace {
thread_local std::string g_error;
thread_local int g_default_device = -1;

// The calling thread's current CUDA device is left as it was found (a model binds its own device inside).
struct DeviceRestore {
  int device = -1;
  DeviceRestore() {
    if (cudaGetDevice(&device) != cudaSuccess) {
      cudaGetLastError();
      device = -1;
    }
  }
  ~DeviceRestore() {
    if (device >= 0 && cudaSetDevice(device) != cudaSuccess) cudaGetLastError();
  }
};

template <class F>
int guarded(F &&f) {
  DeviceRestore restore;
  try {
    f();
    return PSV_OK;
  } catch (const psv::length_error &e) {
    g_error = e.what();
    return PSV_ERR_LENGTH;
  } catch (const psv::invalid_argument &e) {
    g_error = e.what();
    return PSV_ERR_INVALID_ARGUMENT;
  } catch (const psv::cuda_error &e) {
    g_error = e.what();
    return PSV_ERR_CUDA;
  } catch (const psv::nccl_error &e) {
    g_error = e.what();
    return PSV_ERR_NCCL;
  } catch (const psv::state_error &e) {
    g_error = e.what();
    return PSV_ERR_STATE;
  } catch (const psv::io_error &e) {
    g_error = e.what();
    return PSV_ERR_IO;
  } catch (const std::bad_alloc &) {
    g_error = "out of host memory";
    return PSV_ERR_LENGTH;
  } catch (const std::exception &e) {
    g_error = e.what();
    return PSV_ERR_INVALID_ARGUMENT;
  }
}

psv::Model &model(psv_handle h) {
  if (!h || !h->m) throw psv::invalid_argument("Null psv_handle.");
  return *h->m;
}

// The group of a multi-GPU handle, or null (also checks the handle).
psv::DeviceGroup *group(psv_handle h) {
  model(h);
  return h->g;
}

void single_device_only(psv_handle h, const char *what) {
  if (group(h))
    throw psv::state_error(std::string(what) + " is not available on an object that spans several GPUs "
                                                "(psv_set_devices); use a single-device handle.");
}

void set_devices(psv_handle h, const int *ids, int n) {
  psv::Model &m = model(h);
  if (!ids || n < 1) throw psv::invalid_argument("psv_set_devices needs at least one device.");
  int count = 0;
  PSV_CUDA(cudaGetDeviceCount(&count));
  for (int i = 0; i < n; ++i)
    if (ids[i] < 0 || ids[i] >= count) throw psv::invalid_argument("CUDA device index out of range.");
  if (h->g) {  // back to one device first: synthetics and kernels onto the primary
    h->g->collect();
    delete h->g;
    h->g = nullptr;
  }
  if (ids[0] != m.device()) {
    psv::Model *moved = new psv::Model(m, ids[0], /*with_snapshots=*/true);
    delete h->m;
    h->m = moved;
  }
  if (n > 1) h->g = new psv::DeviceGroup(h->m, std::vector<int>(ids, ids + n));
}

// PSVWAVE_DEVICES = "all" | "0,1,2,...": every object created afterwards spans these GPUs, so that an unchanged
// psvWave script (one object, a loop over shots) fills the box.
void apply_device_environment(psv_handle h) {
  const char *e = std::getenv("PSVWAVE_DEVICES");
  if (!e || !*e) return;
  std::vector<int> ids;
  if (std::string(e) == "all") {
    int count = 0;
    PSV_CUDA(cudaGetDeviceCount(&count));
    const int first = h->m->device();
    for (int i = 0; i < count; ++i) ids.push_back((first + i) % count);
  } else {
    std::stringstream ss(e);
    std::string tok;
    while (std::getline(ss, tok, ',')) {
      try {
        ids.push_back(std::stoi(tok));
      } catch (...) {
        throw psv::invalid_argument("PSVWAVE_DEVICES must be \"all\" or a comma-separated list of device indices.");
      }
    }
  }
  if (ids.size() > 1 || (ids.size() == 1 && ids[0] != h->m->device())) set_devices(h, ids.data(), (int)ids.size());
}

psv_handle adopt(psv::Model *m) {
  psv_handle h = new psv_model{m};
  try {
    apply_device_environment(h);
  } catch (...) {
    delete h->g;
    delete h->m;
    delete h;
    throw;
  }
  return h;
}

int pick_device() {
  int count = 0;
  const cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count == 0) {
    cudaGetLastError();
    throw psv::cuda_error(std::string("psvwave_b200 needs a CUDA device (sm_100a); none is usable: ") +
                          (e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0") +
                          ". There is no CPU fallback.");
  }
  int dev = g_default_device;
  if (dev < 0) PSV_CUDA(cudaGetDevice(&dev));
  if (dev >= count) throw psv::invalid_argument("CUDA device index out of range.");
  return dev;
}
}  // namespace

extern "C" {

const char *psv_last_error(void) { return g_error.c_str(); }
const char *psv_version(void) { return "psvwave_b200 0.1 (sm_100a)"; }

int psv_variant_count(void) { return psv::Model::kVariantCount; }

int psv_set_default_device(int device) {
  g_default_device = device;
  return PSV_OK;
}

int psv_device_count(int *count) {
  return guarded([&] {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
      cudaGetLastError();
      n = 0;
    }
    *count = n;
  });
}

int psv_create_from_ini(const char *ini_path, psv_handle *out) {
  return guarded([&] {
    if (!ini_path || !out) throw psv::invalid_argument("Null argument.");
    *out = nullptr;
    psv::Config cfg = psv::config_from_ini(ini_path);
    const int dev = pick_device();
    *out = adopt(new psv::Model(std::move(cfg), dev));
  });
}

int psv_check_ini(const char *ini_path, char *summary, size_t capacity) {
  return guarded([&] {
    if (!ini_path) throw psv::invalid_argument("Null argument.");
    const psv::Config c = psv::config_from_ini(ini_path);
    std::ostringstream os;
    os.precision(17);
    auto list = [&](const char *name, const auto &v) {
      os << name << "=";
      for (size_t i = 0; i < v.size(); ++i) os << (i ? "," : "") << v[i];
      os << "\n";
    };
    os << "nt=" << c.nt << "\nnx_inner=" << c.nx_inner << "\nnz_inner=" << c.nz_inner
       << "\nnx_inner_boundary=" << c.nx_inner_boundary << "\nnz_inner_boundary=" << c.nz_inner_boundary
       << "\ndx=" << c.dx << "\ndz=" << c.dz << "\ndt=" << c.dt << "\nnp_boundary=" << c.np_boundary
       << "\nnp_factor=" << c.np_factor << "\nscalar_rho=" << c.scalar_rho << "\nscalar_vp=" << c.scalar_vp
       << "\nscalar_vs=" << c.scalar_vs << "\nnpx=" << c.npx << "\nnpz=" << c.npz
       << "\npeak_frequency=" << c.peak_frequency << "\nsource_timeshift=" << c.source_timeshift
       << "\ndelay_cycles_per_shot=" << c.delay_cycles_per_shot << "\nn_sources=" << c.n_sources
       << "\nn_shots=" << c.n_shots << "\nnr=" << c.nr << "\nsnapshot_interval=" << c.snapshot_interval
       << "\nnx=" << c.nx << "\nnz=" << c.nz << "\nnx_free_parameters=" << c.nx_free
       << "\nnz_free_parameters=" << c.nz_free << "\nfree_parameters=" << c.free_parameters
       << "\nsnapshots=" << c.snapshots << "\nobserved_data_folder=" << c.observed_data_folder
       << "\nstf_folder=" << c.stf_folder << "\n";
    list("ix_sources", c.ix_sources);
    list("iz_sources", c.iz_sources);
    list("moment_angles", c.moment_angles);
    list("ix_receivers", c.ix_receivers);
    list("iz_receivers", c.iz_receivers);
    os << "which_source_to_fire_in_which_shot=";
    for (size_t s = 0; s < c.which_source_to_fire_in_which_shot.size(); ++s) {
      const auto &v = c.which_source_to_fire_in_which_shot[s];
      os << (s ? "|" : "");
      for (size_t i = 0; i < v.size(); ++i) os << (i ? "," : "") << v[i];
    }
    os << "\n";
    const std::string text = os.str();
    if (summary && capacity) {
      const size_t n = std::min(capacity - 1, text.size());
      std::memcpy(summary, text.data(), n);
      summary[n] = 0;
    }
  });
}

int psv_create(const psv_config *c, psv_handle *out) {
  return guarded([&] {
    if (!c || !out) throw psv::invalid_argument("Null argument.");
    *out = nullptr;
    psv::Config g;
    g.nt = c->nt;
    g.nx_inner = c->nx_inner;
    g.nz_inner = c->nz_inner;
    g.nx_inner_boundary = c->nx_inner_boundary;
    g.nz_inner_boundary = c->nz_inner_boundary;
    g.dx = c->dx;
    g.dz = c->dz;
    g.dt = c->dt;
    g.np_boundary = c->np_boundary;
    g.np_factor = c->np_factor;
    g.scalar_rho = c->scalar_rho;
    g.scalar_vp = c->scalar_vp;
    g.scalar_vs = c->scalar_vs;
    g.npx = c->npx;
    g.npz = c->npz;
    g.peak_frequency = c->peak_frequency;
    g.source_timeshift = c->source_timeshift;
    g.delay_cycles_per_shot = c->delay_cycles_per_shot;
    g.n_sources = c->n_sources;
    g.n_shots = c->n_shots;
    if (c->n_sources < 0 || c->n_shots < 0 || c->nr < 0) throw psv::invalid_argument("Negative count.");
    if ((c->n_sources && (!c->ix_sources || !c->iz_sources || !c->moment_angles)) || !c->shot_offsets ||
        (c->nr && (!c->ix_receivers || !c->iz_receivers)))
      throw psv::invalid_argument("Null index list.");
    g.ix_sources.assign(c->ix_sources, c->ix_sources + c->n_sources);
    g.iz_sources.assign(c->iz_sources, c->iz_sources + c->n_sources);
    g.moment_angles.assign(c->moment_angles, c->moment_angles + c->n_sources);
    for (int s = 0; s < c->n_shots; ++s) {
      if (c->shot_offsets[s + 1] < c->shot_offsets[s]) throw psv::invalid_argument("shot_offsets must be non-decreasing.");
      g.which_source_to_fire_in_which_shot.emplace_back(c->shot_sources + c->shot_offsets[s],
                                                        c->shot_sources + c->shot_offsets[s + 1]);
    }
    g.nr = c->nr;
    g.ix_receivers.assign(c->ix_receivers, c->ix_receivers + c->nr);
    g.iz_receivers.assign(c->iz_receivers, c->iz_receivers + c->nr);
    g.snapshot_interval = c->snapshot_interval;
    if (c->observed_data_folder) g.observed_data_folder = c->observed_data_folder;
    if (c->stf_folder) g.stf_folder = c->stf_folder;
    g.finalize();
    const int dev = pick_device();
    *out = adopt(new psv::Model(std::move(g), dev));
  });
}

int psv_clone(psv_handle src, psv_handle *out) {
  return guarded([&] {
    if (!out) throw psv::invalid_argument("Null argument.");
    *out = nullptr;
    if (psv::DeviceGroup *g = group(src)) {
      g->drain();
      psv_handle h = new psv_model{new psv::Model(model(src))};
      try {
        h->g = new psv::DeviceGroup(*g, h->m);
      } catch (...) {
        delete h->m;
        delete h;
        throw;
      }
      *out = h;
      return;
    }
    *out = new psv_model{new psv::Model(model(src))};
  });
}

int psv_destroy(psv_handle h) {
  return guarded([&] {
    if (!h) return;
    delete h->g;  // joins the workers before the primary replica goes
    delete h->m;
    delete h;
  });
}

int psv_set_devices(psv_handle h, const int *device_ids, int n) {
  return guarded([&] { set_devices(h, device_ids, n); });
}
int psv_get_devices(psv_handle h, int *device_ids, int capacity, int *n) {
  return guarded([&] {
    psv::Model &m = model(h);
    std::vector<int> ids = h->g ? h->g->devices() : std::vector<int>{m.device()};
    if (n) *n = (int)ids.size();
    for (int i = 0; device_ids && i < capacity && i < (int)ids.size(); ++i) device_ids[i] = ids[i];
  });
}

int psv_forward_simulate(psv_handle h, int i_shot, int store_fields, int verbose, int output_wavefields) {
  return guarded([&] {
    if (psv::DeviceGroup *g = group(h)) g->forward_simulate(i_shot, store_fields != 0, verbose != 0, output_wavefields != 0);
    else model(h).forward_simulate(i_shot, store_fields != 0, verbose != 0, output_wavefields != 0);
  });
}
int psv_adjoint_simulate(psv_handle h, int i_shot, int verbose) {
  return guarded([&] {
    if (psv::DeviceGroup *g = group(h)) g->adjoint_simulate(i_shot, verbose != 0);
    else model(h).adjoint_simulate(i_shot, verbose != 0);
  });
}
int psv_calculate_l2_misfit(psv_handle h, double *misfit_out) {
  return guarded([&] {
    const double v = group(h) ? h->g->calculate_l2_misfit() : model(h).calculate_l2_misfit();
    if (misfit_out) *misfit_out = v;
  });
}
int psv_calculate_l2_misfit_shot(psv_handle h, int i_shot, double *misfit_out) {
  return guarded([&] {
    const double v = group(h) ? h->g->calculate_l2_misfit_shot(i_shot) : model(h).calculate_l2_misfit_shot(i_shot);
    if (misfit_out) *misfit_out = v;
  });
}
int psv_calculate_l2_adjoint_sources(psv_handle h) {
  return guarded([&] {
    if (psv::DeviceGroup *g = group(h)) g->calculate_l2_adjoint_sources();
    else model(h).calculate_l2_adjoint_sources();
  });
}
int psv_reset_kernels(psv_handle h) {
  return guarded([&] {
    if (psv::DeviceGroup *g = group(h)) g->reset_kernels();
    else model(h).reset_kernels();
  });
}
int psv_map_kernels_to_velocity(psv_handle h) {
  return guarded([&] {
    if (psv::DeviceGroup *g = group(h)) g->map_kernels_to_velocity();
    else model(h).map_kernels_to_velocity();
  });
}
int psv_update_from_velocity(psv_handle h) {
  return guarded([&] {
    if (psv::DeviceGroup *g = group(h)) g->update_from_velocity();
    else model(h).update_from_velocity();
  });
}
int psv_run_model(psv_handle h, int verbose, int simulate_adjoint) {
  return guarded([&] {
    if (psv::DeviceGroup *g = group(h)) g->run_model(verbose != 0, simulate_adjoint != 0);
    else model(h).run_model(verbose != 0, simulate_adjoint != 0);
  });
}

int psv_get_model_vector(psv_handle h, double *out, size_t n) {
  return guarded([&] {
    if (psv::DeviceGroup *g = group(h)) g->get_model_vector(out, n);
    else model(h).get_model_vector(out, n);
  });
}
int psv_set_model_vector(psv_handle h, const double *in, size_t n) {
  return guarded([&] {
    if (psv::DeviceGroup *g = group(h)) g->set_model_vector(in, n);
    else model(h).set_model_vector(in, n);
  });
}
int psv_get_gradient_vector(psv_handle h, double *out, size_t n) {
  return guarded([&] {
    if (psv::DeviceGroup *g = group(h)) g->get_gradient_vector(out, n);
    else model(h).get_gradient_vector(out, n);
  });
}

int psv_get_int(psv_handle h, psv_int_key key, long long *out) {
  return guarded([&] {
    const psv::Model &m = model(h);
    const psv::Config &c = m.config();
    switch (key) {
      case PSV_INT_NT: *out = c.nt; break;
      case PSV_INT_NX: *out = c.nx; break;
      case PSV_INT_NZ: *out = c.nz; break;
      case PSV_INT_NX_INNER: *out = c.nx_inner; break;
      case PSV_INT_NZ_INNER: *out = c.nz_inner; break;
      case PSV_INT_NX_INNER_BOUNDARY: *out = c.nx_inner_boundary; break;
      case PSV_INT_NZ_INNER_BOUNDARY: *out = c.nz_inner_boundary; break;
      case PSV_INT_NP_BOUNDARY: *out = c.np_boundary; break;
      case PSV_INT_N_SOURCES: *out = c.n_sources; break;
      case PSV_INT_N_SHOTS: *out = c.n_shots; break;
      case PSV_INT_NR: *out = c.nr; break;
      case PSV_INT_SNAPSHOT_INTERVAL: *out = c.snapshot_interval; break;
      case PSV_INT_SNAPSHOTS: *out = c.snapshots; break;
      case PSV_INT_NX_FREE_PARAMETERS: *out = c.nx_free; break;
      case PSV_INT_NZ_FREE_PARAMETERS: *out = c.nz_free; break;
      case PSV_INT_FREE_PARAMETERS: *out = c.free_parameters; break;
      case PSV_INT_BASIS_GRIDPOINTS_X: *out = c.npx; break;
      case PSV_INT_BASIS_GRIDPOINTS_Z: *out = c.npz; break;
      case PSV_INT_DEVICE: *out = m.device(); break;
      case PSV_INT_N_DEVICES: *out = h->g ? h->g->size() : 1; break;
      default: throw psv::invalid_argument("Unknown integer key.");
    }
  });
}

int psv_get_real(psv_handle h, psv_real_key key, double *out) {
  return guarded([&] {
    const psv::Model &m = model(h);
    const psv::Config &c = m.config();
    switch (key) {
      case PSV_REAL_DX: *out = c.dx; break;
      case PSV_REAL_DZ: *out = c.dz; break;
      case PSV_REAL_DT: *out = c.dt; break;
      case PSV_REAL_NP_FACTOR: *out = c.np_factor; break;
      case PSV_REAL_SCALAR_RHO: *out = c.scalar_rho; break;
      case PSV_REAL_SCALAR_VP: *out = c.scalar_vp; break;
      case PSV_REAL_SCALAR_VS: *out = c.scalar_vs; break;
      case PSV_REAL_PEAK_FREQUENCY: *out = c.peak_frequency; break;
      case PSV_REAL_ALPHA: *out = c.alpha; break;
      case PSV_REAL_T0: *out = c.source_timeshift; break;
      case PSV_REAL_DELAY_CYCLES_PER_SHOT: *out = c.delay_cycles_per_shot; break;
      case PSV_REAL_MISFIT: *out = h->g ? h->g->misfit() : m.misfit(); break;
      default: throw psv::invalid_argument("Unknown real key.");
    }
  });
}

int psv_array_size(psv_handle h, psv_array which, size_t *count) {
  return guarded([&] { *count = model(h).array_size(which); });
}
int psv_get_array(psv_handle h, psv_array which, double *out, size_t count) {
  return guarded([&] {
    if (psv::DeviceGroup *g = group(h)) g->get_array(which, out, count);
    else model(h).get_array(which, out, count);
  });
}
int psv_set_array(psv_handle h, psv_array which, const double *in, size_t count) {
  return guarded([&] {
    if (psv::DeviceGroup *g = group(h)) g->set_array(which, in, count);
    else model(h).set_array(which, in, count);
  });
}
int psv_set_parameter_fields(psv_handle h, const double *vp, const double *vs, const double *rho, size_t count) {
  return guarded([&] {
    if (psv::DeviceGroup *g = group(h)) g->set_parameter_fields(vp, vs, rho, count);
    else model(h).set_parameter_fields(vp, vs, rho, count);
  });
}

namespace {
std::vector<int> index_list(const psv::Config &c, psv_index_list which) {
  switch (which) {
    case PSV_IX_SOURCES: return c.ix_sources;
    case PSV_IZ_SOURCES: return c.iz_sources;
    case PSV_IX_RECEIVERS: return c.ix_receivers;
    case PSV_IZ_RECEIVERS: return c.iz_receivers;
    case PSV_SHOT_OFFSETS: {
      std::vector<int> off(1, 0);
      for (const auto &s : c.which_source_to_fire_in_which_shot) off.push_back(off.back() + (int)s.size());
      return off;
    }
    case PSV_SHOT_SOURCES: {
      std::vector<int> flat;
      for (const auto &s : c.which_source_to_fire_in_which_shot) flat.insert(flat.end(), s.begin(), s.end());
      return flat;
    }
  }
  throw psv::invalid_argument("Unknown index list.");
}
}  // namespace

int psv_index_list_size(psv_handle h, psv_index_list which, size_t *count) {
  return guarded([&] { *count = index_list(model(h).config(), which).size(); });
}
int psv_get_index_list(psv_handle h, psv_index_list which, int *out, size_t count) {
  return guarded([&] {
    const std::vector<int> v = index_list(model(h).config(), which);
    if (count != v.size()) throw psv::invalid_argument("Index list size mismatch.");
    if (count) std::memcpy(out, v.data(), count * sizeof(int));
  });
}

int psv_write_receivers(psv_handle h) {
  return guarded([&] {
    if (psv::DeviceGroup *g = group(h)) g->write_receivers();
    else model(h).write_receivers();
  });
}
int psv_load_receivers(psv_handle h, int verbose) {
  return guarded([&] {
    if (psv::DeviceGroup *g = group(h)) g->load_receivers(verbose != 0);
    else model(h).load_receivers(verbose != 0);
  });
}
int psv_write_sources(psv_handle h) {
  return guarded([&] { model(h).write_sources(); });  // host tables only
}
int psv_write_kernels(psv_handle h) {
  return guarded([&] {
    if (psv::DeviceGroup *g = group(h)) g->write_kernels();
    else model(h).write_kernels();
  });
}
int psv_load_model(psv_handle h, const char *de_path, const char *vp_path, const char *vs_path, int verbose) {
  return guarded([&] {
    if (psv::DeviceGroup *g = group(h)) g->load_model(de_path, vp_path, vs_path, verbose != 0);
    else model(h).load_model(de_path, vp_path, vs_path, verbose != 0);
  });
}
int psv_load_vector(psv_handle h, const char *path, int verbose, double *out, size_t n) {
  return guarded([&] { model(h).load_vector(path, verbose != 0, out, n); });
}

int psv_run_gradient(psv_handle h, int release_each_shot, double *misfit_out) {
  return guarded([&] {
    const double m = group(h) ? h->g->run_gradient(release_each_shot != 0) : model(h).run_gradient(release_each_shot != 0);
    if (misfit_out) *misfit_out = m;
  });
}
int psv_descent_step(psv_handle h, double step_length) {
  return guarded([&] {
    if (psv::DeviceGroup *g = group(h)) g->descent_step(step_length);
    else model(h).descent_step(step_length);
  });
}
int psv_gradient_descent(psv_handle h, int iterations, double step_length, int release_each_shot,
                         double *misfits_out) {
  return guarded([&] {
    if (psv::DeviceGroup *g = group(h)) g->gradient_descent(iterations, step_length, release_each_shot != 0, misfits_out);
    else model(h).gradient_descent(iterations, step_length, release_each_shot != 0, misfits_out);
  });
}

int psv_get_gradient_vector_smoothed(psv_handle h, double sigma, double *out, size_t n) {
  return guarded([&] {
    if (psv::DeviceGroup *g = group(h)) g->get_gradient_vector_smoothed(out, n, sigma);
    else model(h).get_gradient_vector_smoothed(out, n, sigma);
  });
}
int psv_gaussian_weights(double sigma, double *weights_out, size_t capacity, int *radius_out) {
  return guarded([&] {
    if (!(sigma > 0.0)) throw psv::invalid_argument("Smoothing sigma must be > 0.");
    int radius = 0;
    const std::vector<double> w = psv::gaussian_weights(sigma, &radius);
    if (radius_out) *radius_out = radius;
    if (weights_out) {
      if (capacity < w.size()) throw psv::length_error("psv_gaussian_weights: capacity too small.");
      std::copy(w.begin(), w.end(), weights_out);
    }
  });
}
int psv_lbfgs(psv_handle h, int max_iterations, int history, double gtol, double smoothing_sigma,
              int release_each_shot, double *misfits_out, int *iterations_done, int *evaluations_done) {
  return guarded([&] {
    psv::LbfgsOptions opt;
    opt.max_iterations = max_iterations;
    opt.history = history;
    opt.gtol = gtol;
    const psv::LbfgsResult r = group(h) ? h->g->lbfgs(opt, smoothing_sigma, release_each_shot != 0, misfits_out)
                                        : model(h).lbfgs(opt, smoothing_sigma, release_each_shot != 0, misfits_out);
    if (iterations_done) *iterations_done = r.iterations;
    if (evaluations_done) *evaluations_done = r.evaluations;
  });
}

namespace {
std::vector<psv::Model *> batch_models(const psv_handle *handles, int n) {
  if (!handles || n < 1) throw psv::invalid_argument("Batched simulate needs at least one model.");
  std::vector<psv::Model *> v(n);
  for (int i = 0; i < n; ++i) {
    single_device_only(handles[i], "A batched simulate call");
    v[i] = &model(handles[i]);
  }
  return v;
}
}  // namespace

int psv_forward_batch(const psv_handle *handles, int n, int i_shot, int store_fields) {
  return guarded([&] {
    const std::vector<psv::Model *> v = batch_models(handles, n);
    psv::Model::forward_batch(v.data(), n, i_shot, store_fields != 0);
  });
}

int psv_adjoint_batch(const psv_handle *handles, int n, int i_shot) {
  return guarded([&] {
    const std::vector<psv::Model *> v = batch_models(handles, n);
    psv::Model::adjoint_batch(v.data(), n, i_shot);
  });
}

int psv_create_slab_from_ini(const char *ini_path, int rank, int nranks, int halo_rows, psv_handle *out) {
  return guarded([&] {
    if (!ini_path || !out) throw psv::invalid_argument("Null argument.");
    *out = nullptr;
    psv::Config cfg = psv::config_from_ini(ini_path);
    const psv::SlabSpec slab = psv::slab_partition(cfg.nx, rank, nranks, halo_rows);
    const int dev = pick_device();
    *out = new psv_model{new psv::Model(std::move(cfg), dev, &slab)};
  });
}
int psv_slab_info(psv_handle h, long long out[7]) {
  return guarded([&] {
    single_device_only(h, "psv_slab_info");
    model(h).slab_info(out);
  });
}
int psv_march_begin(psv_handle h, int i_shot, int adjoint, int store_fields) {
  return guarded([&] {
    single_device_only(h, "The domain split (psv_march_*)");
    model(h).march_begin(i_shot, adjoint != 0, store_fields != 0);
  });
}
int psv_march_steps(psv_handle h, int n, int *done) {
  return guarded([&] {
    single_device_only(h, "The domain split (psv_march_*)");
    const int d = model(h).march_steps(n);
    if (done) *done = d;
  });
}
int psv_march_end(psv_handle h) {
  return guarded([&] {
    single_device_only(h, "The domain split (psv_march_*)");
    model(h).march_end();
  });
}
int psv_march_slabs(const psv_handle *slabs, int n, int i_shot, int adjoint, int store_fields) {
  return guarded([&] {
    if (!slabs || n < 1) throw psv::invalid_argument("psv_march_slabs needs the slabs of one grid.");
    std::vector<psv::Model *> v(n);
    for (int i = 0; i < n; ++i) {
      single_device_only(slabs[i], "The domain split (psv_march_slabs)");
      v[i] = &model(slabs[i]);
    }
    psv::Model::march_slabs(v.data(), n, i_shot, adjoint != 0, store_fields != 0);
  });
}
int psv_halo_pack(psv_handle h) {
  return guarded([&] {
    single_device_only(h, "The halo exchange (psv_halo_*)");
    model(h).halo_pack();
  });
}
int psv_halo_unpack(psv_handle h) {
  return guarded([&] {
    single_device_only(h, "The halo exchange (psv_halo_*)");
    model(h).halo_unpack();
  });
}
int psv_halo_buffers(psv_handle h, void **send_lo, void **send_hi, void **recv_lo, void **recv_hi, size_t *count) {
  return guarded([&] {
    if (!send_lo || !send_hi || !recv_lo || !recv_hi || !count) throw psv::invalid_argument("Null argument.");
    single_device_only(h, "The halo exchange (psv_halo_*)");
    model(h).halo_buffers(send_lo, send_hi, recv_lo, recv_hi, count);
  });
}
int psv_stream(psv_handle h, void **stream) {
  return guarded([&] {
    if (!stream) throw psv::invalid_argument("Null argument.");
    if (psv::DeviceGroup *g = group(h)) g->synchronize();  // queued runs finish before the primary's stream is handed out
    *stream = (void *)model(h).stream();
  });
}

int psv_release_snapshots(psv_handle h, int i_shot) {
  return guarded([&] {
    if (psv::DeviceGroup *g = group(h)) g->release_snapshots(i_shot);
    else model(h).release_snapshots(i_shot);
  });
}
int psv_device_kernel_buffer(psv_handle h, void **device_ptr, size_t *count) {
  return guarded([&] {
    if (psv::DeviceGroup *g = group(h)) g->kernel_buffer(device_ptr, count);
    else model(h).kernel_buffer(device_ptr, count);
  });
}
int psv_synchronize(psv_handle h) {
  return guarded([&] {
    if (psv::DeviceGroup *g = group(h)) g->synchronize();
    else model(h).synchronize();
  });
}
int psv_timer_start(psv_handle h) {
  return guarded([&] {
    if (psv::DeviceGroup *g = group(h)) g->timer_start();
    else model(h).timer_start();
  });
}
int psv_timer_stop(psv_handle h, double *milliseconds) {
  return guarded([&] { *milliseconds = group(h) ? h->g->timer_stop() : model(h).timer_stop(); });
}
int psv_host_alloc(size_t bytes, void **host_ptr) {
  return guarded([&] {
    if (!host_ptr) throw psv::invalid_argument("Null argument.");
    *host_ptr = nullptr;
    if (bytes == 0) return;
    const cudaError_t e = cudaHostAlloc(host_ptr, bytes, cudaHostAllocPortable);
    if (e != cudaSuccess) {
      cudaGetLastError();
      *host_ptr = nullptr;
      throw psv::cuda_error(std::string("cudaHostAlloc: ") + cudaGetErrorString(e));
    }
  });
}
int psv_host_free(void *host_ptr) {
  return guarded([&] {
    if (host_ptr) PSV_CUDA(cudaFreeHost(host_ptr));
  });
}
int psv_set_option(psv_handle h, const char *name, long long value) {
  return guarded([&] {
    if (psv::DeviceGroup *g = group(h)) g->set_option(name, value);
    else model(h).set_option(name, value);
  });
}
int psv_get_option(psv_handle h, const char *name, long long *value) {
  return guarded([&] {
    if (psv::DeviceGroup *g = group(h)) g->synchronize();  // a queued run may still be choosing (tuner) what is read here
    *value = model(h).get_option(name);
  });
}
int psv_get_stat(psv_handle h, const char *name, double *value) {
  return guarded([&] { *value = group(h) ? h->g->get_stat(name) : model(h).get_stat(name); });
}

}  // extern "C"

// Other translation units of the library (no_alloc.cu) report their failures through psv_last_error() too.
namespace psv {
void set_last_error(const char *message) { g_error = message ? message : ""; }
}  // namespace psv
