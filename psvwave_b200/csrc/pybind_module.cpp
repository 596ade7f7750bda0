// Python extension `__psvWave_cpp` -- same module name, class name, method names, keyword
// arguments, defaults and read-only attributes as the reference binding
// (src/psvWave.cpp:460-658), implemented purely on top of the C ABI in
// include/psvwave_b200.h (no solver code is compiled into this module).
//
// Deviations (all supersets): array getters return correctly shaped C-contiguous copies --
// at the surveyed reference HEAD they raise (src/psvWave.cpp:57-63 passes a 1-D stride list
// for N-D shapes); inputs are accepted in any layout/dtype and converted (the reference
// reads the raw buffer assuming C-contiguous float64, src/psvWave.cpp:286-293);
// `omp_threads_override` is accepted and ignored.
#include <pybind11/numpy.h>
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>

#include <iostream>
#include <stdexcept>
#include <map>
#include <string>
#include <vector>

#include "../../include/psvwave_b200.h"

namespace py = pybind11;
using array_d = py::array_t<double, py::array::c_style | py::array::forcecast>;

namespace {

void check(int status) {
  if (status == PSV_OK) return;
  const std::string msg = psv_last_error();
  switch (status) {
    case PSV_ERR_INVALID_ARGUMENT:
    case PSV_ERR_LENGTH:
      throw py::value_error(msg);  // std::invalid_argument / std::length_error -> ValueError
    default:
      throw std::runtime_error(msg);
  }
}

// Large arrays handed to Python live in page-locked host memory, so the library's device-to-host copy is one DMA at
// link speed instead of a staged copy into freshly mapped pageable pages (the 86 MB gradient vector of the Marmousi-
// scale workload: ~15 ms -> ~2 ms).  Every array owns its buffer (getters return copies, like the reference's,
// src/psvWave.cpp:248-322); a buffer goes back to a small pool when its array is garbage-collected and is reused by the
// next getter of the same size.  If memory cannot be pinned the array falls back to ordinary numpy memory.
class PinnedPool {
 public:
  static PinnedPool &instance() {
    static PinnedPool *p = new PinnedPool;  // never destroyed: arrays may outlive module teardown
    return *p;
  }
  void *acquire(size_t bytes) {
    auto range = free_.equal_range(bytes);
    if (range.first != range.second) {
      void *p = range.first->second;
      free_.erase(range.first);
      cached_ -= bytes;
      return p;
    }
    void *p = nullptr;
    if (psv_host_alloc(bytes, &p) != PSV_OK) return nullptr;
    size_[p] = bytes;
    return p;
  }
  void release(void *p) {
    auto f = size_.find(p);
    if (f == size_.end()) return;
    if (cached_ + f->second <= kMaxCached) {
      free_.emplace(f->second, p);
      cached_ += f->second;
    } else {
      psv_host_free(p);
      size_.erase(f);
    }
  }

 private:
  static constexpr size_t kMaxCached = (size_t)2 << 30;
  std::multimap<size_t, void *> free_;
  std::map<void *, size_t> size_;
  size_t cached_ = 0;
};

py::array_t<double> host_array(const std::vector<py::ssize_t> &shape) {
  size_t n = 1;
  for (py::ssize_t d : shape) n *= (size_t)d;
  if (n * sizeof(double) >= ((size_t)1 << 20) && n * sizeof(double) <= ((size_t)1 << 29)) {  // 1 MiB .. 512 MiB
    if (void *p = PinnedPool::instance().acquire(n * sizeof(double))) {
      py::capsule owner(p, [](void *q) { PinnedPool::instance().release(q); });
      return py::array_t<double>(shape, static_cast<double *>(p), owner);
    }
  }
  return py::array_t<double>(shape);
}

class FdModel {
 public:
  explicit FdModel(const std::string &path) { check(psv_create_from_ini(path.c_str(), &h_)); }
  explicit FdModel(psv_handle h) : h_(h) {}
  FdModel(const FdModel &) = delete;
  FdModel &operator=(const FdModel &) = delete;
  ~FdModel() {
    if (h_) psv_destroy(h_);
  }

  FdModel *copy() const {
    psv_handle c = nullptr;
    check(psv_clone(h_, &c));
    return new FdModel(c);
  }

  long long geti(psv_int_key k) const {
    long long v = 0;
    check(psv_get_int(h_, k, &v));
    return v;
  }
  double getr(psv_real_key k) const {
    double v = 0;
    check(psv_get_real(h_, k, &v));
    return v;
  }

  py::array_t<double> fetch(psv_array which, std::vector<py::ssize_t> shape) const {
    py::array_t<double> out = host_array(shape);
    size_t n = 0;
    check(psv_array_size(h_, which, &n));
    if ((size_t)out.size() != n) throw std::runtime_error("internal shape mismatch");
    check(psv_get_array(h_, which, out.mutable_data(), n));
    return out;
  }
  std::vector<py::ssize_t> grid_shape() const { return {geti(PSV_INT_NX), geti(PSV_INT_NZ)}; }
  std::vector<py::ssize_t> data_shape() const {
    return {geti(PSV_INT_N_SHOTS), geti(PSV_INT_NR), geti(PSV_INT_NT)};
  }
  std::vector<py::ssize_t> accu_shape() const {
    return {geti(PSV_INT_N_SHOTS), geti(PSV_INT_SNAPSHOTS), geti(PSV_INT_NX), geti(PSV_INT_NZ)};
  }

  void forward_simulate(int i_shot, bool store_fields, bool verbose, bool output_wavefields,
                        int /*omp_threads_override*/) {
    py::gil_scoped_release release;
    check(psv_forward_simulate(h_, i_shot, store_fields, verbose, output_wavefields));
  }
  void adjoint_simulate(int i_shot, bool verbose, int /*omp_threads_override*/) {
    py::gil_scoped_release release;
    check(psv_adjoint_simulate(h_, i_shot, verbose));
  }

  py::array_t<double> get_model_vector() const {
    py::array_t<double> out = host_array({(py::ssize_t)geti(PSV_INT_FREE_PARAMETERS)});
    check(psv_get_model_vector(h_, out.mutable_data(), (size_t)out.size()));
    return out;
  }
  void set_model_vector(array_d m) { check(psv_set_model_vector(h_, m.data(), (size_t)m.size())); }
  py::array_t<double> get_gradient_vector() const {
    py::array_t<double> out = host_array({(py::ssize_t)geti(PSV_INT_FREE_PARAMETERS)});
    check(psv_get_gradient_vector(h_, out.mutable_data(), (size_t)out.size()));
    return out;
  }
  py::array_t<double> load_vector(const std::string &path, bool verbose) const {
    py::array_t<double> out((py::ssize_t)(3 * geti(PSV_INT_NX_FREE_PARAMETERS) * geti(PSV_INT_NZ_FREE_PARAMETERS)));
    check(psv_load_vector(h_, path.c_str(), verbose, out.mutable_data(), (size_t)out.size()));
    return out;
  }

  py::tuple get_snapshots() const {
    return py::make_tuple(fetch(PSV_ACCU_VX, accu_shape()), fetch(PSV_ACCU_VZ, accu_shape()),
                          fetch(PSV_ACCU_TXX, accu_shape()), fetch(PSV_ACCU_TZZ, accu_shape()),
                          fetch(PSV_ACCU_TXZ, accu_shape()));
  }

  // src/psvWave.cpp:99-110 (mixed units are reference behaviour)
  py::tuple get_extent(bool include_absorbing_boundary) const {
    const long long np = geti(PSV_INT_NP_BOUNDARY);
    const double dx = getr(PSV_REAL_DX), dz = getr(PSV_REAL_DZ);
    if (include_absorbing_boundary)
      return py::make_tuple(-np, dx * (geti(PSV_INT_NX_INNER) + np), -np, dz * (geti(PSV_INT_NZ_INNER) + np));
    return py::make_tuple(0, dx * geti(PSV_INT_NX_INNER), 0, dz * geti(PSV_INT_NZ_INNER));
  }

  // src/psvWave.cpp:207-246
  py::tuple get_coordinates(bool in_units) const {
    const long long nx = geti(PSV_INT_NX), nz = geti(PSV_INT_NZ), np = geti(PSV_INT_NP_BOUNDARY);
    const double dx = getr(PSV_REAL_DX), dz = getr(PSV_REAL_DZ);
    py::array_t<double> IX(grid_shape()), IZ(grid_shape());
    double *px = IX.mutable_data(), *pz = IZ.mutable_data();
    for (long long ix = 0; ix < nx; ++ix)
      for (long long iz = 0; iz < nz; ++iz) {
        px[ix * nz + iz] = in_units ? (ix - np) * dx : (double)ix;
        pz[ix * nz + iz] = in_units ? (iz - np) * dz : (double)iz;
      }
    return py::make_tuple(IX, IZ);
  }

  py::tuple get_parameter_fields() const {
    return py::make_tuple(fetch(PSV_VP, grid_shape()), fetch(PSV_VS, grid_shape()), fetch(PSV_RHO, grid_shape()));
  }
  void set_parameter_fields(array_d vp, array_d vs, array_d rho) {
    const auto shape = grid_shape();
    auto same = [&](const array_d &a) {
      return a.ndim() == 2 && a.shape(0) == shape[0] && a.shape(1) == shape[1];
    };
    if (!same(vp)) throw py::value_error("The input ndarray _vp does not have the right shape.");
    if (!same(vs)) throw py::value_error("The input ndarray _vs does not have the right shape.");
    if (!same(rho)) throw py::value_error("The input ndarray _rho does not have the right shape.");
    check(psv_set_parameter_fields(h_, vp.data(), vs.data(), rho.data(), (size_t)vp.size()));
  }
  py::tuple get_kernels() const {
    return py::make_tuple(fetch(PSV_VP_KERNEL, grid_shape()), fetch(PSV_VS_KERNEL, grid_shape()),
                          fetch(PSV_DENSITY_V_KERNEL, grid_shape()));
  }
  py::tuple get_synthetic_data() const {
    return py::make_tuple(fetch(PSV_RTF_UX, data_shape()), fetch(PSV_RTF_UZ, data_shape()));
  }
  py::tuple get_observed_data() const {
    return py::make_tuple(fetch(PSV_RTF_UX_TRUE, data_shape()), fetch(PSV_RTF_UZ_TRUE, data_shape()));
  }
  void set_data(psv_array wx, psv_array wz, array_d ux, array_d uz) {
    const auto shape = data_shape();
    auto same = [&](const array_d &a) {
      return a.ndim() == 3 && a.shape(0) == shape[0] && a.shape(1) == shape[1] && a.shape(2) == shape[2];
    };
    if (!same(ux)) throw py::value_error("The input ndarray ux does not have the right shape.");
    if (!same(uz)) throw py::value_error("The input ndarray uz does not have the right shape.");
    check(psv_set_array(h_, wx, ux.data(), (size_t)ux.size()));
    check(psv_set_array(h_, wz, uz.data(), (size_t)uz.size()));
  }
  void set_synthetic_data(array_d ux, array_d uz) { set_data(PSV_RTF_UX, PSV_RTF_UZ, ux, uz); }
  void set_observed_data(array_d ux, array_d uz) { set_data(PSV_RTF_UX_TRUE, PSV_RTF_UZ_TRUE, ux, uz); }

  // src/psvWave.cpp:324-387
  py::tuple stations(psv_index_list wx, psv_index_list wz, bool in_units, bool include_boundary) const {
    size_t n = 0;
    check(psv_index_list_size(h_, wx, &n));
    std::vector<int> ix(n), iz(n);
    check(psv_get_index_list(h_, wx, ix.data(), n));
    check(psv_get_index_list(h_, wz, iz.data(), n));
    const long long np = geti(PSV_INT_NP_BOUNDARY);
    const double dx = getr(PSV_REAL_DX), dz = getr(PSV_REAL_DZ);
    py::array_t<double> x((py::ssize_t)n), z((py::ssize_t)n);
    for (size_t i = 0; i < n; ++i) {
      double vx = ix[i], vz = iz[i];
      if (!include_boundary || in_units) {
        vx -= np;
        vz -= np;
        if (in_units) {
          vx *= dx;
          vz *= dz;
        }
      }
      x.mutable_data()[i] = vx;
      z.mutable_data()[i] = vz;
    }
    return py::make_tuple(x, z);
  }
  py::tuple get_sources(bool in_units, bool include_boundary) const {
    return stations(PSV_IX_SOURCES, PSV_IZ_SOURCES, in_units, include_boundary);
  }
  py::tuple get_receivers(bool in_units, bool include_boundary) const {
    return stations(PSV_IX_RECEIVERS, PSV_IZ_RECEIVERS, in_units, include_boundary);
  }

  std::vector<std::vector<int>> which_source_to_fire_in_which_shot() const {
    size_t no = 0, ns = 0;
    check(psv_index_list_size(h_, PSV_SHOT_OFFSETS, &no));
    check(psv_index_list_size(h_, PSV_SHOT_SOURCES, &ns));
    std::vector<int> off(no), src(ns);
    check(psv_get_index_list(h_, PSV_SHOT_OFFSETS, off.data(), no));
    check(psv_get_index_list(h_, PSV_SHOT_SOURCES, src.data(), ns));
    std::vector<std::vector<int>> out;
    for (size_t s = 0; s + 1 < no; ++s) out.emplace_back(src.begin() + off[s], src.begin() + off[s + 1]);
    return out;
  }

  void calculate_l2_misfit() { check(psv_calculate_l2_misfit(h_, nullptr)); }
  double calculate_l2_misfit_shot(int i_shot) {
    double v = 0;
    check(psv_calculate_l2_misfit_shot(h_, i_shot, &v));
    return v;
  }
  void calculate_l2_adjoint_sources() { check(psv_calculate_l2_adjoint_sources(h_)); }
  void reset_kernels() { check(psv_reset_kernels(h_)); }
  void map_kernels_to_velocity() { check(psv_map_kernels_to_velocity(h_)); }

  // C++-only members of the reference class, bound here as additive methods
  void run_model(bool verbose, bool simulate_adjoint) {
    py::gil_scoped_release release;
    check(psv_run_model(h_, verbose, simulate_adjoint));
  }
  double run_gradient(bool release_each_shot) {
    double m = 0;
    py::gil_scoped_release release;
    check(psv_run_gradient(h_, release_each_shot, &m));
    return m;
  }
  void descent_step(double step) { check(psv_descent_step(h_, step)); }
  py::array_t<double> gradient_descent(int iterations, double step, bool release_each_shot) {
    if (iterations < 0) throw py::value_error("iterations must be >= 0");
    py::array_t<double> misfits((py::ssize_t)iterations);
    {
      py::gil_scoped_release release;
      check(psv_gradient_descent(h_, iterations, step, release_each_shot, misfits.mutable_data()));
    }
    return misfits;
  }
  void write_receivers() { check(psv_write_receivers(h_)); }
  void load_receivers(bool verbose) { check(psv_load_receivers(h_, verbose)); }
  void write_sources() { check(psv_write_sources(h_)); }
  void write_kernels() { check(psv_write_kernels(h_)); }
  void load_model(const std::string &de, const std::string &vp, const std::string &vs, bool verbose) {
    check(psv_load_model(h_, de.c_str(), vp.c_str(), vs.c_str(), verbose));
  }
  void update_from_velocity() { check(psv_update_from_velocity(h_)); }
  void release_snapshots(int i_shot) { check(psv_release_snapshots(h_, i_shot)); }
  py::array_t<double> get_array(int which) const {
    size_t n = 0;
    check(psv_array_size(h_, (psv_array)which, &n));
    py::array_t<double> out((py::ssize_t)n);
    check(psv_get_array(h_, (psv_array)which, out.mutable_data(), n));
    return out;
  }
  py::array_t<double> smoothed_gradient(double sigma) const {
    const py::ssize_t n = geti(PSV_INT_FREE_PARAMETERS);
    py::array_t<double> out = host_array({n});
    check(psv_get_gradient_vector_smoothed(h_, sigma, out.mutable_data(), (size_t)n));
    return out;
  }
  py::tuple lbfgs(int iterations, int history, double gtol, double smoothing, bool release_each_shot) {
    std::vector<double> misfits((size_t)std::max(iterations, 0) + 1, 0.0);
    int done = 0, evals = 0;
    {
      py::gil_scoped_release release;
      check(psv_lbfgs(h_, iterations, history, gtol, smoothing, release_each_shot, misfits.data(), &done, &evals));
    }
    py::array_t<double> out(done + 1);
    std::copy(misfits.begin(), misfits.begin() + done + 1, out.mutable_data());
    return py::make_tuple(out, evals);
  }
  void set_option(const std::string &name, long long v) { check(psv_set_option(h_, name.c_str(), v)); }
  long long get_option(const std::string &name) const {
    long long v = 0;
    check(psv_get_option(h_, name.c_str(), &v));
    return v;
  }
  double get_stat(const std::string &name) const {
    double v = 0;
    check(psv_get_stat(h_, name.c_str(), &v));
    return v;
  }
  void timer_start() { check(psv_timer_start(h_)); }
  double timer_stop() {
    double ms = 0;
    check(psv_timer_stop(h_, &ms));
    return ms;
  }
  void synchronize() { check(psv_synchronize(h_)); }
  void set_devices(const std::vector<int> &ids) { check(psv_set_devices(h_, ids.data(), (int)ids.size())); }
  std::vector<int> devices() const {
    int n = 0;
    check(psv_get_devices(h_, nullptr, 0, &n));
    std::vector<int> ids(n);
    check(psv_get_devices(h_, ids.data(), n, &n));
    return ids;
  }
  py::tuple device_kernel_buffer() const {
    void *p = nullptr;
    size_t n = 0;
    check(psv_device_kernel_buffer(h_, &p, &n));
    return py::make_tuple((size_t)(uintptr_t)p, n);
  }

  psv_handle handle() const { return h_; }

  // domain split (psv_create_slab_from_ini ... psv_stream)
  static FdModel *slab(const std::string &path, int rank, int nranks, int halo_rows) {
    psv_handle h = nullptr;
    check(psv_create_slab_from_ini(path.c_str(), rank, nranks, halo_rows, &h));
    return new FdModel(h);
  }
  py::dict slab_info() const {
    long long v[7];
    check(psv_slab_info(h_, v));
    py::dict d;
    d["nx_global"] = v[0];
    d["x_offset"] = v[1];
    d["nx_local"] = v[2];
    d["halo_lo"] = v[3];
    d["halo_hi"] = v[4];
    d["pitch"] = v[5];
    d["is_slab"] = v[6] != 0;
    return d;
  }
  void march_begin(int i_shot, bool adjoint, bool store_fields) {
    check(psv_march_begin(h_, i_shot, adjoint, store_fields));
  }
  int march_steps(int n) {
    int done = 0;
    check(psv_march_steps(h_, n, &done));
    return done;
  }
  void march_end() {
    py::gil_scoped_release release;
    check(psv_march_end(h_));
  }
  void halo_pack() { check(psv_halo_pack(h_)); }
  void halo_unpack() { check(psv_halo_unpack(h_)); }
  py::tuple halo_buffers() const {
    void *p[4] = {nullptr, nullptr, nullptr, nullptr};
    size_t n = 0;
    check(psv_halo_buffers(h_, &p[0], &p[1], &p[2], &p[3], &n));
    return py::make_tuple((size_t)(uintptr_t)p[0], (size_t)(uintptr_t)p[1], (size_t)(uintptr_t)p[2],
                          (size_t)(uintptr_t)p[3], n);
  }
  size_t stream() const {
    void *s = nullptr;
    check(psv_stream(h_, &s));
    return (size_t)(uintptr_t)s;
  }

 private:
  psv_handle h_ = nullptr;
};

std::vector<psv_handle> batch_handles(const std::vector<FdModel *> &models) {
  std::vector<psv_handle> h;
  h.reserve(models.size());
  for (const FdModel *m : models) {
    if (!m) throw py::value_error("None in the list of models.");
    h.push_back(m->handle());
  }
  return h;
}

}  // namespace

PYBIND11_MODULE(__psvWave_cpp, m) {
  m.doc() = "B200-native drop-in for psvWave's __psvWave_cpp extension (C ABI: include/psvwave_b200.h)";
  m.def("set_default_device", [](int d) { check(psv_set_default_device(d)); });
  m.def("device_count", [] {
    int n = 0;
    check(psv_device_count(&n));
    return n;
  });
  m.def("max_variant", [] { return psv_variant_count() - 1; });
  m.def("version", [] { return std::string(psv_version()); });
  // psv_array enumerators under the reference's field names (src/fdModel.h:209-252), for _get_array
  m.def("_array_ids", [] {
    py::dict d;
    const char *names[PSV_ARRAY_COUNT_] = {
        "vx", "vz", "txx", "tzz", "txz", "lm", "la", "mu", "b_vx", "b_vz", "rho", "vp", "vs", "lambda_kernel",
        "mu_kernel", "density_l_kernel", "vp_kernel", "vs_kernel", "density_v_kernel", "taper", "t", "stf", "moment",
        "rtf_ux", "rtf_uz", "rtf_ux_true", "rtf_uz_true", "a_stf_ux", "a_stf_uz", "accu_vx", "accu_vz", "accu_txx",
        "accu_tzz", "accu_txz"};
    for (int i = 0; i < PSV_ARRAY_COUNT_; ++i) d[names[i]] = i;
    return d;
  });
  // additive: all slabs of a split grid inside one process, peer-to-peer halos (psv_march_slabs)
  m.def(
      "march_slabs",
      [](const std::vector<FdModel *> &slabs, int i_shot, bool adjoint, bool store_fields) {
        std::vector<psv_handle> h;
        for (FdModel *s : slabs) h.push_back(s->handle());
        check(psv_march_slabs(h.data(), (int)h.size(), i_shot, adjoint, store_fields));
      },
      py::arg("slabs"), py::arg("i_shot"), py::arg("adjoint"), py::arg("store_fields"));
  // additive: Monte-Carlo ensembles, one launch per time step for all models (psv_forward_batch)
  m.def(
      "forward_batch",
      [](const std::vector<FdModel *> &models, int i_shot, bool store_fields) {
        const std::vector<psv_handle> h = batch_handles(models);
        py::gil_scoped_release release;
        check(psv_forward_batch(h.data(), (int)h.size(), i_shot, store_fields));
      },
      py::arg("models"), py::arg("i_shot"), py::arg("store_fields") = true);
  m.def(
      "adjoint_batch",
      [](const std::vector<FdModel *> &models, int i_shot) {
        const std::vector<psv_handle> h = batch_handles(models);
        py::gil_scoped_release release;
        check(psv_adjoint_batch(h.data(), (int)h.size(), i_shot));
      },
      py::arg("models"), py::arg("i_shot"));

#define RO_INT(name, key) def_property_readonly(name, [](const FdModel &s) { return s.geti(key); })
#define RO_REAL(name, key) def_property_readonly(name, [](const FdModel &s) { return s.getr(key); })

  py::class_<FdModel>(m, "fdModel",
                      "fdModel(configuration_file_path: str)\n"
                      "    Class to simulate P-SV wave physics and its adjoint state (CUDA, sm_100a).\n")
      .def(py::init<const std::string &>())
      .def("copy", &FdModel::copy, py::return_value_policy::take_ownership)
      .def("forward_simulate", &FdModel::forward_simulate, py::arg("i_shot"), py::arg("store_fields") = true,
           py::arg("verbose") = false, py::arg("output_wavefields") = false, py::arg("omp_threads_override") = 0)
      .def("adjoint_simulate", &FdModel::adjoint_simulate, py::arg("i_shot"), py::arg("verbose") = false,
           py::arg("omp_threads_override") = 0)
      .def("get_model_vector", &FdModel::get_model_vector)
      .def("set_model_vector", &FdModel::set_model_vector)
      .def("get_gradient_vector", &FdModel::get_gradient_vector)
      .def("load_vector", &FdModel::load_vector)
      .def("get_snapshots", &FdModel::get_snapshots)
      .def("get_extent", &FdModel::get_extent)
      .def("get_coordinates", &FdModel::get_coordinates)
      .def("get_parameter_fields", &FdModel::get_parameter_fields)
      .def("set_parameter_fields", &FdModel::set_parameter_fields)
      .def("get_kernels", &FdModel::get_kernels)
      .def("get_synthetic_data", &FdModel::get_synthetic_data)
      .def("get_observed_data", &FdModel::get_observed_data)
      .def("set_synthetic_data", &FdModel::set_synthetic_data)
      .def("set_observed_data", &FdModel::set_observed_data)
      .def("get_sources", &FdModel::get_sources, py::arg("in_units") = true,
           py::arg("include_absorbing_boundary_as_index") = true)
      .def("get_receivers", &FdModel::get_receivers, py::arg("in_units") = true,
           py::arg("include_absorbing_boundary_as_index") = true)
      .def("calculate_l2_misfit", &FdModel::calculate_l2_misfit)
      .def("calculate_l2_adjoint_sources", &FdModel::calculate_l2_adjoint_sources)
      .def("reset_kernels", &FdModel::reset_kernels)
      .def("map_kernels_to_velocity", &FdModel::map_kernels_to_velocity)
      .RO_INT("n_sources", PSV_INT_N_SOURCES)
      .RO_REAL("dt", PSV_REAL_DT)
      .RO_REAL("dz", PSV_REAL_DZ)
      .RO_REAL("dx", PSV_REAL_DX)
      .RO_INT("nt", PSV_INT_NT)
      .RO_INT("nz", PSV_INT_NZ)
      .RO_INT("nx", PSV_INT_NX)
      .RO_INT("nx_free_parameters", PSV_INT_NX_FREE_PARAMETERS)
      .RO_INT("nz_free_parameters", PSV_INT_NZ_FREE_PARAMETERS)
      .RO_INT("nx_inner", PSV_INT_NX_INNER)
      .RO_INT("nz_inner", PSV_INT_NZ_INNER)
      .RO_INT("nx_inner_boundary", PSV_INT_NX_INNER_BOUNDARY)
      .RO_INT("nz_inner_boundary", PSV_INT_NZ_INNER_BOUNDARY)
      .RO_INT("np_boundary", PSV_INT_NP_BOUNDARY)
      .RO_INT("free_parameters", PSV_INT_FREE_PARAMETERS)
      .RO_INT("snapshot_interval", PSV_INT_SNAPSHOT_INTERVAL)
      .RO_INT("snapshots", PSV_INT_SNAPSHOTS)
      .RO_INT("n_shots", PSV_INT_N_SHOTS)
      .def_property_readonly("which_source_to_fire_in_which_shot", &FdModel::which_source_to_fire_in_which_shot)
      .RO_REAL("misfit", PSV_REAL_MISFIT)
      // additive (not bound by the reference)
      .RO_INT("nr", PSV_INT_NR)
      .RO_INT("device", PSV_INT_DEVICE)
      .RO_INT("basis_gridpoints_x", PSV_INT_BASIS_GRIDPOINTS_X)
      .RO_INT("basis_gridpoints_z", PSV_INT_BASIS_GRIDPOINTS_Z)
      .def("run_model", &FdModel::run_model, py::arg("verbose") = false, py::arg("simulate_adjoint") = true)
      .def("run_gradient", &FdModel::run_gradient, py::arg("release_each_shot") = true)
      .def("descent_step", &FdModel::descent_step, py::arg("step_length"))
      .def("gradient_descent", &FdModel::gradient_descent, py::arg("iterations"), py::arg("step_length"),
           py::arg("release_each_shot") = true)
      .def("get_smoothed_gradient_vector", &FdModel::smoothed_gradient, py::arg("sigma"),
           "get_gradient_vector() after scipy.ndimage.gaussian_filter on its three (nz_free, nx_free) panels, on the device.")
      .def("lbfgs", &FdModel::lbfgs, py::arg("iterations"), py::arg("history") = 10, py::arg("gtol") = 1e-6,
           py::arg("smoothing") = 0.0, py::arg("release_each_shot") = true,
           "Device-resident L-BFGS on the model vector; returns (misfits[0..iterations done], evaluations).")
      .def("write_receivers", &FdModel::write_receivers)
      .def("load_receivers", &FdModel::load_receivers, py::arg("verbose") = false)
      .def("write_sources", &FdModel::write_sources)
      .def("write_kernels", &FdModel::write_kernels)
      .def("load_model", &FdModel::load_model, py::arg("de_path"), py::arg("vp_path"), py::arg("vs_path"),
           py::arg("verbose") = false)
      .def("update_from_velocity", &FdModel::update_from_velocity)
      .def("calculate_l2_misfit_shot", &FdModel::calculate_l2_misfit_shot)
      .def("release_snapshots", &FdModel::release_snapshots, py::arg("i_shot") = -1)
      .def("_get_array", &FdModel::get_array)
      .def("_set_option", &FdModel::set_option)
      .def("_get_option", &FdModel::get_option)
      .def("_get_stat", &FdModel::get_stat)
      .def("_timer_start", &FdModel::timer_start)
      .def("_timer_stop", &FdModel::timer_stop)
      .def("synchronize", &FdModel::synchronize)
      .def("set_devices", &FdModel::set_devices, py::arg("device_ids"),
           "Spread the object over several GPUs of this box: shots are dealt round-robin, simulate calls are queued "
           "on the shot's GPU and return at once, getters wait; the kernels are summed by one NCCL all-reduce.")
      .def("devices", &FdModel::devices)
      .def("_device_kernel_buffer", &FdModel::device_kernel_buffer)
      .def_static("_slab", &FdModel::slab, py::return_value_policy::take_ownership, py::arg("path"), py::arg("rank"),
                  py::arg("nranks"), py::arg("halo_rows"))
      .def("_slab_info", &FdModel::slab_info)
      .def("_march_begin", &FdModel::march_begin)
      .def("_march_steps", &FdModel::march_steps)
      .def("_march_end", &FdModel::march_end)
      .def("_halo_pack", &FdModel::halo_pack)
      .def("_halo_unpack", &FdModel::halo_unpack)
      .def("_halo_buffers", &FdModel::halo_buffers)
      .def("_stream", &FdModel::stream);
}
