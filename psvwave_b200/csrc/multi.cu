// psv::DeviceGroup: one fdModel object over several GPUs of one box (see multi.hpp).
#include "multi.hpp"

#include <dlfcn.h>
#include <nccl.h>

#include <chrono>
#include <cstdlib>
#include <sstream>

namespace psv {

// ---------------------------------------------------------------------------------------------
// NCCL, bound at first use.  The library is looked up by its SONAME, so a process that already
// holds a libnccl.so.2 (e.g. PyTorch's) shares that one instead of loading a second copy.
// ---------------------------------------------------------------------------------------------
namespace {
struct Nccl {
  decltype(&ncclCommInitAll) CommInitAll = nullptr;
  decltype(&ncclCommDestroy) CommDestroy = nullptr;
  decltype(&ncclAllReduce) AllReduce = nullptr;
  decltype(&ncclGroupStart) GroupStart = nullptr;
  decltype(&ncclGroupEnd) GroupEnd = nullptr;
  decltype(&ncclGetErrorString) GetErrorString = nullptr;
  decltype(&ncclGetVersion) GetVersion = nullptr;
};

const Nccl &nccl() {
  static Nccl api;
  static std::once_flag once;
  static std::string failure;
  std::call_once(once, [] {
    // 1. a libnccl.so.2 the process already holds (PyTorch's); 2. the one PSVWAVE_NCCL_LIBRARY names (the Python
    // package points it at the NCCL wheel PyTorch would load later, so that both end up with the same, newest copy --
    // an older system libnccl loaded first would leave PyTorch's import with missing symbols); 3. the system library.
    void *lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);
    if (!lib)
      if (const char *path = std::getenv("PSVWAVE_NCCL_LIBRARY"))
        if (*path) lib = dlopen(path, RTLD_NOW | RTLD_LOCAL);
    for (const char *name : {"libnccl.so.2", "libnccl.so"}) {
      if (lib) break;
      lib = dlopen(name, RTLD_NOW | RTLD_LOCAL);
    }
    if (!lib) {
      failure = std::string("NCCL is needed to sum the kernels of several GPUs and could not be loaded: ") + dlerror();
      return;
    }
    auto sym = [&](const char *n) {
      void *p = dlsym(lib, n);
      if (!p && failure.empty()) failure = std::string("NCCL symbol missing: ") + n;
      return p;
    };
    api.CommInitAll = reinterpret_cast<decltype(api.CommInitAll)>(sym("ncclCommInitAll"));
    api.CommDestroy = reinterpret_cast<decltype(api.CommDestroy)>(sym("ncclCommDestroy"));
    api.AllReduce = reinterpret_cast<decltype(api.AllReduce)>(sym("ncclAllReduce"));
    api.GroupStart = reinterpret_cast<decltype(api.GroupStart)>(sym("ncclGroupStart"));
    api.GroupEnd = reinterpret_cast<decltype(api.GroupEnd)>(sym("ncclGroupEnd"));
    api.GetErrorString = reinterpret_cast<decltype(api.GetErrorString)>(sym("ncclGetErrorString"));
    api.GetVersion = reinterpret_cast<decltype(api.GetVersion)>(sym("ncclGetVersion"));
  });
  if (!failure.empty()) throw nccl_error(failure);
  return api;
}

void nccl_check(ncclResult_t r, const char *what) {
  if (r == ncclSuccess) return;
  throw nccl_error(std::string(what) + ": " + nccl().GetErrorString(r));
}
#define PSV_NCCL(expr) nccl_check((expr), #expr)

__global__ void add_into_kernel(double *__restrict__ dst, const double *__restrict__ src, size_t n) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    dst[i] += src[i];
}
}  // namespace

// ---------------------------------------------------------------------------------------------
// construction, workers
// ---------------------------------------------------------------------------------------------
DeviceGroup::DeviceGroup(Model *primary, const std::vector<int> &devices) : primary_(primary), devices_(devices) {
  if (devices_.empty() || devices_[0] != primary_->device())
    throw invalid_argument("The first device of the list must be the device the object lives on.");
  int count = 0;
  PSV_CUDA(cudaGetDeviceCount(&count));
  // A device may be listed twice only on request (PSVWAVE_ALLOW_DUPLICATE_DEVICES=1): several replicas on one GPU
  // exercise the whole group machinery on a single-GPU box; NCCL refuses two ranks on one device, so the sum of such
  // a group is a plain device add in replica order instead of the all-reduce.
  const char *dup = std::getenv("PSVWAVE_ALLOW_DUPLICATE_DEVICES");
  for (size_t i = 0; i < devices_.size(); ++i) {
    if (devices_[i] < 0 || devices_[i] >= count) throw invalid_argument("CUDA device index out of range.");
    for (size_t j = 0; j < i; ++j)
      if (devices_[j] == devices_[i]) {
        if (!dup || std::atoi(dup) == 0) throw invalid_argument("A device is listed twice.");
        shared_device_ = true;
      }
  }
  primary_->synchronize();
  // direct loads/stores between the GPUs where the box allows it (NVLink): the copies below and NCCL use it
  for (int a : devices_)
    for (int b : devices_) {
      if (a == b) continue;
      int can = 0;
      PSV_CUDA(cudaDeviceCanAccessPeer(&can, a, b));
      if (!can) continue;
      PSV_CUDA(cudaSetDevice(a));
      const cudaError_t e = cudaDeviceEnablePeerAccess(b, 0);
      if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) PSV_CUDA(e);
      cudaGetLastError();
    }
  for (size_t k = 1; k < devices_.size(); ++k)
    replicas_.emplace_back(new Model(*primary_, devices_[k], /*with_snapshots=*/false));
  for (auto &r : replicas_) r->reset_kernels();  // the primary keeps what the object has accumulated so far
  syn_elsewhere_.assign(primary_->config().n_shots, 0);
  start_workers();
}

DeviceGroup::DeviceGroup(DeviceGroup &o, Model *primary_clone) : primary_(primary_clone), devices_(o.devices_) {
  o.drain();
  for (auto &r : o.replicas_) replicas_.emplace_back(new Model(*r));
  syn_elsewhere_ = o.syn_elsewhere_;
  kernels_spread_ = o.kernels_spread_.load();
  shared_device_ = o.shared_device_;
  last_run_ = o.last_run_;
  start_workers();
}

void DeviceGroup::start_workers() {
  for (int k = 0; k < size(); ++k) workers_.emplace_back(new Worker);
  for (int k = 0; k < size(); ++k) {
    Worker *w = workers_[k].get();
    const int device = devices_[k];
    w->thread = std::thread([w, device] {
      cudaSetDevice(device);
      for (;;) {
        std::function<void()> task;
        {
          std::unique_lock<std::mutex> lock(w->mutex);
          w->wake.wait(lock, [&] { return w->stop || !w->queue.empty(); });
          if (w->queue.empty()) return;  // stop requested and nothing left
          task = std::move(w->queue.front());
          w->queue.pop_front();
          w->busy = true;
        }
        std::exception_ptr err;
        try {
          task();
        } catch (...) {
          err = std::current_exception();
        }
        {
          std::lock_guard<std::mutex> lock(w->mutex);
          if (err) {
            if (!w->error) w->error = err;
            w->queue.clear();  // what was queued behind a failed call depended on it
          }
          w->busy = false;
        }
        w->idle.notify_all();
      }
    });
  }
}

DeviceGroup::~DeviceGroup() {
  for (auto &w : workers_) {
    {
      std::lock_guard<std::mutex> lock(w->mutex);
      w->stop = true;
    }
    w->wake.notify_all();
  }
  for (auto &w : workers_)
    if (w->thread.joinable()) w->thread.join();
  if (!comms_.empty()) {
    try {
      for (void *c : comms_)
        if (c) nccl().CommDestroy(static_cast<ncclComm_t>(c));
    } catch (...) {
    }
  }
}

void DeviceGroup::post(int k, std::function<void()> task) {
  Worker &w = *workers_[k];
  {
    std::lock_guard<std::mutex> lock(w.mutex);
    if (w.error) return;  // reported by the next call that waits
    w.queue.push_back(std::move(task));
  }
  w.wake.notify_one();
}

void DeviceGroup::drain() {
  std::exception_ptr first;
  for (auto &wp : workers_) {
    Worker &w = *wp;
    std::unique_lock<std::mutex> lock(w.mutex);
    w.idle.wait(lock, [&] { return w.queue.empty() && !w.busy; });
    if (w.error && !first) first = w.error;
    w.error = nullptr;
  }
  if (first) std::rethrow_exception(first);
}

void DeviceGroup::on_all(const std::function<void(Model &)> &f) {
  drain();
  for (int k = 0; k < size(); ++k) {
    Model *m = &rep(k);
    post(k, [m, f] { f(*m); });
  }
  drain();
}

void DeviceGroup::check_shot(int i_shot) const {
  if (i_shot < 0 || i_shot >= primary_->config().n_shots) throw invalid_argument("Shot index out of range.");
}

// ---------------------------------------------------------------------------------------------
// the hot path: queued on the GPU that owns the shot
// ---------------------------------------------------------------------------------------------
void DeviceGroup::forward_simulate(int i_shot, bool store_fields, bool verbose, bool output_wavefields) {
  check_shot(i_shot);
  const int k = owner(i_shot);
  Model *m = &rep(k);
  last_run_ = k;
  // The book-keeping flags are raised by the task itself, after the run succeeded: a run that throws (or one that
  // post() drops behind an earlier failure) must not make gather_synthetics() / reduce_kernels() pull rows a replica
  // never computed over the primary's valid ones.  The caller reads the flags only after drain(), which takes the
  // worker's mutex after the task finished.
  post(k, [=] {
    m->forward_simulate(i_shot, store_fields, verbose, output_wavefields);
    if (k != 0) syn_elsewhere_[i_shot] = 1;  // one writer per shot: its owner's worker
  });
}

void DeviceGroup::adjoint_simulate(int i_shot, bool verbose) {
  check_shot(i_shot);
  const int k = owner(i_shot);
  Model *m = &rep(k);
  last_run_ = k;
  post(k, [=] {
    m->adjoint_simulate(i_shot, verbose);
    if (k != 0) kernels_spread_ = true;
  });
}

void DeviceGroup::gather_synthetics() {
  drain();
  for (int s = 0; s < (int)syn_elsewhere_.size(); ++s)
    if (syn_elsewhere_[s]) {
      primary_->pull_shot_synthetics(rep(owner(s)), s);
      syn_elsewhere_[s] = 0;
    }
}

double DeviceGroup::calculate_l2_misfit() {
  gather_synthetics();
  return primary_->calculate_l2_misfit();  // one device, the reference's summation domain (src/fdModel.cpp:1128-1145)
}

double DeviceGroup::calculate_l2_misfit_shot(int i_shot) {
  check_shot(i_shot);
  drain();
  return rep(owner(i_shot)).calculate_l2_misfit_shot(i_shot);
}

void DeviceGroup::calculate_l2_adjoint_sources() {
  // every replica holds the observed data of all shots and the synthetics of its own: syn - obs is local
  gather_synthetics();  // the primary's copy serves get_array(PSV_A_STF_*)
  on_all([](Model &m) { m.calculate_l2_adjoint_sources(); });
}

void DeviceGroup::reset_kernels() {
  on_all([](Model &m) { m.reset_kernels(); });
  kernels_spread_ = false;
}

void DeviceGroup::ensure_comms() {
  if (!comms_.empty()) return;
  std::vector<ncclComm_t> c(size());
  PSV_NCCL(nccl().CommInitAll(c.data(), size(), devices_.data()));
  comms_.assign(c.begin(), c.end());
}

// The one collective of the path: sum of the per-GPU lambda/mu/rho accumulations (src/fdModel.cpp:832-858 adds the
// shots into one array; the sum is associative up to rounding, O(1e-16) relative, fixed shot->GPU map => deterministic).
void DeviceGroup::reduce_kernels() {
  drain();
  if (!kernels_spread_) return;
  const auto t0 = std::chrono::steady_clock::now();
  if (shared_device_) {  // test configuration (see the constructor): replicas share a GPU, NCCL cannot be used
    void *dst = nullptr;
    size_t count = 0;
    primary_->kernel_buffer(&dst, &count);
    for (int k = 1; k < size(); ++k) {
      void *src = nullptr;
      rep(k).kernel_buffer(&src, &count);
      PSV_CUDA(cudaSetDevice(devices_[0]));
      add_into_kernel<<<1184, 256, 0, primary_->stream()>>>(static_cast<double *>(dst), static_cast<const double *>(src), count);
      PSV_CUDA(cudaGetLastError());
    }
    primary_->synchronize();
    for (int k = 1; k < size(); ++k) rep(k).reset_kernels();
    kernels_spread_ = false;
    ++stat_local_sums_;
    return;
  }
  ensure_comms();
  PSV_NCCL(nccl().GroupStart());
  for (int k = 0; k < size(); ++k) {
    void *buf = nullptr;
    size_t count = 0;
    rep(k).kernel_buffer(&buf, &count);
    PSV_NCCL(nccl().AllReduce(buf, buf, count, ncclDouble, ncclSum, static_cast<ncclComm_t>(comms_[k]), rep(k).stream()));
  }
  PSV_NCCL(nccl().GroupEnd());
  for (int k = 0; k < size(); ++k) rep(k).synchronize();
  for (int k = 1; k < size(); ++k) rep(k).reset_kernels();  // the total lives on the primary; the sum stays the total
  kernels_spread_ = false;
  ++stat_allreduces_;
  stat_allreduce_ms_ = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
}

void DeviceGroup::map_kernels_to_velocity() {
  reduce_kernels();
  primary_->map_kernels_to_velocity();
}

void DeviceGroup::update_from_velocity() {
  on_all([](Model &m) { m.update_from_velocity(); });
}

// fdModel::run_model (src/fdModel.cpp:1291-1308)
void DeviceGroup::run_model(bool verbose, bool simulate_adjoint) {
  const int n_shots = primary_->config().n_shots;
  for (int s = 0; s < n_shots; ++s) forward_simulate(s, true, verbose, false);
  calculate_l2_misfit();
  if (simulate_adjoint) {
    calculate_l2_adjoint_sources();
    reset_kernels();
    for (int s = 0; s < n_shots; ++s) adjoint_simulate(s, verbose);
    map_kernels_to_velocity();
  }
}

// Model::run_gradient with the shots of every GPU running back to back on that GPU.
double DeviceGroup::run_gradient(bool release_each_shot) {
  const int n_shots = primary_->config().n_shots;
  reset_kernels();
  auto per_shot = std::make_shared<std::vector<double>>(n_shots, 0.0);
  for (int s = 0; s < n_shots; ++s) {
    const int k = owner(s);
    Model *m = &rep(k);
    last_run_ = k;
    post(k, [=] {
      m->forward_simulate(s, true, false, false);
      if (k != 0) syn_elsewhere_[s] = 1;
      (*per_shot)[s] = m->calculate_l2_misfit_shot(s);
      m->calculate_l2_adjoint_sources();
      m->adjoint_simulate(s, false);
      if (k != 0) kernels_spread_ = true;
      if (release_each_shot) m->release_snapshots(s);
    });
  }
  drain();
  double total = 0.0;
  for (double v : *per_shot) total += v;  // shot-major like the reference
  map_kernels_to_velocity();
  primary_->set_misfit(total);
  return total;
}

void DeviceGroup::broadcast_medium() {
  drain();
  for (int k = 1; k < size(); ++k) {
    Model *m = &rep(k), *p = primary_;
    post(k, [m, p] { m->pull_medium(*p); });
  }
  drain();
}

void DeviceGroup::descent_step(double step) {
  reduce_kernels();
  primary_->descent_step(step);
  broadcast_medium();
}

void DeviceGroup::gradient_descent(int iterations, double step, bool release_each_shot, double *misfits) {
  if (iterations < 0) throw invalid_argument("iterations must be >= 0");
  reduce_kernels();
  primary_->descent_load();
  for (int i = 0; i < iterations; ++i) {
    primary_->descent_apply(step);
    broadcast_medium();
    const double m = run_gradient(release_each_shot);
    if (misfits) misfits[i] = m;
  }
}

LbfgsResult DeviceGroup::lbfgs(const LbfgsOptions &opt, double sigma, bool release_each_shot, double *misfits) {
  if (sigma < 0.0) throw invalid_argument("Smoothing sigma must be >= 0.");
  reduce_kernels();
  const size_t n = (size_t)primary_->config().free_parameters;
  PSV_CUDA(cudaSetDevice(devices_[0]));
  DeviceBuffer x(n * sizeof(double));
  primary_->vector_from_fields(x.as<>());
  const LbfgsResult r = lbfgs_minimize(
      x.as<>(), n,
      [&](const double *m, double *g) {
        primary_->fields_from_vector(m);
        broadcast_medium();
        const double f = run_gradient(release_each_shot);  // all GPUs, one all-reduce
        primary_->gradient_to_vector(g, sigma);
        return f;
      },
      opt, misfits, primary_->stream());
  primary_->fields_from_vector(x.as<>());
  broadcast_medium();
  primary_->set_misfit(r.misfit);
  return r;
}

void DeviceGroup::get_gradient_vector_smoothed(double *out, size_t n, double sigma) {
  drain();
  primary_->get_gradient_vector_smoothed(out, n, sigma);
}

// ---------------------------------------------------------------------------------------------
// vectors, arrays, files
// ---------------------------------------------------------------------------------------------
void DeviceGroup::get_model_vector(double *out, size_t n) {
  drain();
  primary_->get_model_vector(out, n);
}

void DeviceGroup::set_model_vector(const double *in, size_t n) {
  drain();
  primary_->set_model_vector(in, n);  // validates; the replicas take the medium device to device
  broadcast_medium();
}

void DeviceGroup::get_gradient_vector(double *out, size_t n) {
  drain();
  primary_->get_gradient_vector(out, n);  // the vp/vs/rho kernels: written by map_kernels_to_velocity only
}

void DeviceGroup::get_array(psv_array which, double *out, size_t count) {
  if (count != primary_->array_size(which)) throw invalid_argument("Array size mismatch.");
  switch (which) {
    case PSV_VX:
    case PSV_VZ:
    case PSV_TXX:
    case PSV_TZZ:
    case PSV_TXZ:  // the state the most recent simulate call left behind
      drain();
      return rep(last_run_).get_array(which, out, count);
    case PSV_LAMBDA_KERNEL:
    case PSV_MU_KERNEL:
    case PSV_DENSITY_L_KERNEL:
      reduce_kernels();
      return primary_->get_array(which, out, count);
    case PSV_RTF_UX:
    case PSV_RTF_UZ:
    case PSV_A_STF_UX:
    case PSV_A_STF_UZ:
      gather_synthetics();
      return primary_->get_array(which, out, count);
    case PSV_ACCU_VX:
    case PSV_ACCU_VZ:
    case PSV_ACCU_TXX:
    case PSV_ACCU_TZZ:
    case PSV_ACCU_TXZ: {
      drain();
      const Config &c = primary_->config();
      const size_t per_shot = (size_t)c.snapshots * c.nx * c.nz;
      for (int s = 0; s < c.n_shots; ++s)
        rep(owner(s)).get_snapshots_shot(which - PSV_ACCU_VX, s, out + (size_t)s * per_shot);
      return;
    }
    default:
      drain();
      return primary_->get_array(which, out, count);
  }
}

void DeviceGroup::set_array(psv_array which, const double *in, size_t count) {
  if (count != primary_->array_size(which)) throw invalid_argument("Array size mismatch.");
  switch (which) {
    case PSV_LAMBDA_KERNEL:
    case PSV_MU_KERNEL:
    case PSV_DENSITY_L_KERNEL:
      reduce_kernels();  // the total of all three sits on the primary, the other replicas hold zeros
      return primary_->set_array(which, in, count);
    case PSV_RTF_UX:
    case PSV_RTF_UZ:
      gather_synthetics();
    // fall through
    default:
      drain();
      primary_->set_array(which, in, count);  // validates before anything is queued
      for (int k = 1; k < size(); ++k) {
        Model *m = &rep(k);
        post(k, [=] { m->set_array(which, in, count); });
      }
      drain();  // `in` is the caller's buffer
  }
}

void DeviceGroup::set_parameter_fields(const double *vp, const double *vs, const double *rho, size_t count) {
  drain();
  primary_->set_parameter_fields(vp, vs, rho, count);
  broadcast_medium();
}

void DeviceGroup::write_receivers() {
  gather_synthetics();
  primary_->write_receivers();
}

void DeviceGroup::load_receivers(bool verbose) {
  drain();
  primary_->load_receivers(verbose);
  for (int k = 1; k < size(); ++k) {
    Model *m = &rep(k), *p = primary_;
    post(k, [m, p] { m->pull_observed(*p); });
  }
  drain();
}

void DeviceGroup::write_kernels() {
  reduce_kernels();
  primary_->write_kernels();
}

void DeviceGroup::load_model(const std::string &de, const std::string &vp, const std::string &vs, bool verbose) {
  drain();
  primary_->load_model(de, vp, vs, verbose);
  broadcast_medium();
}

void DeviceGroup::release_snapshots(int i_shot) {
  drain();
  if (i_shot < 0) {
    for (int k = 0; k < size(); ++k) rep(k).release_snapshots(-1);
  } else {
    check_shot(i_shot);
    rep(owner(i_shot)).release_snapshots(i_shot);
  }
}

void DeviceGroup::kernel_buffer(void **ptr, size_t *count) {
  reduce_kernels();
  primary_->kernel_buffer(ptr, count);
}

void DeviceGroup::synchronize() {
  drain();
  for (int k = 0; k < size(); ++k) rep(k).synchronize();
}

// A group's stopwatch is the host clock between two points at which every GPU is idle.
void DeviceGroup::timer_start() {
  synchronize();
  timer_wall_ = std::chrono::steady_clock::now();
}

double DeviceGroup::timer_stop() {
  synchronize();
  return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - timer_wall_).count();
}

void DeviceGroup::set_option(const std::string &name, long long value) {
  drain();
  primary_->set_option(name, value);  // validates
  for (int k = 1; k < size(); ++k) rep(k).set_option(name, value);
}

double DeviceGroup::get_stat(const std::string &name) {
  drain();
  if (name == "devices") return size();
  if (name == "allreduces") return (double)stat_allreduces_;
  if (name == "local_sums") return (double)stat_local_sums_;
  if (name == "allreduce_ms") return stat_allreduce_ms_;
  return rep(last_run_).get_stat(name);
}

double DeviceGroup::misfit() {
  drain();
  return primary_->misfit();
}

}  // namespace psv
