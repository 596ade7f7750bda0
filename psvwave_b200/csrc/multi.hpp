// psv::DeviceGroup -- ONE fdModel object driving several GPUs of one box (psv_set_devices).
//
// The reference's user loop (docs-source/source/typical-workflow.rst:5-13,42-59; fdModel::run_model,
// src/fdModel.cpp:1291-1308) is   for shot: forward_simulate(shot) ... for shot: adjoint_simulate(shot)
// on one object.  Shots are independent (the only coupling is the additive lambda/mu/rho accumulation,
// src/fdModel.cpp:832-858, and the misfit sum, :1128-1145), so the group keeps one replica of the object per GPU,
// deals the shots round-robin (shot s -> device s mod n, a fixed map), and lets every simulate call return as soon as
// it is queued on the worker thread of the shot's GPU; calls that read results wait.  An unchanged psvWave script
// fills all GPUs.
//   * media / observed data / options: written to every replica;
//   * synthetics: gathered onto the primary replica (device-to-device over NVLink) before anything reads them, so
//     the misfit is summed on one device in the reference's order;
//   * Lame kernels: every replica accumulates the shots it ran; before anything reads them ONE
//     ncclAllReduce(sum, double) over the contiguous [lambda|mu|rho] buffer (native NCCL, loaded at first use)
//     leaves the total everywhere, after which the non-primary copies are zeroed again -- so "the sum over the
//     replicas is the object's kernel" holds at all times and further adjoint runs keep accumulating;
//   * snapshots stay on the GPU that produced them.
#pragma once

#include <atomic>
#include <chrono>
#include <condition_variable>
#include <deque>
#include <exception>
#include <functional>
#include <memory>
#include <mutex>
#include <thread>
#include <vector>

#include "model.hpp"

namespace psv {

struct nccl_error : std::runtime_error {
  using std::runtime_error::runtime_error;
};

class DeviceGroup {
 public:
  // `primary` (owned by the caller's handle) becomes replica 0; devices[0] must be its device.
  DeviceGroup(Model *primary, const std::vector<int> &devices);
  // deep copy of a whole group around an already cloned primary (psv_clone)
  DeviceGroup(DeviceGroup &other, Model *primary_clone);
  ~DeviceGroup();
  DeviceGroup(const DeviceGroup &) = delete;
  DeviceGroup &operator=(const DeviceGroup &) = delete;

  int size() const { return (int)devices_.size(); }
  const std::vector<int> &devices() const { return devices_; }
  int owner(int i_shot) const { return i_shot % size(); }

  // the fdModel surface (same meaning as the members of psv::Model)
  void forward_simulate(int i_shot, bool store_fields, bool verbose, bool output_wavefields);
  void adjoint_simulate(int i_shot, bool verbose);
  double calculate_l2_misfit();
  double calculate_l2_misfit_shot(int i_shot);
  void calculate_l2_adjoint_sources();
  void reset_kernels();
  void map_kernels_to_velocity();
  void update_from_velocity();
  void run_model(bool verbose, bool simulate_adjoint);
  double run_gradient(bool release_each_shot);
  void descent_step(double step);
  void gradient_descent(int iterations, double step, bool release_each_shot, double *misfits);
  LbfgsResult lbfgs(const LbfgsOptions &opt, double sigma, bool release_each_shot, double *misfits);
  void get_gradient_vector_smoothed(double *out, size_t n, double sigma);
  void get_model_vector(double *out, size_t n);
  void set_model_vector(const double *in, size_t n);
  void get_gradient_vector(double *out, size_t n);
  void get_array(psv_array which, double *out, size_t count);
  void set_array(psv_array which, const double *in, size_t count);
  void set_parameter_fields(const double *vp, const double *vs, const double *rho, size_t count);
  void write_receivers();
  void load_receivers(bool verbose);
  void write_kernels();
  void load_model(const std::string &de, const std::string &vp, const std::string &vs, bool verbose);
  void release_snapshots(int i_shot);
  void kernel_buffer(void **ptr, size_t *count);
  void synchronize();
  void timer_start();
  double timer_stop();
  void set_option(const std::string &name, long long value);
  double get_stat(const std::string &name);
  double misfit();

  // Wait for every queued simulate call; the first failure of a worker is re-thrown here.
  void drain();
  // Everything onto the primary replica: synthetics gathered, kernels summed (before the group is dissolved).
  void collect() {
    gather_synthetics();
    reduce_kernels();
  }

 private:
  struct Worker {
    std::thread thread;
    std::mutex mutex;
    std::condition_variable wake, idle;
    std::deque<std::function<void()>> queue;
    bool busy = false, stop = false;
    std::exception_ptr error;
  };
  void start_workers();
  void post(int k, std::function<void()> task);
  void on_all(const std::function<void(Model &)> &f);  // drain, run f on every replica in parallel, drain
  void gather_synthetics();                            // drain + pull the rows of shots simulated elsewhere
  void reduce_kernels();                               // drain + the one all-reduce, if any replica accumulated
  void broadcast_medium();
  void ensure_comms();
  Model &rep(int k) { return k == 0 ? *primary_ : *replicas_[k - 1]; }
  void check_shot(int i_shot) const;

  Model *primary_;
  std::vector<int> devices_;
  std::vector<std::unique_ptr<Model>> replicas_;  // replicas 1..n-1
  std::vector<std::unique_ptr<Worker>> workers_;
  std::vector<char> syn_elsewhere_;  // per shot: the newest synthetics are on the owner, not yet on the primary
  std::atomic<bool> kernels_spread_{false};  // a non-primary replica accumulated since the last all-reduce (set by workers)
  int last_run_ = 0;                 // replica of the most recent simulate call (dynamic fields, counters)
  std::vector<void *> comms_;        // ncclComm_t per replica, created at the first all-reduce
  bool shared_device_ = false;       // test configuration: several replicas on one GPU (no NCCL)
  long long stat_allreduces_ = 0, stat_local_sums_ = 0;
  double stat_allreduce_ms_ = 0;
  std::chrono::steady_clock::time_point timer_wall_;
};

}  // namespace psv
