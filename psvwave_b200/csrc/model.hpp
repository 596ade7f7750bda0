// psv::Model -- one fdModel object living on one CUDA device.
// Replaces the state + methods of the reference class fdModel (src/fdModel.h:35-318).
#pragma once

#include <cuda_runtime.h>

#include <map>
#include <memory>
#include <string>
#include <vector>

#include "../../include/psvwave_b200.h"
#include "config.hpp"
#include "optim.hpp"

namespace psv {

struct cuda_error : std::runtime_error {
  using std::runtime_error::runtime_error;
};
struct state_error : std::runtime_error {
  using std::runtime_error::runtime_error;
};
struct io_error : std::runtime_error {
  using std::runtime_error::runtime_error;
};

void cuda_check(cudaError_t e, const char *what, const char *file, int line);
#define PSV_CUDA(expr) ::psv::cuda_check((expr), #expr, __FILE__, __LINE__)

// Owning device allocation (zero-initialised).
class DeviceBuffer {
 public:
  DeviceBuffer() = default;
  explicit DeviceBuffer(size_t bytes) { allocate(bytes); }
  ~DeviceBuffer() { release(); }
  DeviceBuffer(const DeviceBuffer &) = delete;
  DeviceBuffer &operator=(const DeviceBuffer &) = delete;
  DeviceBuffer(DeviceBuffer &&o) noexcept : p_(o.p_), bytes_(o.bytes_) { o.p_ = nullptr; o.bytes_ = 0; }
  DeviceBuffer &operator=(DeviceBuffer &&o) noexcept {
    if (this != &o) {
      release();
      p_ = o.p_;
      bytes_ = o.bytes_;
      o.p_ = nullptr;
      o.bytes_ = 0;
    }
    return *this;
  }
  void allocate(size_t bytes);
  void release();
  template <class T = double>
  T *as() const { return static_cast<T *>(p_); }
  size_t bytes() const { return bytes_; }
  bool empty() const { return p_ == nullptr; }

 private:
  void *p_ = nullptr;
  size_t bytes_ = 0;
};

// A pitched (nx, ld) grid array with its guard bands; data() points at element (0, 0).
struct GridArray {
  DeviceBuffer buf;
  double *view = nullptr;  // set instead of `buf` when the array lives inside a shared allocation
  double *data() const { return view ? view : buf.as<double>() + guard; }
  size_t guard = 0;
};

// x-slab of a larger grid (SURVEY 8e "large single grids"): global rows [x0, x1) including
// halo_lo / halo_hi rows that belong to the neighbouring slabs.
struct SlabSpec {
  int x0 = 0, x1 = 0, halo_lo = 0, halo_hi = 0;
};
// Even split of nx_global rows over nranks slabs, each widened by `halo` rows towards its neighbours.
SlabSpec slab_partition(int nx_global, int rank, int nranks, int halo);

class Model {
 public:
  static constexpr int kVariantCount = 5;
  explicit Model(Config cfg, int device, const SlabSpec *slab = nullptr);
  Model(const Model &other);  // deep copy (src/fdModel.cpp:64-97,301-388)
  // The same deep copy onto another device (psv_set_devices: one replica of the object per GPU); a replica made
  // without the stored snapshots re-creates them with its own forward runs.
  Model(const Model &other, int device, bool with_snapshots);
  ~Model();

  const Config &config() const { return cfg_; }
  int device() const { return device_; }

  // hot path
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
  void descent_load();
  void descent_apply(double step);
  void gradient_descent(int iterations, double step, bool release_each_shot, double *misfits);
  // Device-resident L-BFGS on the model vector (optim.hpp; the reference's notebooks use SciPy's L-BFGS-B on the
  // host): every evaluation is run_gradient; sigma > 0 smooths the gradient panels like the reference's _fwi class.
  LbfgsResult lbfgs(const LbfgsOptions &opt, double sigma, bool release_each_shot, double *misfits);
  void get_gradient_vector_smoothed(double *out, size_t n, double sigma);
  // device-side halves of get_model_vector / set_model_vector / get_gradient_vector (vectors of free_parameters)
  void vector_from_fields(double *dev_m);
  void fields_from_vector(const double *dev_m);
  void gradient_to_vector(double *dev_g, double sigma);
  // Monte-Carlo ensembles (SURVEY 8e "replicas only", config C4): n models that share one
  // experiment (grid, time axis, sources, receivers) advance together, ONE launch per time step
  // for the whole ensemble (blockIdx.z = model).  Each model keeps its own medium, fields,
  // seismograms, snapshots and kernels; results are those of n separate simulate calls.
  // Domain split: the time loop in pieces, so that the host layer can exchange halos between
  // them (psv_march_* / psv_halo_* in the C ABI).  Fused variants 2 and 3 only.
  void march_begin(int i_shot, bool adjoint, bool store_fields);
  int march_steps(int n);  // enqueues up to n time steps, returns how many (0 = the loop is done)
  void march_end();
  void halo_pack();    // owned edge rows of the 5 fields -> send buffers
  void halo_unpack();  // receive buffers -> halo rows of the 5 fields
  void halo_buffers(void **send_lo, void **send_hi, void **recv_lo, void **recv_hi, size_t *count) const;
  void slab_info(long long out[7]) const;
  // All slabs of one grid inside ONE process (one per GPU, or several on one): the whole march of shot i_shot with
  // peer-to-peer halo pulls between the slabs' state arrays, ordered by CUDA events (no staging, no NCCL, no host
  // synchronisation inside the march).  `slabs` in slab order, all of them.
  static void march_slabs(Model *const *slabs, int n, int i_shot, bool adjoint, bool store_fields);
  cudaStream_t stream() const { return stream_; }
  static void forward_batch(Model *const *models, int n, int i_shot, bool store_fields);
  static void adjoint_batch(Model *const *models, int n, int i_shot);

  void get_model_vector(double *out, size_t n);
  void set_model_vector(const double *in, size_t n);
  void get_gradient_vector(double *out, size_t n);

  size_t array_size(psv_array which) const;
  void get_array(psv_array which, double *out, size_t count);
  void set_array(psv_array which, const double *in, size_t count);
  void set_parameter_fields(const double *vp, const double *vs, const double *rho, size_t count);

  // text I/O
  void write_receivers();
  void load_receivers(bool verbose);
  void write_sources();
  void write_kernels();
  void load_model(const std::string &de, const std::string &vp, const std::string &vs, bool verbose);
  void load_vector(const std::string &path, bool verbose, double *out, size_t n);

  void release_snapshots(int i_shot);
  void kernel_buffer(void **ptr, size_t *count) const;
  void synchronize();
  void timer_start();
  double timer_stop();

  void set_option(const std::string &name, long long value);
  long long get_option(const std::string &name) const;
  double get_stat(const std::string &name) const;

  double misfit() const { return misfit_; }

  // ---- support for psv::DeviceGroup (multi.cu): one replica of the object per GPU, shots dealt round-robin ----
  void set_misfit(double v) { misfit_ = v; }
  void pull_shot_synthetics(const Model &src, int i_shot);  // rtf_ux / rtf_uz rows of one shot, device to device
  void pull_medium(const Model &src);                       // vp vs rho lm mu b, device to device
  void pull_observed(const Model &src);                     // observed seismograms of all shots, device to device
  void get_snapshots_shot(int field, int i_shot, double *dst);  // (snapshots, nx, nz) of one shot, zeros if not stored

 private:
  struct ShotPlan {  // per-shot device lists for the point operations
    DeviceBuffer ops, heads;
    int n_heads = 0;
    DeviceBuffer row_ops, op_row_start;  // the same operations bucketed by grid row (fused.cuh)
    DeviceBuffer op_rank;                // position of a row-sorted operation in its (field, point) chain
    int op_rounds = 1;                   // 1 + largest rank
    std::vector<int> op_row_start_host;
  };

  void create_stream_and_events();
  void destroy_stream_and_events();
  void allocate();
  void initialize_host_tables();
  void build_plans();
  void bind() const;  // cudaSetDevice
  GridArray make_grid() const;
  void copy_sync(void *dst, const void *src, size_t bytes, cudaMemcpyKind kind) const;
  void grid_to_host(const double *dev, double *host) const;
  void grid_from_host(double *dev, const double *host) const;
  void zero_dynamic_fields();
  void ensure_snapshots(int i_shot);
  double *snapshot_ptr(int i_shot, int field, int snap) const;
  void check_shot(int i_shot) const;
  double misfit_of_range(size_t first, size_t count);
  void launch_stress(double sdt, int mode, int i_shot, int snap);
  void launch_velocity(double sdt);
  // variant 2: one fused launch per time step (fused.cuh)
  void fused_begin(int i_shot, int point_mode);
  bool receiver_increments() const;
  void finish_receivers(int i_shot, cudaStream_t st);
  void fill_fused_model(int i_shot, void *out) const;
  void fill_fused_model32(int i_shot, void *out) const;
  void ensure_fp32();
  void refresh_medium32();
  const double *field_as_double(int k);
  float *field32_ptr(int p, int k) const { return fields32_.as<float>() + (size_t)(5 * p + k) * field_stride_ + guard_; }
  float *medium32_ptr(int k) const { return medium32_.as<float>() + (size_t)k * field_stride_ + guard_; }
  static void check_batch(Model *const *models, int n, int i_shot);
  void adopt_batch_state(const Model &leader);
  int fused_auto_strip(int n_zwarps, int batch);
  int pipe_warps() const;
  void tune_pipe_geometry(int i_shot, int point_mode);  // first use on a large grid: time a few (warps, strip) candidates, keep the best
  void pipe_geometry(int batch, int *tiles, int *strip, size_t *smem);
  void launch_fused(double sdt, int mode, int point_mode, int i_shot, int snap, int it, int pair_step = 0);
  void fill_fused_common(void *out, double sdt, int point_mode, int i_shot, int snap, int it);
  // resident whole-loop kernel (resident.cuh)
  bool resident_plan(int i_shot, bool adjoint, int *rows, int *ctas, size_t *smem);
  void run_resident(int i_shot, bool adjoint, bool store, int rows, int ctas, size_t smem);
  void resident_check();
  void pair_geometry(int batch, int *tiles, int *tile_pairs, int *strip, size_t *smem);
  int pair_stages() const;
  // Slack (in doubles) before and after the snapshot and Lame-kernel allocations: the bulk copies of
  // the correlation launches read whole row segments, which start 2 pairs before a row and may run
  // past its end by a tile's tail.
  static constexpr size_t kBulkPad = 1024;
  double *lame() const { return lame_kernels_.as<double>() + kBulkPad; }
  double *field_ptr(int p, int k) const { return fields_.as<double>() + (size_t)(5 * p + k) * field_stride_ + guard_; }
  double *F(int k) const { return field_ptr(cur_, k); }  // current state: vx vz txx tzz txz
  void write_wavefield_files(int it);

  Config cfg_;  // of a slab: nx is the LOCAL row count, source/receiver indices stay global
  bool slab_ = false;
  int x_off_ = 0, nx_global_ = 0, halo_lo_ = 0, halo_hi_ = 0;
  DeviceBuffer halo_send_, halo_recv_;  // [lo | hi][5 fields][halo rows * ld]
  struct March {
    bool active = false, adjoint = false, store = false;
    int i_shot = 0, it = 0;
  } march_;
  void pull_halos(const Model *lower, const Model *upper);  // neighbours' owned edge rows -> my halo rows
  cudaEvent_t slab_done_[2] = {nullptr, nullptr}, slab_pulled_[2] = {nullptr, nullptr};  // by exchange parity
  bool slab_peer_checked_ = false;
  int device_ = 0;
  cudaStream_t stream_ = nullptr;
  cudaEvent_t ev_begin_ = nullptr, ev_end_ = nullptr, ev_t0_ = nullptr, ev_t1_ = nullptr;
  std::vector<cudaEvent_t> prof_events_;

  // geometry of the device layout
  int ld_ = 0;            // row pitch in doubles
  size_t grid_elems_ = 0; // nx * ld
  size_t guard_ = 0;
  int n_snap_alloc_ = 0;  // ceil(nt / snapshot_interval)

  // grid arrays
  DeviceBuffer fields_;                       // dynamic fields [2 buffers][vx vz txx tzz txz], each padded grid
  size_t field_stride_ = 0;                   // doubles between consecutive state arrays
  int cur_ = 0;                               // which buffer holds the current state
  GridArray vp_, vs_, rho_, lm_, mu_, b_;     // medium (lm mu b: views into medium_, one L2 policy window)
  DeviceBuffer medium_;                       // [lm | mu | b], each a padded grid
  DeviceBuffer lame_kernels_;                 // [lambda | mu | density_l], each nx*ld
  GridArray vp_kernel_, vs_kernel_, density_v_kernel_;
  DeviceBuffer taper_table_;                  // np_boundary + 1 values
  std::vector<double> taper_host_;            // the same table on the host

  // signals
  std::vector<double> t_host_, stf_host_, moment_host_;
  DeviceBuffer stf_;                                      // [n_sources][nt]
  DeviceBuffer rtf_ux_, rtf_uz_, obs_ux_, obs_uz_, a_stf_ux_, a_stf_uz_;  // [n_shots][nr][nt]
  DeviceBuffer rec_index_, rec_next_, rec_heads_;
  int n_rec_heads_ = 0;
  DeviceBuffer rec_row_start_, rec_row_z_, rec_row_id_, rec_row_x_;  // receivers bucketed by grid row (fused.cuh)
  // optional FP32 wavefield mode (fused32.cuh): float images of the state and of lm mu b
  bool fp32_ = false;
  int fused32_resident_ = 0;
  DeviceBuffer fields32_, medium32_, fused_model32_;
  GridArray conv_;  // double image of one float grid array (getters)
  DeviceBuffer fused_model_;                             // FusedModel per batch member (device copy of the pointers)
  std::vector<Model *> batch_followers_;                 // non-empty only inside forward_batch / adjoint_batch
  std::vector<ShotPlan> plans_;
  std::vector<DeviceBuffer> snapshots_;  // per shot: [5][n_snap_alloc][nx][ld]
  std::vector<DeviceBuffer> snapshot_pool_;  // released buffers, reused by ensure_snapshots
  DeviceBuffer smooth_tmp_;              // second buffer of the separable Gaussian smoothing
  DeviceBuffer descent_vec_;             // [model vector | gradient vector] of the device-resident descent
  DeviceBuffer scratch_;                 // reduction partials / vectors

  double misfit_ = 0.0;

  // options / statistics
  int variant_ = 3;
  bool exact_ = false;
  bool profile_ = false;
  int strip_ = 16, block_ = 128;
  int fused_strip_ = 0 /* 0 = whole waves */, fused_block_ = 128;
  int fused_resident_blocks_ = 0;
  int pipe_stages_ = 4, pipe_warps_ = 0 /* 0 = auto */, pipe_resident_ = 0;
  int tune_ = -1 /* -1 = auto (large grids), 0 = off, 1 = on */, tuned_batch_ = 0, tuned_warps_ = 0, tuned_strip_ = 0;
  double stat_tune_ms_ = 0;
  int l2_persist_mb_ = 0;  // persisting-L2 carve-out for the medium arrays the step kernel re-reads (0 = off)
  void apply_l2_window();
  bool pdl_ = false;  // programmatic dependent launch of consecutive variant-3 step kernels
  int l2_mode_ = 0;  // L2 eviction hints of the variant-3 bulk copies (fused.cuh)
  // CUDA-graph replay of a whole time loop (option "graph"): a small grid's step kernel is shorter than the gap
  // between two stream launches, so the loop of one (direction, shot, store) is captured once and replayed.
  struct LoopGraph {
    cudaGraphExec_t exec = nullptr;
    int end_cur = 0;
    long long launches = 0, pair_launches = 0;
  };
  int resident_ = -1 /* -1 = auto (grids that fit on chip), 0 = off, 1 = required */, resident_rows_ = 0 /* 0 = auto */;
  DeviceBuffer resident_packets_;  // velocity rows in flight between strips (resident.cuh) + the abort word
  unsigned long long resident_flag_base_ = 0;
  unsigned *resident_abort_word_ = nullptr;
  std::map<int, int> resident_capacity_;  // rows per CTA -> co-resident CTAs
  bool resident_timing_ = false;
  // Debug option "poison_guards": every guard band and pad column the kernels may read but must never use or write
  // holds a NaN canary (tests/test_gpu_guards.py; stands in for compute-sanitizer, which is closed on the GPU pool).
  bool poison_ = false;
  void poison_arrays(bool state_only);
  long long count_guard_violations();
  DeviceBuffer resident_timing_buf_;
  double stat_resident_phase_[8] = {}, stat_resident_phase_max_[8] = {};
  DeviceBuffer rec_rank_;
  int rec_rounds_ = 1;
  std::vector<int> rec_row_start_host_;
  int graph_ = -1;  // -1 = auto (small grids only, see graph_wanted), 0 = off, 1 = on
  bool graph_wanted() const;
  std::map<std::vector<unsigned long long>, LoopGraph> graphs_;
  std::map<std::vector<unsigned long long>, int> graph_seen_;  // auto mode: a loop is captured when it comes back
  std::vector<unsigned long long> graph_key(int direction, int i_shot, bool store) const;
  void graph_capture_begin();
  void graph_capture_abort();
  const LoopGraph &graph_capture_end(const std::vector<unsigned long long> &key);
  void graph_replay(const LoopGraph &g);
  void drop_graphs();
  int pair_warps_ = 6, pair_stages_ = 0 /* 0 = as many as fit */, pair_strip_ = 0;
  long long stat_pair_launches_ = 0;
  int stat_resident_ = 0;  // CTAs of the resident loop kernel in the last simulate call (0: per-step launches)
  double stat_pair_ms_ = 0;
  double stat_elapsed_ms_ = 0, stat_stress_ms_ = 0, stat_velocity_ms_ = 0;
  long long stat_launches_ = 0;
};

}  // namespace psv
