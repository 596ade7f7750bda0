#include "model.hpp"

#include <algorithm>
#include <chrono>
#include <mutex>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iomanip>
#include <iostream>
#include <limits>
#include <type_traits>
#include <sstream>

#include "kernels.cuh"
#include "launchers.hpp"
#include "pointwise.cuh"

namespace psv {

void cuda_check(cudaError_t e, const char *what, const char *file, int line) {
  if (e == cudaSuccess) return;
  std::ostringstream os;
  os << "CUDA error '" << cudaGetErrorString(e) << "' in " << what << " (" << file << ":" << line << ")";
  throw cuda_error(os.str());
}

void DeviceBuffer::allocate(size_t bytes) {
  release();
  if (bytes == 0) bytes = 8;
  PSV_CUDA(cudaMalloc(&p_, bytes));
  bytes_ = bytes;
  PSV_CUDA(cudaMemset(p_, 0, bytes));
  PSV_CUDA(cudaStreamSynchronize(cudaStreamLegacy));  // the model's stream does not order against the legacy stream
}

void DeviceBuffer::release() {
  if (p_) cudaFree(p_);
  p_ = nullptr;
  bytes_ = 0;
}

namespace {
constexpr double kPiTruncated = 3.14159265;  // the reference's PI macro (src/fdModel.cpp:14)

inline int round_up(int v, int m) { return (v + m - 1) / m * m; }
inline unsigned blocks_for(size_t n, int threads) { return (unsigned)((n + threads - 1) / threads); }

std::string zero_pad(int v, int width) {
  std::ostringstream os;
  os << std::setw(width) << std::setfill('0') << v;
  return os.str();
}
}  // namespace

// ---------------------------------------------------------------------------------------------
// construction
// ---------------------------------------------------------------------------------------------
SlabSpec slab_partition(int nx_global, int rank, int nranks, int halo) {
  if (nranks < 1 || rank < 0 || rank >= nranks) throw invalid_argument("Slab rank out of range.");
  if (halo < 4 || halo % 4 != 0)
    throw invalid_argument("The halo must be a positive multiple of 4 rows (one time step consumes 4).");
  const int base = nx_global / nranks, rem = nx_global % nranks;
  const int own0 = rank * base + std::min(rank, rem), own1 = own0 + base + (rank < rem ? 1 : 0);
  if (nranks > 1 && own1 - own0 < halo)
    throw invalid_argument("Too many slabs: every slab must own at least `halo` rows.");
  SlabSpec s;
  s.halo_lo = rank > 0 ? halo : 0;
  s.halo_hi = rank < nranks - 1 ? halo : 0;
  s.x0 = own0 - s.halo_lo;
  s.x1 = own1 + s.halo_hi;
  return s;
}

Model::Model(Config cfg, int device, const SlabSpec *slab) : cfg_(std::move(cfg)), device_(device) {
  nx_global_ = cfg_.nx;
  if (slab) {
    if (slab->x0 < 0 || slab->x1 > cfg_.nx || slab->x1 - slab->x0 < slab->halo_lo + slab->halo_hi + 1)
      throw invalid_argument("Slab rows out of range.");
    slab_ = true;
    x_off_ = slab->x0;
    halo_lo_ = slab->halo_lo;
    halo_hi_ = slab->halo_hi;
    cfg_.nx = slab->x1 - slab->x0;  // everything below sizes by the local row count
  }
  bind();
  // process-wide default of the "resident" option (test suites pin the per-step kernels with PSVWAVE_RESIDENT=0)
  if (const char *e = std::getenv("PSVWAVE_RESIDENT")) resident_ = std::max(-1, std::min(1, std::atoi(e)));
  try {
    create_stream_and_events();
    allocate();
    initialize_host_tables();
    build_plans();
    if (slab_) {
      const size_t bytes = 2 * 5 * (size_t)std::max(halo_lo_, halo_hi_) * ld_ * sizeof(double);
      halo_send_.allocate(bytes);
      halo_recv_.allocate(bytes);
    }
    PSV_CUDA(cudaStreamSynchronize(stream_));
  } catch (...) {  // the destructor does not run for a half-built object
    destroy_stream_and_events();
    throw;
  }
}

void Model::create_stream_and_events() {
  PSV_CUDA(cudaStreamCreateWithFlags(&stream_, cudaStreamNonBlocking));
  PSV_CUDA(cudaEventCreate(&ev_begin_));
  PSV_CUDA(cudaEventCreate(&ev_end_));
  PSV_CUDA(cudaEventCreate(&ev_t0_));
  PSV_CUDA(cudaEventCreate(&ev_t1_));
}

void Model::destroy_stream_and_events() {
  cudaSetDevice(device_);
  for (auto &g : graphs_)
    if (g.second.exec) cudaGraphExecDestroy(g.second.exec);
  graphs_.clear();
  for (cudaEvent_t e : prof_events_) cudaEventDestroy(e);
  prof_events_.clear();
  for (cudaEvent_t *e : {&slab_done_[0], &slab_done_[1], &slab_pulled_[0], &slab_pulled_[1], &ev_begin_, &ev_end_,
                         &ev_t0_, &ev_t1_})
    if (*e) {
      cudaEventDestroy(*e);
      *e = nullptr;
    }
  if (stream_) cudaStreamDestroy(stream_);
  stream_ = nullptr;
}

Model::~Model() { destroy_stream_and_events(); }

void Model::bind() const { PSV_CUDA(cudaSetDevice(device_)); }

GridArray Model::make_grid() const {
  GridArray g;
  g.guard = guard_;
  g.buf.allocate((grid_elems_ + 2 * guard_) * sizeof(double));
  return g;
}

void Model::allocate() {
  const Config &c = cfg_;
  ld_ = round_up(c.nz, 16);
  grid_elems_ = (size_t)c.nx * ld_;
  guard_ = 6 * (size_t)ld_ + 1024;  // fused.cuh reads rows -4..nx+3 and up to 30 pairs past a row end
  n_snap_alloc_ = (c.nt + c.snapshot_interval - 1) / c.snapshot_interval;

  GridArray *grids[] = {&vp_, &vs_, &rho_, &vp_kernel_, &vs_kernel_, &density_v_kernel_};
  for (GridArray *g : grids) *g = make_grid();
  // lm, mu, b -- what the step kernel re-reads every time step -- share one allocation
  medium_.allocate(3 * (grid_elems_ + 2 * guard_) * sizeof(double));
  GridArray *med[] = {&lm_, &mu_, &b_};
  for (int k = 0; k < 3; ++k) {
    med[k]->guard = guard_;
    med[k]->view = medium_.as<double>() + (size_t)k * (grid_elems_ + 2 * guard_) + guard_;
  }
  // the ten state arrays (2 buffers x vx vz txx tzz txz) sit in ONE allocation at a constant stride
  field_stride_ = grid_elems_ + 2 * guard_;
  fields_.allocate(10 * field_stride_ * sizeof(double));
  lame_kernels_.allocate((3 * grid_elems_ + 2 * kBulkPad) * sizeof(double));
  taper_table_.allocate((c.np_boundary + 1) * sizeof(double));
  stf_.allocate((size_t)c.n_sources * c.nt * sizeof(double));
  const size_t rec = (size_t)c.n_shots * c.nr * c.nt * sizeof(double);
  DeviceBuffer *recs[] = {&rtf_ux_, &rtf_uz_, &obs_ux_, &obs_uz_, &a_stf_ux_, &a_stf_uz_};
  for (DeviceBuffer *b : recs) b->allocate(rec);
  snapshots_.resize(c.n_shots);
  plans_.resize(c.n_shots);
}

void Model::initialize_host_tables() {
  const Config &c = cfg_;
  const int nt = c.nt;
  // time axis and Ricker wavelets (src/fdModel.cpp:215-231); the delay of a source is its
  // position inside its shot's list times delay_cycles_per_shot periods.
  t_host_.assign(nt, 0.0);
  stf_host_.assign((size_t)c.n_sources * nt, 0.0);
  for (int shot = 0; shot < c.n_shots; ++shot) {
    const std::vector<int> &list = c.which_source_to_fire_in_which_shot[shot];
    for (int k = 0; k < (int)list.size(); ++k) {
      for (unsigned int it = 0; it < (unsigned int)nt; ++it) {
        t_host_[it] = it * c.dt;
        const double f = 1.0 / c.alpha;
        const double tau = t_host_[it] - 1.4 / f - c.delay_cycles_per_shot * k / f;
        stf_host_[(size_t)list[k] * nt + it] =
            (1 - 2 * std::pow(M_PI * f * tau, 2)) * std::exp(-std::pow(M_PI * f * tau, 2));
      }
    }
  }
  copy_sync(stf_.as<>(), stf_host_.data(), stf_host_.size() * sizeof(double), cudaMemcpyHostToDevice);
  // moment tensors (src/fdModel.cpp:234-246)
  moment_host_.assign((size_t)c.n_sources * 4, 0.0);
  for (int s = 0; s < c.n_sources; ++s) {
    const double ang = c.moment_angles[s];
    moment_host_[s * 4 + 0] = std::cos(ang * kPiTruncated / 180.0) * 1e15;
    moment_host_[s * 4 + 1] = -std::sin(ang * kPiTruncated / 180.0) * 1e15;
    moment_host_[s * 4 + 2] = -std::sin(ang * kPiTruncated / 180.0) * 1e15;
    moment_host_[s * 4 + 3] = -std::cos(ang * kPiTruncated / 180.0) * 1e15;
  }
  // taper table (src/fdModel.cpp:263-298): value for `n` enclosing frames, n = 0..np
  taper_host_.assign(c.np_boundary + 1, 1.0);
  for (int n = 0; n <= c.np_boundary; ++n) {
    const double cnt = (double)n;
    taper_host_[n] = std::exp(-std::pow(c.np_factor * (c.np_boundary - cnt), 2));
  }
  copy_sync(taper_table_.as<>(), taper_host_.data(), taper_host_.size() * sizeof(double), cudaMemcpyHostToDevice);
  // homogeneous medium (src/fdModel.cpp:249-261); pad columns get the same values so that
  // 1/rho stays finite everywhere.
  std::vector<double> fill(grid_elems_);
  const std::pair<GridArray *, double> scal[] = {{&vp_, c.scalar_vp}, {&vs_, c.scalar_vs}, {&rho_, c.scalar_rho}};
  for (const auto &p : scal) {
    std::fill(fill.begin(), fill.end(), p.second);
    copy_sync(p.first->data(), fill.data(), grid_elems_ * sizeof(double), cudaMemcpyHostToDevice);
  }
  update_from_velocity();
}

void Model::build_plans() {
  const Config &c = cfg_;
  // receivers: linear index + chains of receivers sharing a grid point
  std::vector<int> idx(std::max(c.nr, 1), 0), next(std::max(c.nr, 1), -1), heads;
  std::map<int, int> last;
  // rows of a slab: global row - x_off_; points outside the slab take no part (local row -1)
  auto local_row = [&](int gx) {
    const int x = gx - x_off_;
    return x >= 0 && x < c.nx ? x : -1;
  };
  for (int r = 0; r < c.nr; ++r) {
    idx[r] = std::max(local_row(c.ix_receivers[r]), 0) * ld_ + c.iz_receivers[r];  // (variants 0/1: whole grids only)
    auto it = last.find(idx[r]);
    if (it == last.end()) heads.push_back(r);
    else next[it->second] = r;
    last[idx[r]] = r;
  }
  n_rec_heads_ = (int)heads.size();
  if (heads.empty()) heads.push_back(0);
  rec_index_.allocate(idx.size() * sizeof(int));
  rec_next_.allocate(next.size() * sizeof(int));
  rec_heads_.allocate(heads.size() * sizeof(int));
  copy_sync(rec_index_.as<int>(), idx.data(), idx.size() * sizeof(int), cudaMemcpyHostToDevice);
  copy_sync(rec_next_.as<int>(), next.data(), next.size() * sizeof(int), cudaMemcpyHostToDevice);
  copy_sync(rec_heads_.as<int>(), heads.data(), heads.size() * sizeof(int), cudaMemcpyHostToDevice);

  // receivers bucketed by grid row for the fused step kernel: stable sort by ix keeps the
  // receiver order inside a row (adjoint injection at coinciding receivers is order-sensitive)
  {
    std::vector<int> order, row_start(c.nx + 1, 0), rz(std::max(c.nr, 1), 0), rid(std::max(c.nr, 1), 0);
    for (int r = 0; r < c.nr; ++r)
      if (local_row(c.ix_receivers[r]) >= 0) order.push_back(r);
    std::stable_sort(order.begin(), order.end(),
                     [&](int a, int b) { return c.ix_receivers[a] < c.ix_receivers[b]; });
    for (int r : order) ++row_start[local_row(c.ix_receivers[r]) + 1];
    for (int x = 0; x < c.nx; ++x) row_start[x + 1] += row_start[x];
    std::vector<int> rx(std::max(c.nr, 1), 0);
    for (size_t q = 0; q < order.size(); ++q) {
      rz[q] = c.iz_receivers[order[q]];
      rid[q] = order[q];
      rx[q] = local_row(c.ix_receivers[order[q]]);
    }
    rec_row_x_.allocate(rx.size() * sizeof(int));
    copy_sync(rec_row_x_.as<int>(), rx.data(), rx.size() * sizeof(int), cudaMemcpyHostToDevice);
    // position of every sorted receiver in the chain of receivers on its grid point (resident.cuh applies
    // coinciding adjoint sources round by round, in receiver order)
    std::vector<int> rank(std::max(c.nr, 1), 0);
    std::map<std::pair<int, int>, int> seen;
    rec_rounds_ = 1;
    for (size_t q = 0; q < order.size(); ++q) {
      rank[q] = seen[std::make_pair(rx[q], rz[q])]++;
      rec_rounds_ = std::max(rec_rounds_, rank[q] + 1);
    }
    rec_rank_.allocate(rank.size() * sizeof(int));
    copy_sync(rec_rank_.as<int>(), rank.data(), rank.size() * sizeof(int), cudaMemcpyHostToDevice);
    rec_row_start_host_ = row_start;
    rec_row_start_.allocate(row_start.size() * sizeof(int));
    rec_row_z_.allocate(rz.size() * sizeof(int));
    rec_row_id_.allocate(rid.size() * sizeof(int));
    copy_sync(rec_row_start_.as<int>(), row_start.data(), row_start.size() * sizeof(int), cudaMemcpyHostToDevice);
    copy_sync(rec_row_z_.as<int>(), rz.data(), rz.size() * sizeof(int), cudaMemcpyHostToDevice);
    copy_sync(rec_row_id_.as<int>(), rid.data(), rid.size() * sizeof(int), cudaMemcpyHostToDevice);
  }

  // sources: the 12 point updates per source in the order of src/fdModel.cpp:702-750
  const double dx4 = c.dx * c.dx * c.dx * c.dx, dz4 = c.dz * c.dz * c.dz * c.dz;
  for (int shot = 0; shot < c.n_shots; ++shot) {
    std::vector<InjectOp> ops;
    std::vector<int> op_heads;
    std::map<std::pair<int, int>, int> last_op;
    auto push = [&](int field, int gx, int z, int source, double weight, double sign, double d4) {
      const int x = local_row(gx);
      if (x < 0) return;  // target row belongs to another slab
      InjectOp op;
      op.point = x * ld_ + z;
      op.field = field;
      op.source = source;
      op.next = -1;
      op.weight = weight;
      op.sign = sign;
      op.d4 = d4;
      const int id = (int)ops.size();
      const auto key = std::make_pair(field, op.point);
      auto it = last_op.find(key);
      if (it == last_op.end()) op_heads.push_back(id);
      else ops[it->second].next = id;
      last_op[key] = id;
      ops.push_back(op);
    };
    for (int s : c.which_source_to_fire_in_which_shot[shot]) {
      const int x = c.ix_sources[s], z = c.iz_sources[s];
      const double *M = &moment_host_[(size_t)s * 4];
      push(0, x - 1, z, s, M[0], -1, dx4);             // (x,x)
      push(0, x, z, s, M[0], +1, dx4);
      push(1, x, z - 1, s, M[3], -1, dz4);             // (z,z)
      push(1, x, z, s, M[3], +1, dz4);
      push(0, x - 1, z + 1, s, 0.25 * M[1], +1, dx4);  // (x,z)
      push(0, x, z + 1, s, 0.25 * M[1], +1, dx4);
      push(0, x - 1, z - 1, s, 0.25 * M[1], -1, dx4);
      push(0, x, z - 1, s, 0.25 * M[1], -1, dx4);
      push(1, x + 1, z - 1, s, 0.25 * M[2], +1, dz4);  // (z,x)
      push(1, x + 1, z, s, 0.25 * M[2], +1, dz4);
      push(1, x - 1, z - 1, s, 0.25 * M[2], -1, dz4);
      push(1, x - 1, z, s, 0.25 * M[2], -1, dz4);
    }
    ShotPlan &p = plans_[shot];
    p.n_heads = (int)op_heads.size();
    {  // the same operations bucketed by target row (stable: original order kept inside a row)
      std::vector<int> order(ops.size()), row_start(c.nx + 1, 0);
      for (size_t i = 0; i < ops.size(); ++i) order[i] = (int)i;
      auto row_of = [&](int i) { return ops[i].point / ld_; };
      std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return row_of(a) < row_of(b); });
      std::vector<RowOp> rops(std::max<size_t>(ops.size(), 1));
      for (size_t q = 0; q < ops.size(); ++q) {
        const InjectOp &o = ops[order[q]];
        ++row_start[row_of(order[q]) + 1];
        rops[q] = RowOp{o.point % ld_, o.field, o.source, o.point / ld_, o.weight, o.sign, o.d4, 1.0 / o.d4};
      }
      for (int x = 0; x < c.nx; ++x) row_start[x + 1] += row_start[x];
      std::vector<int> rank(rops.size(), 0);
      std::map<std::pair<int, int>, int> seen;  // (field, point) -> operations so far
      p.op_rounds = 1;
      for (size_t q = 0; q < ops.size(); ++q) {
        rank[q] = seen[std::make_pair(ops[order[q]].field, ops[order[q]].point)]++;
        p.op_rounds = std::max(p.op_rounds, rank[q] + 1);
      }
      p.op_rank.allocate(rank.size() * sizeof(int));
      copy_sync(p.op_rank.as<int>(), rank.data(), rank.size() * sizeof(int), cudaMemcpyHostToDevice);
      p.op_row_start_host = row_start;
      p.row_ops.allocate(rops.size() * sizeof(RowOp));
      p.op_row_start.allocate(row_start.size() * sizeof(int));
      copy_sync(p.row_ops.as<RowOp>(), rops.data(), rops.size() * sizeof(RowOp), cudaMemcpyHostToDevice);
      copy_sync(p.op_row_start.as<int>(), row_start.data(), row_start.size() * sizeof(int), cudaMemcpyHostToDevice);
    }
    if (ops.empty()) continue;
    p.ops.allocate(ops.size() * sizeof(InjectOp));
    p.heads.allocate(op_heads.size() * sizeof(int));
    copy_sync(p.ops.as<InjectOp>(), ops.data(), ops.size() * sizeof(InjectOp), cudaMemcpyHostToDevice);
    copy_sync(p.heads.as<int>(), op_heads.data(), op_heads.size() * sizeof(int), cudaMemcpyHostToDevice);
  }
}

// Deep copy: every array, the stored snapshots included (src/fdModel.cpp:301-388).
Model::Model(const Model &o) : Model(o, o.device_, true) {}

Model::Model(const Model &o, int device, bool with_snapshots) : cfg_(o.cfg_), device_(device) {
  bind();
  try {
  create_stream_and_events();
  slab_ = o.slab_;
  x_off_ = o.x_off_;
  nx_global_ = o.nx_global_;
  halo_lo_ = o.halo_lo_;
  halo_hi_ = o.halo_hi_;
  allocate();
  if (slab_) {
    halo_send_.allocate(o.halo_send_.bytes());
    halo_recv_.allocate(o.halo_recv_.bytes());
  }
  t_host_ = o.t_host_;
  stf_host_ = o.stf_host_;
  moment_host_ = o.moment_host_;
  taper_host_ = o.taper_host_;
  misfit_ = o.misfit_;
  variant_ = o.variant_;
  exact_ = o.exact_;
  profile_ = o.profile_;
  strip_ = o.strip_;
  block_ = o.block_;
  fused_strip_ = o.fused_strip_;
  fused_block_ = o.fused_block_;
  pipe_stages_ = o.pipe_stages_;
  pipe_warps_ = o.pipe_warps_;
  l2_mode_ = o.l2_mode_;
  pdl_ = o.pdl_;
  graph_ = o.graph_;
  pair_warps_ = o.pair_warps_;
  resident_ = o.resident_;
  resident_rows_ = o.resident_rows_;
  tune_ = o.tune_;
  if (device_ == o.device_) {  // a measurement of that GPU
    tuned_batch_ = o.tuned_batch_;
    tuned_warps_ = o.tuned_warps_;
    tuned_strip_ = o.tuned_strip_;
  }
  pair_stages_ = o.pair_stages_;
  pair_strip_ = o.pair_strip_;
  l2_persist_mb_ = o.l2_persist_mb_;
  fp32_ = o.fp32_;
  PSV_CUDA(cudaStreamSynchronize(o.stream_));
  auto dcopy = [&](const DeviceBuffer &src, DeviceBuffer &dst) {
    if (src.empty()) return;
    if (dst.bytes() != src.bytes()) dst.allocate(src.bytes());
    copy_sync(dst.as<char>(), src.as<char>(), src.bytes(), cudaMemcpyDefault);  // unified addressing: any device
  };
  const GridArray *gs[] = {&o.vp_, &o.vs_, &o.rho_, &o.vp_kernel_, &o.vs_kernel_, &o.density_v_kernel_};
  GridArray *gd[] = {&vp_, &vs_, &rho_, &vp_kernel_, &vs_kernel_, &density_v_kernel_};
  for (size_t i = 0; i < sizeof(gs) / sizeof(gs[0]); ++i) dcopy(gs[i]->buf, gd[i]->buf);
  dcopy(o.medium_, medium_);
  cur_ = o.cur_;
  dcopy(o.fields_, fields_);
  dcopy(o.fields32_, fields32_);
  dcopy(o.medium32_, medium32_);
  dcopy(o.lame_kernels_, lame_kernels_);
  dcopy(o.taper_table_, taper_table_);
  dcopy(o.stf_, stf_);
  dcopy(o.rtf_ux_, rtf_ux_);
  dcopy(o.rtf_uz_, rtf_uz_);
  dcopy(o.obs_ux_, obs_ux_);
  dcopy(o.obs_uz_, obs_uz_);
  dcopy(o.a_stf_ux_, a_stf_ux_);
  dcopy(o.a_stf_uz_, a_stf_uz_);
  if (with_snapshots)
    for (int s = 0; s < cfg_.n_shots; ++s) dcopy(o.snapshots_[s], snapshots_[s]);
  build_plans();
  if (l2_persist_mb_ > 0) apply_l2_window();  // a stream attribute: not inherited from the source's stream
  PSV_CUDA(cudaDeviceSynchronize());
  } catch (...) {  // the destructor does not run for a half-built object
    destroy_stream_and_events();
    throw;
  }
}

// ---- device-to-device transfers between the replicas of one object (psv::DeviceGroup, multi.cu) ----
void Model::pull_shot_synthetics(const Model &src, int i_shot) {
  check_shot(i_shot);
  bind();
  const size_t per_shot = (size_t)cfg_.nr * cfg_.nt, off = (size_t)i_shot * per_shot;
  if (per_shot == 0) return;
  PSV_CUDA(cudaMemcpyAsync(rtf_ux_.as<>() + off, src.rtf_ux_.as<>() + off, per_shot * sizeof(double), cudaMemcpyDefault, stream_));
  PSV_CUDA(cudaMemcpyAsync(rtf_uz_.as<>() + off, src.rtf_uz_.as<>() + off, per_shot * sizeof(double), cudaMemcpyDefault, stream_));
  PSV_CUDA(cudaStreamSynchronize(stream_));
}

void Model::pull_medium(const Model &src) {
  bind();
  const GridArray *gs[] = {&src.vp_, &src.vs_, &src.rho_};
  GridArray *gd[] = {&vp_, &vs_, &rho_};
  for (int i = 0; i < 3; ++i)
    PSV_CUDA(cudaMemcpyAsync(gd[i]->buf.as<char>(), gs[i]->buf.as<char>(), gs[i]->buf.bytes(), cudaMemcpyDefault, stream_));
  PSV_CUDA(cudaMemcpyAsync(medium_.as<char>(), src.medium_.as<char>(), medium_.bytes(), cudaMemcpyDefault, stream_));
  PSV_CUDA(cudaStreamSynchronize(stream_));
  if (fp32_) refresh_medium32();
}

void Model::pull_observed(const Model &src) {
  bind();
  if (obs_ux_.bytes() == 0) return;
  PSV_CUDA(cudaMemcpyAsync(obs_ux_.as<char>(), src.obs_ux_.as<char>(), obs_ux_.bytes(), cudaMemcpyDefault, stream_));
  PSV_CUDA(cudaMemcpyAsync(obs_uz_.as<char>(), src.obs_uz_.as<char>(), obs_uz_.bytes(), cudaMemcpyDefault, stream_));
  PSV_CUDA(cudaStreamSynchronize(stream_));
}

// ---------------------------------------------------------------------------------------------
// host <-> device helpers
// ---------------------------------------------------------------------------------------------
void Model::grid_to_host(const double *dev, double *host) const {
  PSV_CUDA(cudaMemcpy2DAsync(host, cfg_.nz * sizeof(double), dev, ld_ * sizeof(double),
                             cfg_.nz * sizeof(double), cfg_.nx, cudaMemcpyDeviceToHost, stream_));
  PSV_CUDA(cudaStreamSynchronize(stream_));
}

void Model::grid_from_host(double *dev, const double *host) const {
  // stream-ordered: a plain cudaMemcpy2D from pageable memory may return before the DMA lands,
  // and the model's non-blocking stream would not wait for it
  PSV_CUDA(cudaMemcpy2DAsync(dev, ld_ * sizeof(double), host, cfg_.nz * sizeof(double),
                             cfg_.nz * sizeof(double), cfg_.nx, cudaMemcpyHostToDevice, stream_));
  PSV_CUDA(cudaStreamSynchronize(stream_));
}

// Copy on the model's stream and wait for it (every buffer of a model is only ever touched
// on that stream, so this is the one ordering domain).
void Model::copy_sync(void *dst, const void *src, size_t bytes, cudaMemcpyKind kind) const {
  if (bytes == 0) return;
  PSV_CUDA(cudaMemcpyAsync(dst, src, bytes, kind, stream_));
  PSV_CUDA(cudaStreamSynchronize(stream_));
}

void Model::synchronize() {
  bind();
  PSV_CUDA(cudaStreamSynchronize(stream_));
}

void Model::timer_start() {
  bind();
  PSV_CUDA(cudaEventRecord(ev_t0_, stream_));
}

double Model::timer_stop() {
  bind();
  PSV_CUDA(cudaEventRecord(ev_t1_, stream_));
  PSV_CUDA(cudaEventSynchronize(ev_t1_));
  float ms = 0;
  PSV_CUDA(cudaEventElapsedTime(&ms, ev_t0_, ev_t1_));
  return ms;
}

void Model::check_shot(int i_shot) const {
  if (i_shot < 0 || i_shot >= cfg_.n_shots) throw invalid_argument("Shot index out of range.");
}

void Model::zero_dynamic_fields() {
  if (fp32_) PSV_CUDA(cudaMemsetAsync(fields32_.as<char>(), 0, fields32_.bytes(), stream_));
  PSV_CUDA(cudaMemsetAsync(fields_.as<char>(), 0, fields_.bytes(), stream_));
  cur_ = 0;
  if (poison_) poison_arrays(/*state_only=*/true);
}

// ---- debug: NaN canaries in everything the kernels may touch outside the grid proper ----
namespace {
constexpr unsigned long long kCanary = 0x7ff8dead0000beefull;  // a quiet NaN with a recognisable payload
// One padded grid array [guard | nx rows of ld | guard]: canaries into both guards and the pad columns nz..ld-1.
__global__ void poison_kernel(double *base, size_t guard, int nx, int ld, int nz) {
  const size_t total = 2 * guard + (size_t)nx * ld;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    bool hit = i < guard || i >= guard + (size_t)nx * ld;
    if (!hit) hit = (int)((i - guard) % ld) >= nz;
    if (hit) reinterpret_cast<unsigned long long *>(base)[i] = kCanary;
  }
}
__global__ void poison_range_kernel(double *p, size_t n) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    reinterpret_cast<unsigned long long *>(p)[i] = kCanary;
}
// guards only (pads are legitimately rewritten with whatever sat there): anything but the canary is a stray write
__global__ void count_guard_kernel(const double *base, size_t guard, size_t grid, unsigned long long *bad) {
  unsigned long long mine = 0;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < 2 * guard; i += (size_t)gridDim.x * blockDim.x) {
    const size_t at = i < guard ? i : grid + i;
    if (reinterpret_cast<const unsigned long long *>(base)[at] != kCanary) ++mine;
  }
  if (mine) atomicAdd(bad, mine);
}
}  // namespace

void Model::poison_arrays(bool state_only) {
  bind();
  const Config &c = cfg_;
  for (int k = 0; k < 10; ++k)
    poison_kernel<<<256, 256, 0, stream_>>>(fields_.as<double>() + (size_t)k * field_stride_, guard_, c.nx, ld_, c.nz);
  if (!state_only) {
    for (int k = 0; k < 3; ++k)
      poison_kernel<<<256, 256, 0, stream_>>>(medium_.as<double>() + (size_t)k * field_stride_, guard_, c.nx, ld_, c.nz);
    GridArray *grids[] = {&vp_, &vs_, &rho_, &vp_kernel_, &vs_kernel_, &density_v_kernel_};
    for (GridArray *g : grids) poison_kernel<<<256, 256, 0, stream_>>>(g->buf.as<double>(), guard_, c.nx, ld_, c.nz);
    // the slack around the Lame kernels and the snapshots that bulk copies of whole row segments run into
    poison_range_kernel<<<8, 256, 0, stream_>>>(lame_kernels_.as<double>(), kBulkPad);
    poison_range_kernel<<<8, 256, 0, stream_>>>(lame_kernels_.as<double>() + kBulkPad + 3 * grid_elems_, kBulkPad);
    for (DeviceBuffer &s : snapshots_)
      if (!s.empty() && !fp32_) {
        poison_range_kernel<<<8, 256, 0, stream_>>>(s.as<double>(), kBulkPad);
        poison_range_kernel<<<8, 256, 0, stream_>>>(s.as<double>() + kBulkPad + 5 * (size_t)n_snap_alloc_ * grid_elems_,
                                                   kBulkPad);
      }
  }
  PSV_CUDA(cudaGetLastError());
}

long long Model::count_guard_violations() {
  bind();
  PSV_CUDA(cudaStreamSynchronize(stream_));
  DeviceBuffer bad(sizeof(unsigned long long));
  std::vector<const double *> bases;
  for (int k = 0; k < 10; ++k) bases.push_back(fields_.as<double>() + (size_t)k * field_stride_);
  for (int k = 0; k < 3; ++k) bases.push_back(medium_.as<double>() + (size_t)k * field_stride_);
  for (const GridArray *g : {&vp_, &vs_, &rho_, &vp_kernel_, &vs_kernel_, &density_v_kernel_})
    bases.push_back(g->buf.as<double>());
  for (const double *b : bases)
    count_guard_kernel<<<64, 256, 0, stream_>>>(b, guard_, grid_elems_, bad.as<unsigned long long>());
  PSV_CUDA(cudaGetLastError());
  unsigned long long n = 0;
  copy_sync(&n, bad.as<char>(), sizeof n, cudaMemcpyDeviceToHost);
  return (long long)n;
}

void Model::ensure_snapshots(int i_shot) {
  DeviceBuffer &s = snapshots_[i_shot];
  const size_t bytes = (5 * (size_t)n_snap_alloc_ * grid_elems_ + 2 * kBulkPad) * (fp32_ ? sizeof(float) : sizeof(double));
  if (s.empty() && !snapshot_pool_.empty()) {
    // a pooled buffer may hold another shot's snapshots: every snapshot the adjoint reads is
    // rewritten by the stored forward run, and get_array reports only shots stored since
    s = std::move(snapshot_pool_.back());
    snapshot_pool_.pop_back();
  }
  if (s.empty()) {
    try {
      s.allocate(bytes);
    } catch (const cuda_error &e) {
      cudaGetLastError();
      std::ostringstream os;
      os << "Cannot keep the forward snapshots of shot " << i_shot << " in device memory ("
         << (bytes >> 20) << " MiB: 5 fields x " << n_snap_alloc_ << " snapshots x " << cfg_.nx << "x"
         << cfg_.nz << "). Release other shots with psv_release_snapshots / run shots one at a time. ["
         << e.what() << "]";
      throw length_error(os.str());
    }
  }
  if (poison_ && !fp32_) {  // the slack around the snapshots that bulk copies of whole row segments run into
    poison_range_kernel<<<8, 256, 0, stream_>>>(s.as<double>(), kBulkPad);
    poison_range_kernel<<<8, 256, 0, stream_>>>(s.as<double>() + kBulkPad + 5 * (size_t)n_snap_alloc_ * grid_elems_, kBulkPad);
    PSV_CUDA(cudaGetLastError());
  }
}

double *Model::snapshot_ptr(int i_shot, int field, int snap) const {
  return snapshots_[i_shot].as<double>() + kBulkPad + ((size_t)field * n_snap_alloc_ + snap) * grid_elems_;
}

void Model::release_snapshots(int i_shot) {
  bind();
  PSV_CUDA(cudaStreamSynchronize(stream_));
  if (i_shot < 0) {
    for (DeviceBuffer &s : snapshots_) s.release();
    snapshot_pool_.clear();
  } else {
    check_shot(i_shot);
    if (!snapshots_[i_shot].empty()) snapshot_pool_.push_back(std::move(snapshots_[i_shot]));
    snapshots_[i_shot] = DeviceBuffer();
  }
}

void Model::kernel_buffer(void **ptr, size_t *count) const {
  *ptr = lame();
  *count = 3 * grid_elems_;
}

// ---------------------------------------------------------------------------------------------
// sweeps
// ---------------------------------------------------------------------------------------------
namespace {
template <class M, int MODE>
void launch_stress_t(int variant, const StressArgs &a, int block, cudaStream_t st) {
  if (variant == 0) {
    dim3 grid((a.g.nz + 31) / 32, (a.g.nx + 7) / 8), blk(32, 8);
    stress_point_kernel<M, MODE><<<grid, blk, 0, st>>>(a);
  } else {
    const int pairs = a.g.ld / 2;
    dim3 grid((pairs + block - 1) / block, (a.g.nx + a.strip - 1) / a.strip), blk(block);
    stress_march_kernel<M, MODE><<<grid, blk, 0, st>>>(a);
  }
}
template <class M>
void launch_stress_m(int variant, int mode, const StressArgs &a, int block, cudaStream_t st) {
  if (mode == MODE_SNAPSHOT) launch_stress_t<M, MODE_SNAPSHOT>(variant, a, block, st);
  else if (mode == MODE_ACCUMULATE) launch_stress_t<M, MODE_ACCUMULATE>(variant, a, block, st);
  else launch_stress_t<M, MODE_PLAIN>(variant, a, block, st);
}
template <class M>
void launch_velocity_t(int variant, const VelocityArgs &a, int block, cudaStream_t st) {
  if (variant == 0) {
    dim3 grid((a.g.nz + 31) / 32, (a.g.nx + 7) / 8), blk(32, 8);
    velocity_point_kernel<M><<<grid, blk, 0, st>>>(a);
  } else {
    const int pairs = a.g.ld / 2;
    dim3 grid((pairs + block - 1) / block, (a.g.nx + a.strip - 1) / a.strip), blk(block);
    velocity_march_kernel<M><<<grid, blk, 0, st>>>(a);
  }
}
}  // namespace

void Model::launch_stress(double sdt, int mode, int i_shot, int snap) {
  const Config &c = cfg_;
  StressArgs a{};
  a.g = GridShape{c.nx, c.nz, ld_, c.np_boundary, 0, c.nx};
  a.k = SweepConsts{sdt, c.dx, c.dz, 1.0 / c.dx, 1.0 / c.dz, 9.0 / 8.0, 1.0 / 24.0};
  a.vx = F(0);
  a.vz = F(1);
  a.txx = F(2);
  a.tzz = F(3);
  a.txz = F(4);
  a.lm = lm_.data();
  a.mu = mu_.data();
  a.taper_table = taper_table_.as<>();
  if (mode != MODE_PLAIN) {
    a.s_vx = snapshot_ptr(i_shot, 0, snap);
    a.s_vz = snapshot_ptr(i_shot, 1, snap);
    a.s_txx = snapshot_ptr(i_shot, 2, snap);
    a.s_tzz = snapshot_ptr(i_shot, 3, snap);
    a.s_txz = snapshot_ptr(i_shot, 4, snap);
  }
  if (mode == MODE_ACCUMULATE) {
    a.k_lam = lame();
    a.k_mu = lame() + grid_elems_;
    a.k_rho = lame() + 2 * grid_elems_;
    a.wx0 = c.np_boundary + c.nx_inner_boundary;
    a.wx1 = c.np_boundary + c.nx_inner - c.nx_inner_boundary;
    a.wz0 = c.np_boundary + c.nz_inner_boundary;
    a.wz1 = c.np_boundary + c.nz_inner - c.nz_inner_boundary;
    a.sidt = c.snapshot_interval * c.dt;
  }
  a.strip = strip_;
  if (exact_) launch_stress_m<ExactMath>(variant_, mode, a, block_, stream_);
  else launch_stress_m<FastMath>(variant_, mode, a, block_, stream_);
  ++stat_launches_;
}

void Model::launch_velocity(double sdt) {
  const Config &c = cfg_;
  VelocityArgs a{};
  a.g = GridShape{c.nx, c.nz, ld_, c.np_boundary, 0, c.nx};
  a.k = SweepConsts{sdt, c.dx, c.dz, 1.0 / c.dx, 1.0 / c.dz, 9.0 / 8.0, 1.0 / 24.0};
  a.txx = F(2);
  a.tzz = F(3);
  a.txz = F(4);
  a.vx = F(0);
  a.vz = F(1);
  a.b = b_.data();
  a.taper_table = taper_table_.as<>();
  a.strip = strip_;
  if (exact_) launch_velocity_t<ExactMath>(variant_, a, block_, stream_);
  else launch_velocity_t<FastMath>(variant_, a, block_, stream_);
  ++stat_launches_;
}

// ---- variant 2: one fused launch per time step ------------------------------------------------
void Model::fill_fused_model(int i_shot, void *out) const {
  const Config &c = cfg_;
  FusedModel fm{};
  for (int p = 0; p < 2; ++p)
    for (int k = 0; k < 5; ++k) fm.f[p][k] = field_ptr(p, k);
  fm.lm = lm_.data();
  fm.mu = mu_.data();
  fm.b = b_.data();
  fm.snap = snapshots_[i_shot].empty() ? nullptr : snapshots_[i_shot].as<double>() + kBulkPad;
  fm.k_lam = lame();
  fm.k_mu = lame() + grid_elems_;
  fm.k_rho = lame() + 2 * grid_elems_;
  const size_t shot_rows = (size_t)i_shot * c.nr * c.nt;
  fm.ux = rtf_ux_.as<>() + shot_rows;
  fm.uz = rtf_uz_.as<>() + shot_rows;
  fm.ax = a_stf_ux_.as<>() + shot_rows;
  fm.az = a_stf_uz_.as<>() + shot_rows;
  *static_cast<FusedModel *>(out) = fm;
}

// Device copy of the per-model pointers: this model first, then the followers of a batch.
void Model::fused_begin(int i_shot, int point_mode) {
  (void)point_mode;
  std::vector<FusedModel> fm(1 + batch_followers_.size());
  fill_fused_model(i_shot, &fm[0]);
  for (size_t i = 0; i < batch_followers_.size(); ++i) batch_followers_[i]->fill_fused_model(i_shot, &fm[i + 1]);
  const size_t bytes = fm.size() * sizeof(FusedModel);
  if (fused_model_.bytes() < bytes) fused_model_.allocate(bytes);
  // stream-ordered upload from a local object: wait for it before the object dies
  copy_sync(fused_model_.as<FusedModel>(), fm.data(), bytes, cudaMemcpyHostToDevice);
  if (fp32_) {
    std::vector<FusedModel32> f32(1 + batch_followers_.size());
    fill_fused_model32(i_shot, &f32[0]);
    for (size_t i = 0; i < batch_followers_.size(); ++i) batch_followers_[i]->fill_fused_model32(i_shot, &f32[i + 1]);
    const size_t b32 = f32.size() * sizeof(FusedModel32);
    if (fused_model32_.bytes() < b32) fused_model32_.allocate(b32);
    copy_sync(fused_model32_.as<FusedModel32>(), f32.data(), b32, cudaMemcpyHostToDevice);
  }
}

// Persisting-L2 window over [lm | mu | b]: a carve-out of the 126 MB L2 keeps (a fraction of) the medium
// resident across the launches of a time loop, the rest of the window streams.
void Model::apply_l2_window() {
  bind();
  cudaStreamAttrValue attr{};
  if (l2_persist_mb_ > 0) {
    int max_persist = 0, max_window = 0;
    PSV_CUDA(cudaDeviceGetAttribute(&max_persist, cudaDevAttrMaxPersistingL2CacheSize, device_));
    PSV_CUDA(cudaDeviceGetAttribute(&max_window, cudaDevAttrMaxAccessPolicyWindowSize, device_));
    const size_t want = std::min<size_t>((size_t)l2_persist_mb_ << 20, (size_t)max_persist);
    PSV_CUDA(cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, want));
    const size_t bytes = std::min<size_t>(medium_.bytes(), (size_t)max_window);
    attr.accessPolicyWindow.base_ptr = medium_.as<char>();
    attr.accessPolicyWindow.num_bytes = bytes;
    attr.accessPolicyWindow.hitRatio = (float)std::min(1.0, (double)want / (double)bytes);
    attr.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
    attr.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
  } else {
    attr.accessPolicyWindow.num_bytes = 0;
  }
  PSV_CUDA(cudaStreamSetAttribute(stream_, cudaStreamAttributeAccessPolicyWindow, &attr));
}

// ---- optional FP32 wavefield mode (fused32.cuh) -----------------------------------------------
void Model::ensure_fp32() {
  if (fields32_.empty()) {
    fields32_.allocate(10 * field_stride_ * sizeof(float));
    medium32_.allocate(3 * field_stride_ * sizeof(float));
  }
  refresh_medium32();
}

void Model::refresh_medium32() {  // float images of lm, mu, b (pad columns included, guards stay zero)
  const double *src[3] = {lm_.data(), mu_.data(), b_.data()};
  for (int k = 0; k < 3; ++k)
    to_float_kernel<<<blocks_for(grid_elems_, 256), 256, 0, stream_>>>(src[k], medium32_ptr(k), grid_elems_);
  PSV_CUDA(cudaGetLastError());
  PSV_CUDA(cudaStreamSynchronize(stream_));
}

void Model::fill_fused_model32(int i_shot, void *out) const {
  const Config &c = cfg_;
  FusedModel32 fm{};
  for (int p = 0; p < 2; ++p)
    for (int k = 0; k < 5; ++k) fm.f[p][k] = field32_ptr(p, k);
  fm.lm = medium32_ptr(0);
  fm.mu = medium32_ptr(1);
  fm.b = medium32_ptr(2);
  fm.snap = snapshots_[i_shot].empty() ? nullptr : snapshots_[i_shot].as<float>() + kBulkPad;
  fm.k_lam = lame();
  fm.k_mu = lame() + grid_elems_;
  fm.k_rho = lame() + 2 * grid_elems_;
  const size_t shot_rows = (size_t)i_shot * c.nr * c.nt;
  fm.ux = rtf_ux_.as<>() + shot_rows;
  fm.uz = rtf_uz_.as<>() + shot_rows;
  fm.ax = a_stf_ux_.as<>() + shot_rows;
  fm.az = a_stf_uz_.as<>() + shot_rows;
  *static_cast<FusedModel32 *>(out) = fm;
}

// Current state array k as doubles: the array itself, or (FP32 mode) a converted image.
const double *Model::field_as_double(int k) {
  if (!fp32_) return F(k);
  if (conv_.buf.empty()) conv_ = make_grid();
  to_double_kernel<<<blocks_for(grid_elems_, 256), 256, 0, stream_>>>(field32_ptr(cur_, k), conv_.data(), grid_elems_);
  PSV_CUDA(cudaGetLastError());
  return conv_.data();
}

// Rows per warp such that the launch is a whole number of waves of resident blocks (the warps
// of a launch run in lock-step through their strips, so a partial last wave costs a full one).
int Model::fused_auto_strip(int n_zwarps, int batch) {
  if (fused_resident_blocks_ == 0) {
    int per_sm = 0, sms = 0;
    per_sm = fused_step_occupancy(fused_block_);
    PSV_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device_));
    fused_resident_blocks_ = std::max(1, per_sm * sms);
  }
  const int wpb = fused_block_ / 32;
  const int blocks_x = ((n_zwarps + wpb - 1) / wpb) * batch;
  const int nx = cfg_.nx;
  for (int waves = 1;; ++waves) {
    const int n_strips = std::max(1, waves * fused_resident_blocks_ / blocks_x);
    const int strip = (nx + n_strips - 1) / n_strips;
    if (strip <= 96 || n_strips >= nx) return std::max(strip, 8);
  }
}

// Consumer warps per CTA of variant 3: the compiled width (8, 4 or 3 warps of 28 pairs) that covers
// a row with the fewest warps; wide rows tie and take 8, the 80-pair rows of the default
// configuration take 3 (84 pairs) instead of leaving 5 of 8 warps idle.
int Model::pipe_warps() const {
  if (pipe_warps_ != 0) return pipe_warps_;
  if (tuned_warps_ != 0) return tuned_warps_;
  int best = 8, cost = 1 << 30;
  for (int nw : {8, 4, 3}) {
    const int owned = nw * kFusedOwnedPairs;
    const int c = ((ld_ / 2 + owned - 1) / owned) * nw;
    if (c < cost) {
      cost = c;
      best = nw;
    }
  }
  return best;
}

void Model::pipe_geometry(int batch, int *tiles, int *strip, size_t *smem) {
  const int nw = pipe_warps();
  const int owned = nw * kFusedOwnedPairs;
  *tiles = (ld_ / 2 + owned - 1) / owned;
  *smem = pipe_smem_bytes(nw, pipe_stages_, false);
  if (*smem > kMaxDynamicSmem) throw invalid_argument("pipe_stages: the ring does not fit in shared memory");
  if (pipe_resident_ == 0) {
    int sms = 0;
    const int per_sm = pipe_occupancy(nw, *smem);
    PSV_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device_));
    pipe_resident_ = std::max(1, per_sm * sms);
  }
  if (fused_strip_ > 0) {
    *strip = fused_strip_;
    return;
  }
  if (tuned_strip_ > 0 && tuned_batch_ == batch) {
    *strip = tuned_strip_;
    return;
  }
  const int blocks_x = *tiles * batch;
  const int nx = cfg_.nx;
  for (int waves = 1;; ++waves) {
    const int n_strips = std::max(1, waves * pipe_resident_ / blocks_x);
    const int st = (nx + n_strips - 1) / n_strips;
    if (st <= 128 || n_strips >= nx) {
      // small grids are bound by the serial latency of a CTA's march (~0.6 us per row), so the shortest strips
      // win although each re-reads 8 halo rows: 250x150, one model, under graph replay: 9.0 us per step with
      // 6 rows, 5.9 with 2, 5.4 with 1 (profiles/r1o_graph_replay.md)
      *strip = std::max(st, 1);
      return;
    }
  }
}

// Launch geometry of the variant-3 step kernel by measurement.  On large grids the step is bandwidth-bound only when
// the CTAs are balanced over the SMs: too many short strips re-read halo rows and refill their pipelines, too few long
// ones leave SMs idle, and which consumer-warp count wins depends on how the row length tiles (on a 1044 x 4100 slab
// the formula above gives 88.6 us per step, the best candidate 76.9; profiles/r2g_slab_shape.md).  So the first time
// loop of a model (and of every ensemble size) spends a few milliseconds on plain step launches over its own state
// arrays -- CTA counts that are small multiples of the SM count, for each compiled warp count -- and keeps the fastest.
// Results do not depend on the geometry (same arithmetic per point), and the caller zeroes the state afterwards.
// The candidates run with the point operations of the loop that follows (receiver samples / injections at time index
// 0, which that loop writes again): the warp that carries them is the critical path of its CTA row.
namespace {
std::mutex g_tune_mutex;
std::map<std::vector<long long>, std::pair<int, int>> g_tune_cache;  // shape key -> (consumer warps, strip rows)
}  // namespace

void Model::tune_pipe_geometry(int i_shot, int point_mode) {
  const int batch = 1 + (int)batch_followers_.size();
  if (variant_ != 3 || tune_ == 0 || pipe_warps_ != 0 || fused_strip_ != 0) return;
  if (tune_ < 0 && (long long)cfg_.nx * cfg_.nz * batch < (1ll << 21)) return;  // small grids: latency rules, see above
  if (tuned_strip_ > 0 && tuned_batch_ == batch) return;
  // measured once per process for this device and problem shape: fresh objects of a Monte-Carlo run, the replicas of
  // a device group and the slabs of equal shape reuse it
  const std::vector<long long> key = {device_, cfg_.nx, cfg_.nz, ld_, batch, fp32_, exact_, pipe_stages_, l2_mode_,
                                      slab_ ? x_off_ : -1, nx_global_, point_mode};
  {
    std::lock_guard<std::mutex> lock(g_tune_mutex);
    auto f = g_tune_cache.find(key);
    if (f != g_tune_cache.end() && tune_ < 1) {
      tuned_warps_ = f->second.first;
      tuned_strip_ = f->second.second;
      tuned_batch_ = batch;
      return;
    }
  }
  bind();
  const auto wall0 = std::chrono::steady_clock::now();
  int sms = 0;
  PSV_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device_));
  fused_begin(i_shot, point_mode);
  const int cur0 = cur_;
  const long long launches0 = stat_launches_;
  tuned_warps_ = tuned_strip_ = 0;
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  PSV_CUDA(cudaEventCreate(&e0));
  PSV_CUDA(cudaEventCreate(&e1));
  int best_w = 0, best_s = 0;
  struct Candidate {
    int warps, strip;
    float ms;
  };
  std::vector<Candidate> cands;
  try {
    auto time_steps = [&](int nw, int strip, int steps) {
      pipe_warps_ = nw;
      fused_strip_ = strip;
      pipe_resident_ = 0;
      launch_fused(cfg_.dt, MODE_PLAIN, point_mode, i_shot, 0, 0);  // warm-up (attributes, caches)
      PSV_CUDA(cudaEventRecord(e0, stream_));
      for (int k = 0; k < steps; ++k) launch_fused(cfg_.dt, MODE_PLAIN, point_mode, i_shot, 0, 0);
      PSV_CUDA(cudaEventRecord(e1, stream_));
      PSV_CUDA(cudaEventSynchronize(e1));
      float ms = 0;
      PSV_CUDA(cudaEventElapsedTime(&ms, e0, e1));
      return ms / steps;
    };
    for (int nw : kPipeWarpChoices) {
      const int owned = nw * kFusedOwnedPairs, tiles = (ld_ / 2 + owned - 1) / owned;
      if (pipe_smem_bytes(nw, pipe_stages_, false) > kMaxDynamicSmem) continue;
      int last_strip = 0;
      for (int f = 1; f <= 8; ++f) {  // about f CTAs per SM in total
        const int n_strips = std::max(1, f * sms / (tiles * batch));  // at most f CTAs per SM: no ragged last wave
        const int strip = (cfg_.nx + n_strips - 1) / n_strips;
        if (strip == last_strip || strip < 16) continue;
        last_strip = strip;
        cands.push_back({nw, strip, 0.f});
      }
    }
    if (cands.empty()) cands.push_back({8, std::max(cfg_.nx, 1), 0.f});
    time_steps(cands[0].warps, cands[0].strip, 12);  // clocks up before anything is compared
    for (Candidate &c : cands) c.ms = time_steps(c.warps, c.strip, 3);
    std::sort(cands.begin(), cands.end(), [](const Candidate &x, const Candidate &y) { return x.ms < y.ms; });
    float best_ms = 1e30f;
    for (size_t i = 0; i < cands.size() && i < 4; ++i) {  // the leaders again, longer
      const float ms = time_steps(cands[i].warps, cands[i].strip, 10);
      if (ms < best_ms) {
        best_ms = ms;
        best_w = cands[i].warps;
        best_s = cands[i].strip;
      }
    }
  } catch (...) {
    pipe_warps_ = fused_strip_ = 0;
    pipe_resident_ = 0;
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    throw;
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  pipe_warps_ = fused_strip_ = 0;
  pipe_resident_ = 0;
  tuned_warps_ = best_w;
  tuned_strip_ = best_s;
  tuned_batch_ = batch;
  {
    std::lock_guard<std::mutex> lock(g_tune_mutex);
    g_tune_cache[key] = {best_w, best_s};
  }
  cur_ = cur0;
  // the caller zeroes THIS model's state next; the other members of an ensemble were zeroed before it was called
  PSV_CUDA(cudaStreamSynchronize(stream_));
  for (Model *f : batch_followers_) {
    f->zero_dynamic_fields();
    PSV_CUDA(cudaStreamSynchronize(f->stream_));
  }
  bind();
  stat_launches_ = launches0;
  drop_graphs();
  stat_tune_ms_ = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - wall0).count();
}

// Depth of the input ring of variant 4: the option, or (0) as many stages as fit next to the exchange ring.
int Model::pair_stages() const {
  if (pair_stages_ != 0) return pair_stages_;
  int stages = 4;
  while (stages > 2 && pair_smem(pair_warps_, stages) > kMaxDynamicSmem) --stages;
  return stages;
}

// Launch geometry of variant 4 (two steps per launch): even z-tiles over the pairs that hold real columns, and rows per
// CTA such that the launch is a whole number of waves (one CTA per SM: the rings fill its smem).
void Model::pair_geometry(int batch, int *tiles, int *tile_pairs, int *strip, size_t *smem) {
  const int nw = pair_warps_;
  const int max_tile = nw * kFusedOwnedPairs - 4;
  const int pairs = (cfg_.nz + 1) / 2;
  *tiles = (pairs + max_tile - 1) / max_tile;
  *tile_pairs = (pairs + *tiles - 1) / *tiles;
  const int stages = pair_stages();
  *smem = pair_smem(nw, stages);
  if (*smem > kMaxDynamicSmem) throw invalid_argument("pair_stages: the rings do not fit in shared memory");
  if (pair_strip_ > 0) {
    *strip = pair_strip_;
    return;
  }
  int sms = 0;
  PSV_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device_));
  const int blocks_x = *tiles * batch;
  const int nx = cfg_.nx;
  for (int waves = 1;; ++waves) {
    const int n_strips = std::max(1, waves * sms / blocks_x);
    const int st = (nx + n_strips - 1) / n_strips;
    if (st <= 192 || n_strips >= nx) {
      *strip = std::max(st, 8);
      return;
    }
  }
}

// Everything a step launch needs besides the per-model pointers.
void Model::fill_fused_common(void *out, double sdt, int point_mode, int i_shot, int snap, int it) {
  const Config &c = cfg_;
  FusedCommon fc{};
  fc.g = GridShape{c.nx, c.nz, ld_, c.np_boundary, x_off_, nx_global_};
  fc.k = SweepConsts{sdt, c.dx, c.dz, 1.0 / c.dx, 1.0 / c.dz, 9.0 / 8.0, 1.0 / 24.0};
  fc.taper_table = taper_table_.as<>();
  fc.parity = cur_;
  fc.field_stride = field_stride_;
  fc.n_zwarps = (ld_ / 2 + kFusedOwnedPairs - 1) / kFusedOwnedPairs;
  fc.strip = 0;
  fc.tile_pairs = 0;
  fc.it_step = 0;
  fc.l2_mode = l2_mode_;
  fc.pdl = (pdl_ && variant_ == 3 && !fp32_) ? 1 : 0;
  fc.rec_increments = receiver_increments() ? 1 : 0;
  fc.snap_field_stride = (size_t)n_snap_alloc_ * grid_elems_;
  fc.snap_offset = (size_t)snap * grid_elems_;
  // free window in local rows; a slab correlates only on the rows it owns
  fc.wx0 = std::max(c.np_boundary + c.nx_inner_boundary - x_off_, halo_lo_);
  fc.wx1 = std::min(c.np_boundary + c.nx_inner - c.nx_inner_boundary - x_off_, c.nx - halo_hi_);
  fc.wz0 = c.np_boundary + c.nz_inner_boundary;
  fc.wz1 = c.np_boundary + c.nz_inner - c.nz_inner_boundary;
  fc.sidt = c.snapshot_interval * c.dt;
  fc.point_mode = point_mode;
  fc.nt = c.nt;
  fc.it = it;
  fc.dt = c.dt;
  fc.dxdz = c.dx * c.dz;
  fc.rdxdz = 1.0 / (c.dx * c.dz);
  fc.rec_row_start = rec_row_start_.as<int>();
  fc.rec_z = rec_row_z_.as<int>();
  fc.rec_x = rec_row_x_.as<int>();
  fc.rec_id = rec_row_id_.as<int>();
  fc.op_row_start = plans_[i_shot].op_row_start.as<int>();
  fc.ops = plans_[i_shot].row_ops.as<RowOp>();
  fc.stf = stf_.as<>();
  *static_cast<FusedCommon *>(out) = fc;
}

void Model::launch_fused(double sdt, int mode, int point_mode, int i_shot, int snap, int it, int pair_step) {
  const Config &c = cfg_;
  if (slab_ && variant_ != 2 && variant_ != 3)
    throw invalid_argument("A slab of a split grid runs the fused variants 2 and 3 only.");
  FusedCommon fc{};
  fill_fused_common(&fc, sdt, point_mode, i_shot, snap, it);
  fc.strip = fused_strip_ > 0 ? fused_strip_ : (variant_ >= 3 ? 0 : fused_auto_strip(fc.n_zwarps, 1 + (int)batch_followers_.size()));
  const FusedModel *models = fused_model_.as<FusedModel>();
  const int batch = 1 + (int)batch_followers_.size();
  if (fp32_) {
    if (variant_ == 3) {  // the march fed by the TMA ring, geometry like the FP64 default (option, measured, or formula)
      int tiles = 0, strip = 0;
      size_t smem64 = 0;
      pipe_geometry(batch, &tiles, &strip, &smem64);
      const int nw = pipe_warps();
      fc.strip = strip;
      dim3 grid(tiles, (c.nx + strip - 1) / strip, batch);
      launch_pipe32(nw, mode, fc, fused_model32_.as<FusedModel32>(), grid, pipe_stages_,
                    pipe32_smem_bytes(nw, pipe_stages_), stream_);
      cur_ ^= 1;
      ++stat_launches_;
      return;
    }
    const int wpb = 4;
    if (fused_strip_ > 0) {
      fc.strip = fused_strip_;
    } else {  // whole waves of resident blocks, as fused_auto_strip does for variant 2
      if (fused32_resident_ == 0) {
        int sms = 0;
        PSV_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device_));
        fused32_resident_ = std::max(1, fused32_occupancy() * sms);
      }
      const int blocks_x = ((fc.n_zwarps + wpb - 1) / wpb) * batch;
      fc.strip = 8;
      for (int waves = 1;; ++waves) {
        const int n_strips = std::max(1, waves * fused32_resident_ / blocks_x);
        const int st = (c.nx + n_strips - 1) / n_strips;
        if (st <= 64 || n_strips >= c.nx) {
          fc.strip = std::max(st, 2);
          break;
        }
      }
    }
    dim3 grid((fc.n_zwarps + wpb - 1) / wpb, (c.nx + fc.strip - 1) / fc.strip, batch);
    launch_fused32(mode, fc, fused_model32_.as<FusedModel32>(), grid, stream_);
  } else if (variant_ == 4 && pair_step != 0) {
    int tiles = 0, tile_pairs = 0, strip = 0;
    size_t smem = 0;
    pair_geometry(batch, &tiles, &tile_pairs, &strip, &smem);
    fc.strip = strip;
    fc.tile_pairs = tile_pairs;
    fc.it_step = pair_step;
    dim3 grid(tiles, (c.nx + strip - 1) / strip, batch);
    launch_pair(exact_, pair_warps_, mode, fc, models, grid, pair_stages(), smem, stream_);
    ++stat_pair_launches_;
  } else if (variant_ == 3 || variant_ == 4) {
    int tiles = 0, strip = 0;
    size_t smem = 0;
    pipe_geometry(batch, &tiles, &strip, &smem);
    fc.strip = strip;
    dim3 grid(tiles, (c.nx + strip - 1) / strip, batch);
    const int nw = pipe_warps();
    int stages = pipe_stages_;
    if (mode == MODE_ACCUMULATE) {
      stages = pipe_corr_stages(nw, stages);
      smem = pipe_smem_bytes(nw, stages, true);
    }
    launch_pipe(exact_, nw, mode, fc, models, grid, stages, smem, stream_);
  } else {
    const int wpb = fused_block_ / 32;
    dim3 grid((fc.n_zwarps + wpb - 1) / wpb, (c.nx + fc.strip - 1) / fc.strip, batch);
    launch_fused_step(exact_, mode, fc, models, grid, fused_block_, stream_);
  }
  cur_ ^= 1;
  ++stat_launches_;
}

// ---- resident whole-loop kernel (resident.cuh): one launch per simulate call on grids that fit on chip ----
namespace {
constexpr long long kResidentAutoPoints = 1 << 18;
}  // namespace

// Threads per CTA of the loop kernel: one per pair of columns of the rows a strip updates (+ halo), at most the
// compiled maximum.  resident.cuh gives each source / receiver of a strip to ONE thread, so this is also the number
// of point operations a strip can hold.
static int resident_threads(int ld, int rows) { return std::min(kResidentThreads, round_up((rows + 4) * (ld / 2), 32)); }

// Rows per CTA, CTAs and shared memory of the loop kernel for this grid, or false when it does not apply.
// `forced` (option "resident" = 1) turns the reasons into errors.
bool Model::resident_plan(int i_shot, bool adjoint, int *rows, int *ctas, size_t *smem) {
  const bool forced = resident_ == 1;
  auto no = [&](const char *why) {
    if (forced) throw invalid_argument(std::string("resident = 1: ") + why);
    return false;
  };
  if (resident_ == 0) return false;
  if (slab_) return no("not available on a slab of a split grid");
  if (fp32_) return no("not available in the FP32 wavefield mode");
  if (variant_ < 2) return no("needs a fused variant (>= 2)");
  if (profile_) return no("per-launch profiling needs one launch per step");
  if (!batch_followers_.empty()) return no("ensembles share per-step launches");
  if (!forced && (variant_ != 3 || (long long)cfg_.nx * cfg_.nz > kResidentAutoPoints)) return false;
  int r = resident_rows_ > 0 ? resident_rows_ : 2;
  for (;; ++r) {
    const size_t bytes = resident_smem_bytes(ld_, r);
    if (bytes > kResidentSmemBudget) return no("the strips do not fit in shared memory");
    auto f = resident_capacity_.find(r);
    if (f == resident_capacity_.end()) f = resident_capacity_.emplace(r, resident_max_ctas(bytes)).first;
    const int need = (cfg_.nx + r - 1) / r;
    if (need <= f->second) {
      *rows = r;
      *ctas = need;
      *smem = bytes;
      break;
    }
    if (f->second == 0) return no("cooperative launch is not available");
    if (resident_rows_ > 0) return no("resident_rows: too many strips to be co-resident");
  }
  // a strip's point operations are handled one per thread OF THE LAUNCH (narrow grids run fewer than
  // kResidentThreads threads, see resident_threads): the limits are the launched count, not the compiled maximum
  const int threads = resident_threads(ld_, *rows);
  const int max_ops = std::min(threads, kResidentMaxOps);
  const std::vector<int> &ops = plans_[i_shot].op_row_start_host, &rec = rec_row_start_host_;
  for (int x = 0; x < cfg_.nx; x += *rows) {
    const int xe = std::min(x + *rows, cfg_.nx);
    if ((!adjoint && (ops[xe] - ops[x] > max_ops || rec[xe] - rec[x] > threads)) ||
        (adjoint && rec[xe] - rec[x] > max_ops))
      return no("too many sources / receivers in one strip");
  }
  return true;
}

void Model::run_resident(int i_shot, bool adjoint, bool store, int rows, int ctas, size_t smem) {
  const Config &c = cfg_;
  ResidentArgs a{};
  const int point_mode = adjoint ? (c.nr > 0 ? POINT_ADJOINT : POINT_NONE) : POINT_FORWARD;
  fill_fused_common(&a.c, adjoint ? -c.dt : c.dt, point_mode, i_shot, 0, 0);
  a.c.rec_increments = 1;
  a.models = fused_model_.as<FusedModel>();
  // packets of the velocity rows in flight: [2 parities][vx vz][nx][ld] x 16 bytes + the abort word.  Flags are
  // step counts on top of a base that grows from launch to launch, so the buffer is zeroed only when it is
  // (re)allocated or the 32-bit flags would wrap.
  const size_t packet_bytes = 4 * grid_elems_ * 16 + 128;
  if (resident_packets_.bytes() < packet_bytes || resident_flag_base_ + 2ull * (unsigned)c.nt + 4 > 0xffffff00ull) {
    PSV_CUDA(cudaStreamSynchronize(stream_));
    resident_packets_.allocate(packet_bytes);
    resident_flag_base_ = 0;
  }
  a.packets = resident_packets_.as<uint4>();
  a.abort = reinterpret_cast<unsigned *>(resident_packets_.as<char>() + 4 * grid_elems_ * 16);
  resident_abort_word_ = a.abort;
  a.flag_base = (unsigned)resident_flag_base_;
  resident_flag_base_ += (unsigned)c.nt + 1;
  a.rows = rows;
  a.nsteps = c.nt;
  a.direction = adjoint ? -1 : +1;
  a.store = store ? 1 : 0;
  a.si = c.snapshot_interval;
  a.grid_elems = grid_elems_;
  a.timing = nullptr;
  if (resident_timing_) {
    const size_t tb = (size_t)ctas * 8 * sizeof(long long);
    if (resident_timing_buf_.bytes() < tb) resident_timing_buf_.allocate(tb);
    a.timing = resident_timing_buf_.as<long long>();
  }
  const int threads = resident_threads(ld_, rows);
  // a spin-limit abort of an earlier launch (debugger, time-slicing) must not fail every later call of this model
  PSV_CUDA(cudaMemsetAsync(a.abort, 0, sizeof(unsigned), stream_));
  launch_resident(exact_, a, dim3(ctas, 1), threads, smem, stream_);
  cur_ = c.nt & 1;
  ++stat_launches_;
  stat_resident_ = ctas;
}

// After the loop kernel: a wait that timed out (a CTA never published its step) must not pass as a result.
void Model::resident_check() {
  unsigned abort = 0;
  PSV_CUDA(cudaMemcpyAsync(&abort, resident_abort_word_, sizeof(unsigned), cudaMemcpyDeviceToHost, stream_));
  PSV_CUDA(cudaStreamSynchronize(stream_));
  if (abort) throw cuda_error("resident loop kernel: a strip waited for its neighbours longer than the spin limit");
  if (resident_timing_ && stat_resident_ > 0) {  // mean / max over the CTAs of thread 0's cycles per step and phase
    std::vector<long long> t((size_t)stat_resident_ * 8);
    copy_sync(t.data(), resident_timing_buf_.as<long long>(), t.size() * sizeof(long long), cudaMemcpyDeviceToHost);
    for (int k = 0; k < 8; ++k) {
      double sum = 0, mx = 0;
      for (int b = 0; b < stat_resident_; ++b) {
        sum += (double)t[(size_t)b * 8 + k];
        mx = std::max(mx, (double)t[(size_t)b * 8 + k]);
      }
      stat_resident_phase_[k] = sum / stat_resident_ / std::max(cfg_.nt, 1);
      stat_resident_phase_max_[k] = mx / std::max(cfg_.nt, 1);
    }
  }
}

namespace {
struct Profiler {  // CUDA events around sampled sweep launches
  std::vector<cudaEvent_t> &ev;
  cudaStream_t st;
  bool on;
  int every, n = 0;
  Profiler(std::vector<cudaEvent_t> &e, cudaStream_t s, bool enabled, int nt)
      : ev(e), st(s), on(enabled), every(std::max(1, nt / 32)) {
    if (on && ev.empty()) {
      ev.resize(3 * 40);
      for (cudaEvent_t &x : ev) PSV_CUDA(cudaEventCreate(&x));
    }
  }
  int countdown = 0;
  // every `every`-th step, skipping snapshot / correlation steps (`special`): the sampled mean is
  // the plain sweep's duration
  bool sample(bool special) {
    if (!on || 3 * (n + 1) > (int)ev.size()) return false;
    if (countdown > 0) {
      --countdown;
      return false;
    }
    if (special) return false;
    countdown = every - 1;
    return true;
  }
  void mark(int k) { PSV_CUDA(cudaEventRecord(ev[3 * n + k], st)); }
  void next() { ++n; }
  void collect(double &stress_ms, double &velocity_ms) {
    stress_ms = velocity_ms = 0;
    if (!on || n == 0) return;
    for (int i = 0; i < n; ++i) {
      float a = 0, b = 0;
      PSV_CUDA(cudaEventElapsedTime(&a, ev[3 * i], ev[3 * i + 1]));
      PSV_CUDA(cudaEventElapsedTime(&b, ev[3 * i + 1], ev[3 * i + 2]));
      stress_ms += a;
      velocity_ms += b;
    }
    stress_ms /= n;
    velocity_ms /= n;
  }
};
}  // namespace

// The fused FP64 marches (variants 2, 3, 4) store receiver increments; the running sum over time
// is taken once after the time loop (same order, same rounding as the reference's in-loop sum).
bool Model::receiver_increments() const { return variant_ >= 2; }

void Model::finish_receivers(int i_shot, cudaStream_t st) {
  const Config &c = cfg_;
  if (c.nr == 0 || c.nt == 0) return;
  const size_t shot_rows = (size_t)i_shot * c.nr * c.nt;
  receiver_prefix_kernel<<<blocks_for(2 * (size_t)c.nr, 64), 64, 0, st>>>(rtf_ux_.as<>() + shot_rows,
                                                                         rtf_uz_.as<>() + shot_rows, c.nr, c.nt);
  PSV_CUDA(cudaGetLastError());
}

// ---- CUDA-graph replay of a time loop (option "graph") ----
// Everything a captured launch bakes in (FusedCommon by value, the models pointer, the launch geometry) is a function
// of the values below; what the kernels read through FusedModel (state, medium, snapshot and seismogram pointers) is
// uploaded by fused_begin before every run and may change freely.
std::vector<unsigned long long> Model::graph_key(int direction, int i_shot, bool store) const {
  auto p = [](const void *q) { return (unsigned long long)(uintptr_t)q; };
  auto n = [](long long v) { return (unsigned long long)v; };
  const ShotPlan &plan = plans_[i_shot];
  return {n(direction > 0), n(i_shot), n(store), n(variant_), n(exact_), n(fp32_), n(fused_strip_), n(fused_block_),
          n(pipe_stages_), n(pipe_warps_), n(tuned_warps_), n(tuned_strip_), n(pdl_), n(l2_mode_), n(pair_warps_),
          n(pair_stages()), n(pair_strip_), n((long long)batch_followers_.size()), p(fused_model_.as<char>()),
          p(fused_model32_.as<char>()), p(fields_.as<char>()), n((long long)fields_.bytes()), p(fields32_.as<char>()),
          n((long long)fields32_.bytes()), p(taper_table_.as<char>()), p(rec_row_start_.as<char>()),
          p(rec_row_z_.as<char>()), p(rec_row_x_.as<char>()), p(rec_row_id_.as<char>()),
          p(plan.op_row_start.as<char>()), p(plan.row_ops.as<char>()), p(stf_.as<char>()), n(n_snap_alloc_),
          n(cfg_.nt), n(cfg_.snapshot_interval), n(cfg_.nr)};
}

// Replay pays when a step launch is shorter than the ~2.5 us a stream launch costs on top of a graph node: by default
// for launches of up to 2^20 grid points (all models of an ensemble counted), and only while the graphs of all shots
// and both directions fit the node budget of the cache.
namespace {
constexpr long long kGraphAutoPoints = 1 << 20;
constexpr long long kGraphNodeBudget = 400000;
}  // namespace

bool Model::graph_wanted() const {
  if (graph_ == 0 || variant_ < 2 || profile_ || slab_) return false;
  if (graph_ == 1) return true;
  const long long points = (long long)cfg_.nx * cfg_.nz * (1 + (long long)batch_followers_.size());
  return points <= kGraphAutoPoints && 2LL * cfg_.n_shots * (cfg_.nt + 2) <= kGraphNodeBudget;
}

void Model::drop_graphs() {
  for (auto &g : graphs_)
    if (g.second.exec) cudaGraphExecDestroy(g.second.exec);
  graphs_.clear();
  graph_seen_.clear();
}

void Model::graph_capture_begin() { PSV_CUDA(cudaStreamBeginCapture(stream_, cudaStreamCaptureModeRelaxed)); }

void Model::graph_capture_abort() {
  cudaGraph_t g = nullptr;
  cudaStreamEndCapture(stream_, &g);
  if (g) cudaGraphDestroy(g);
  cudaGetLastError();
}

const Model::LoopGraph &Model::graph_capture_end(const std::vector<unsigned long long> &key) {
  cudaGraph_t g = nullptr;
  PSV_CUDA(cudaStreamEndCapture(stream_, &g));
  LoopGraph lg;
  const cudaError_t e = cudaGraphInstantiate(&lg.exec, g, 0);
  cudaGraphDestroy(g);
  PSV_CUDA(e);
  lg.end_cur = cur_;
  lg.launches = stat_launches_;
  lg.pair_launches = stat_pair_launches_;
  long long nodes = lg.launches;  // one graph holds a node per time step
  for (const auto &o : graphs_) nodes += o.second.launches;
  if (nodes > kGraphNodeBudget) drop_graphs();
  return graphs_[key] = lg;
}

void Model::graph_replay(const LoopGraph &g) {
  PSV_CUDA(cudaGraphLaunch(g.exec, stream_));
  cur_ = g.end_cur;
  stat_launches_ = g.launches;
  stat_pair_launches_ = g.pair_launches;
}

// fdModel::forward_simulate (src/fdModel.cpp:526-787)
void Model::forward_simulate(int i_shot, bool store_fields, bool verbose, bool output_wavefields) {
  check_shot(i_shot);
  if (slab_ && (halo_lo_ || halo_hi_))
    throw state_error("This model is one slab of a split grid: drive it with psv_march_* and exchange halos "
                      "(psvwave_b200.SlabDomain does).");
  bind();
  const Config &c = cfg_;
  if (fp32_ && (variant_ < 2 || output_wavefields))
    throw invalid_argument("The FP32 wavefield mode needs a fused variant (>= 2) and no wavefield dumps.");
  if (store_fields) ensure_snapshots(i_shot);
  const auto wall0 = std::chrono::steady_clock::now();
  stat_launches_ = 0;
  stat_pair_launches_ = 0;
  stat_resident_ = 0;
  {  // grids that fit on chip: the whole loop in one launch (resident.cuh)
    int rows = 0, ctas = 0;
    size_t smem = 0;
    if (!output_wavefields && resident_plan(i_shot, false, &rows, &ctas, &smem)) {
      fused_begin(i_shot, POINT_FORWARD);
      zero_dynamic_fields();
      PSV_CUDA(cudaEventRecord(ev_begin_, stream_));
      run_resident(i_shot, false, store_fields, rows, ctas, smem);
      if (c.nr > 0) {
        receiver_prefix_models_kernel<<<dim3(blocks_for(2 * (size_t)c.nr, 64), 1), 64, 0, stream_>>>(
            fused_model_.as<FusedModel>(), c.nr, c.nt);
        ++stat_launches_;
      }
      PSV_CUDA(cudaEventRecord(ev_end_, stream_));
      PSV_CUDA(cudaGetLastError());
      resident_check();
      float rms = 0;
      PSV_CUDA(cudaEventElapsedTime(&rms, ev_begin_, ev_end_));
      stat_elapsed_ms_ = rms;
      if (verbose) {
        const double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - wall0).count();
        std::cout << "Seconds elapsed for forward wave simulation: " << secs << std::endl;
      }
      return;
    }
  }
  // option "graph": the loop of this (shot, store) is captured once, then replayed
  bool use_graph = graph_wanted() && !output_wavefields;
  const bool begun = use_graph;
  const LoopGraph *cached = nullptr;
  std::vector<unsigned long long> gkey;
  if (use_graph) {
    fused_begin(i_shot, POINT_FORWARD);  // synchronous upload: outside the capture
    gkey = graph_key(+1, i_shot, store_fields);
    auto f = graphs_.find(gkey);
    if (f != graphs_.end()) cached = &f->second;
    // auto mode: a loop seen for the first time is launched kernel by kernel (capture + instantiation cost about
    // a third of a small march); it is captured when it comes back
    else if (graph_ < 0 && graph_seen_[gkey]++ == 0) use_graph = false;
  }
  if (cached) {
    PSV_CUDA(cudaEventRecord(ev_begin_, stream_));
    graph_replay(*cached);
    PSV_CUDA(cudaEventRecord(ev_end_, stream_));
    PSV_CUDA(cudaStreamSynchronize(stream_));
    float gms = 0;
    PSV_CUDA(cudaEventElapsedTime(&gms, ev_begin_, ev_end_));
    stat_elapsed_ms_ = gms;
    if (verbose) {
      const double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - wall0).count();
      std::cout << "Seconds elapsed for forward wave simulation: " << secs << std::endl;
    }
    return;
  }
  if (!use_graph) tune_pipe_geometry(i_shot, POINT_FORWARD);
  if (use_graph) graph_capture_begin();
  try {
  zero_dynamic_fields();
  if (!use_graph) PSV_CUDA(cudaEventRecord(ev_begin_, stream_));
  Profiler prof(prof_events_, stream_, profile_, c.nt);

  const size_t shot_rows = (size_t)i_shot * c.nr * c.nt;
  RecordArgs ra{F(0), F(1), rec_index_.as<int>(), rtf_ux_.as<>() + shot_rows,
                rtf_uz_.as<>() + shot_rows, c.nr, c.nt, 0, c.dt, c.dx * c.dz};
  const ShotPlan &plan = plans_[i_shot];
  InjectArgs ia{F(0), F(1), b_.data(), stf_.as<>(), plan.ops.as<InjectOp>(),
                plan.heads.as<int>(), plan.n_heads, c.nt, 0, c.dt};

  if (variant_ >= 2 && !begun) fused_begin(i_shot, POINT_FORWARD);
  for (int it = 0; it < c.nt; ++it) {
    if (variant_ >= 2) {
      const bool snap = store_fields && it % c.snapshot_interval == 0;
      // variant 4 advances two steps per launch unless the second one has a snapshot to store
      const bool pair = variant_ == 4 && !fp32_ && !output_wavefields && it + 1 < c.nt &&
                        !(store_fields && (it + 1) % c.snapshot_interval == 0);
      const bool sample = prof.sample(snap || (variant_ == 4 && !pair));
      if (sample) prof.mark(0);
      launch_fused(c.dt, snap ? MODE_SNAPSHOT : MODE_PLAIN, POINT_FORWARD, i_shot, it / c.snapshot_interval, it,
                   pair ? 1 : 0);
      if (sample) {
        prof.mark(1);
        prof.mark(2);
        prof.next();
      }
      if (output_wavefields && it % 10 == 0) write_wavefield_files(it);
      if (pair) ++it;
      continue;
    }
    // the point kernels take the state pointers per step (the fused variant may have left the
    // current state in either buffer; variants 0/1 update it in place)
    ra.vx = F(0);
    ra.vz = F(1);
    ia.vx = F(0);
    ia.vz = F(1);
    if (c.nr > 0) {
      ra.it = it;
      record_receivers_kernel<<<blocks_for(c.nr, 128), 128, 0, stream_>>>(ra);
      ++stat_launches_;
    }
    const bool snap = store_fields && it % c.snapshot_interval == 0;
    const bool sample = prof.sample(snap);
    if (sample) prof.mark(0);
    launch_stress(c.dt, snap ? MODE_SNAPSHOT : MODE_PLAIN, i_shot, it / c.snapshot_interval);
    if (sample) prof.mark(1);
    launch_velocity(c.dt);
    if (sample) {
      prof.mark(2);
      prof.next();
    }
    if (plan.n_heads > 0) {
      ia.it = it;
      inject_sources_kernel<<<blocks_for(plan.n_heads, 64), 64, 0, stream_>>>(ia);
      ++stat_launches_;
    }
    if (output_wavefields && it % 10 == 0) write_wavefield_files(it);
  }
  if (receiver_increments() && c.nr > 0) {  // running sum of the stored increments, all models of a batch at once
    const int n_models = 1 + (int)batch_followers_.size();
    receiver_prefix_models_kernel<<<dim3(blocks_for(2 * (size_t)c.nr, 64), n_models), 64, 0, stream_>>>(
        fused_model_.as<FusedModel>(), c.nr, c.nt);
    PSV_CUDA(cudaGetLastError());
    ++stat_launches_;
  }
  if (use_graph) {
    const LoopGraph &lg = graph_capture_end(gkey);
    PSV_CUDA(cudaEventRecord(ev_begin_, stream_));
    graph_replay(lg);
  }
  PSV_CUDA(cudaEventRecord(ev_end_, stream_));
  PSV_CUDA(cudaGetLastError());
  PSV_CUDA(cudaStreamSynchronize(stream_));
  float ms = 0;
  PSV_CUDA(cudaEventElapsedTime(&ms, ev_begin_, ev_end_));
  stat_elapsed_ms_ = ms;
  prof.collect(stat_stress_ms_, stat_velocity_ms_);
  } catch (...) {
    if (use_graph) graph_capture_abort();
    throw;
  }
  if (verbose) {
    const double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - wall0).count();
    std::cout << "Seconds elapsed for forward wave simulation: " << secs << std::endl;
  }
}

// fdModel::adjoint_simulate (src/fdModel.cpp:789-969)
void Model::adjoint_simulate(int i_shot, bool verbose) {
  check_shot(i_shot);
  if (slab_ && (halo_lo_ || halo_hi_))
    throw state_error("This model is one slab of a split grid: drive it with psv_march_* and exchange halos "
                      "(psvwave_b200.SlabDomain does).");
  bind();
  const Config &c = cfg_;
  if (snapshots_[i_shot].empty())
    throw state_error("adjoint_simulate needs the forward wavefield of this shot: call "
                      "forward_simulate(i_shot, store_fields=True) first.");
  if (fp32_ && variant_ < 2) throw invalid_argument("The FP32 wavefield mode needs a fused variant (>= 2).");
  const auto wall0 = std::chrono::steady_clock::now();
  stat_launches_ = 0;
  stat_pair_launches_ = 0;
  stat_resident_ = 0;
  {  // grids that fit on chip: the whole loop in one launch (resident.cuh)
    int rows = 0, ctas = 0;
    size_t smem = 0;
    if (resident_plan(i_shot, true, &rows, &ctas, &smem)) {
      fused_begin(i_shot, POINT_ADJOINT);
      zero_dynamic_fields();
      PSV_CUDA(cudaEventRecord(ev_begin_, stream_));
      run_resident(i_shot, true, false, rows, ctas, smem);
      PSV_CUDA(cudaEventRecord(ev_end_, stream_));
      PSV_CUDA(cudaGetLastError());
      resident_check();
      float rms = 0;
      PSV_CUDA(cudaEventElapsedTime(&rms, ev_begin_, ev_end_));
      stat_elapsed_ms_ = rms;
      if (verbose) {
        const double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - wall0).count();
        std::cout << "Seconds elapsed for adjoint wave simulation: " << secs << std::endl;
      }
      return;
    }
  }
  bool use_graph = graph_wanted();
  const bool begun = use_graph;
  const LoopGraph *cached = nullptr;
  std::vector<unsigned long long> gkey;
  if (use_graph) {
    fused_begin(i_shot, POINT_ADJOINT);  // synchronous upload: outside the capture
    gkey = graph_key(-1, i_shot, true);
    gkey.push_back((unsigned long long)(uintptr_t)snapshots_[i_shot].as<char>());
    auto f = graphs_.find(gkey);
    if (f != graphs_.end()) cached = &f->second;
    else if (graph_ < 0 && graph_seen_[gkey]++ == 0) use_graph = false;  // as in forward_simulate
  }
  if (cached) {
    PSV_CUDA(cudaEventRecord(ev_begin_, stream_));
    graph_replay(*cached);
    PSV_CUDA(cudaEventRecord(ev_end_, stream_));
    PSV_CUDA(cudaStreamSynchronize(stream_));
    float gms = 0;
    PSV_CUDA(cudaEventElapsedTime(&gms, ev_begin_, ev_end_));
    stat_elapsed_ms_ = gms;
    if (verbose) {
      const double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - wall0).count();
      std::cout << "Seconds elapsed for adjoint wave simulation: " << secs << std::endl;
    }
    return;
  }
  if (!use_graph) tune_pipe_geometry(i_shot, c.nr > 0 ? POINT_ADJOINT : POINT_NONE);
  if (use_graph) graph_capture_begin();
  try {
  zero_dynamic_fields();
  if (!use_graph) PSV_CUDA(cudaEventRecord(ev_begin_, stream_));
  Profiler prof(prof_events_, stream_, profile_, c.nt);

  const size_t shot_rows = (size_t)i_shot * c.nr * c.nt;
  AdjointInjectArgs aa{F(0), F(1), b_.data(), a_stf_ux_.as<>() + shot_rows,
                       a_stf_uz_.as<>() + shot_rows, rec_index_.as<int>(), rec_next_.as<int>(),
                       rec_heads_.as<int>(), n_rec_heads_, c.nt, 0, c.dt, c.dx * c.dz};
  if (variant_ >= 2 && !begun) fused_begin(i_shot, POINT_ADJOINT);
  for (int it = c.nt - 1; it >= 0; --it) {
    const bool corr = it % c.snapshot_interval == 0;
    // variant 4 retreats two steps per launch unless the second one correlates with a snapshot
    const bool pair = variant_ == 4 && !fp32_ && it - 1 >= 0 && (it - 1) % c.snapshot_interval != 0;
    const bool sample = prof.sample(corr || (variant_ == 4 && !pair));
    if (variant_ >= 2) {
      if (sample) prof.mark(0);
      launch_fused(-c.dt, corr ? MODE_ACCUMULATE : MODE_PLAIN, c.nr > 0 ? POINT_ADJOINT : POINT_NONE, i_shot,
                   it / c.snapshot_interval, it, pair ? -1 : 0);
      if (sample) {
        prof.mark(1);
        prof.mark(2);
        prof.next();
      }
      if (pair) --it;
      continue;
    }
    if (sample) prof.mark(0);
    launch_stress(-c.dt, corr ? MODE_ACCUMULATE : MODE_PLAIN, i_shot, it / c.snapshot_interval);
    if (sample) prof.mark(1);
    launch_velocity(-c.dt);
    if (sample) {
      prof.mark(2);
      prof.next();
    }
    if (n_rec_heads_ > 0) {
      aa.it = it;
      inject_adjoint_kernel<<<blocks_for(n_rec_heads_, 128), 128, 0, stream_>>>(aa);
      ++stat_launches_;
    }
  }
  if (use_graph) {
    const LoopGraph &lg = graph_capture_end(gkey);
    PSV_CUDA(cudaEventRecord(ev_begin_, stream_));
    graph_replay(lg);
  }
  PSV_CUDA(cudaEventRecord(ev_end_, stream_));
  PSV_CUDA(cudaGetLastError());
  PSV_CUDA(cudaStreamSynchronize(stream_));
  float ms = 0;
  PSV_CUDA(cudaEventElapsedTime(&ms, ev_begin_, ev_end_));
  stat_elapsed_ms_ = ms;
  prof.collect(stat_stress_ms_, stat_velocity_ms_);
  } catch (...) {
    if (use_graph) graph_capture_abort();
    throw;
  }
  if (verbose) {
    const double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - wall0).count();
    std::cout << "Seconds elapsed for adjoint wave simulation: " << secs << std::endl;
  }
}

// ---- domain split: the time loop in pieces + halo staging (SURVEY 8e "large single grids") ----
// One fused step consumes 4 halo rows (new velocities of row x need the old state of rows x-4..x+4),
// so a halo of H rows lets H/4 steps run between two exchanges; the rows a step invalidates lie in
// the halo and are overwritten by the next exchange.  Snapshots, correlation and receiver samples
// are pointwise on rows the slab owns, which are always valid.
void Model::march_begin(int i_shot, bool adjoint, bool store_fields) {
  check_shot(i_shot);
  bind();
  if (variant_ != 2 && variant_ != 3)
    throw invalid_argument("psv_march_* needs the fused variant 2 or 3.");
  if (march_.active) throw state_error("A march is already in progress on this model.");
  if (adjoint && snapshots_[i_shot].empty())
    throw state_error("The adjoint march needs the forward wavefield of this shot: run the stored forward march first.");
  if (!adjoint && store_fields) ensure_snapshots(i_shot);
  stat_launches_ = 0;
  stat_pair_launches_ = 0;
  tune_pipe_geometry(i_shot, adjoint ? (cfg_.nr > 0 ? POINT_ADJOINT : POINT_NONE) : POINT_FORWARD);
  zero_dynamic_fields();
  PSV_CUDA(cudaEventRecord(ev_begin_, stream_));
  fused_begin(i_shot, adjoint ? POINT_ADJOINT : POINT_FORWARD);
  march_.active = true;
  march_.adjoint = adjoint;
  march_.store = store_fields;
  march_.i_shot = i_shot;
  march_.it = adjoint ? cfg_.nt - 1 : 0;
}

int Model::march_steps(int n) {
  if (!march_.active) throw state_error("psv_march_steps without psv_march_begin.");
  bind();
  const Config &c = cfg_;
  int done = 0;
  for (; done < n; ++done) {
    const int it = march_.it;
    if (march_.adjoint) {
      if (it < 0) break;
      const bool corr = it % c.snapshot_interval == 0;
      launch_fused(-c.dt, corr ? MODE_ACCUMULATE : MODE_PLAIN, c.nr > 0 ? POINT_ADJOINT : POINT_NONE, march_.i_shot,
                   it / c.snapshot_interval, it, 0);
      --march_.it;
    } else {
      if (it >= c.nt) break;
      const bool snap = march_.store && it % c.snapshot_interval == 0;
      launch_fused(c.dt, snap ? MODE_SNAPSHOT : MODE_PLAIN, POINT_FORWARD, march_.i_shot, it / c.snapshot_interval,
                   it, 0);
      ++march_.it;
    }
  }
  PSV_CUDA(cudaGetLastError());
  return done;
}

void Model::march_end() {
  if (!march_.active) throw state_error("psv_march_end without psv_march_begin.");
  bind();
  march_.active = false;
  if (!march_.adjoint && receiver_increments()) {
    finish_receivers(march_.i_shot, stream_);
    ++stat_launches_;
  }
  PSV_CUDA(cudaEventRecord(ev_end_, stream_));
  PSV_CUDA(cudaStreamSynchronize(stream_));
  float ms = 0;
  PSV_CUDA(cudaEventElapsedTime(&ms, ev_begin_, ev_end_));
  stat_elapsed_ms_ = ms;
}

namespace {
// rows [row0, row0 + rows) of the 5 state arrays <-> one contiguous staging block per side.  Everything in BYTES
// (multiples of 16): the state is double, or float in the FP32 wavefield mode.
struct HaloArgs {
  char *state;  // vx of the current buffer; field k at + k * field_stride
  size_t field_stride;
  char *staging;  // [side][field][chunk]
  size_t chunk;   // rows * ld * sizeof(element)
  size_t row_off[2];  // first byte of the rows of side 0 / 1 inside a field
  int active[2];
};
template <bool PACK>
__global__ void halo_copy_kernel(const HaloArgs a) {
  const int side = blockIdx.y / 5, field = blockIdx.y % 5;
  if (!a.active[side]) return;
  int4 *grid = reinterpret_cast<int4 *>(a.state + field * a.field_stride + a.row_off[side]);
  int4 *stage = reinterpret_cast<int4 *>(a.staging + ((size_t)side * 5 + field) * a.chunk);
  const size_t n = a.chunk >> 4;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    if (PACK) stage[i] = grid[i];
    else grid[i] = stage[i];
  }
}
}  // namespace

void Model::halo_pack() {
  if (!slab_) throw state_error("Not a slab of a split grid.");
  bind();
  const int H = std::max(halo_lo_, halo_hi_);
  if (H == 0) return;
  const size_t es = fp32_ ? sizeof(float) : sizeof(double);
  char *state = fp32_ ? reinterpret_cast<char *>(field32_ptr(cur_, 0)) : reinterpret_cast<char *>(F(0));
  HaloArgs a{state, field_stride_ * es, halo_send_.as<char>(), (size_t)H * ld_ * es,
             {(size_t)halo_lo_ * ld_ * es, (size_t)(cfg_.nx - halo_hi_ - H) * ld_ * es}, {halo_lo_ > 0, halo_hi_ > 0}};
  halo_copy_kernel<true><<<dim3(64, 10), 256, 0, stream_>>>(a);
  PSV_CUDA(cudaGetLastError());
  ++stat_launches_;
}

void Model::halo_unpack() {
  if (!slab_) throw state_error("Not a slab of a split grid.");
  bind();
  const int H = std::max(halo_lo_, halo_hi_);
  if (H == 0) return;
  const size_t es = fp32_ ? sizeof(float) : sizeof(double);
  char *state = fp32_ ? reinterpret_cast<char *>(field32_ptr(cur_, 0)) : reinterpret_cast<char *>(F(0));
  HaloArgs a{state, field_stride_ * es, halo_recv_.as<char>(), (size_t)H * ld_ * es,
             {(size_t)0, (size_t)(cfg_.nx - H) * ld_ * es}, {halo_lo_ > 0, halo_hi_ > 0}};
  halo_copy_kernel<false><<<dim3(64, 10), 256, 0, stream_>>>(a);
  PSV_CUDA(cudaGetLastError());
  ++stat_launches_;
}

// ---- the slabs of one process exchange halos peer to peer ----
namespace {
struct PullArgs {  // bytes throughout (the state is double, or float in the FP32 wavefield mode)
  char *dst;         // my vx (current buffer); field k at + k * dst_stride
  size_t dst_stride;
  const char *src[2];  // the neighbour's vx (current buffer) per side, or null
  size_t src_stride[2];
  size_t dst_off[2], src_off[2];  // first byte of the H rows inside a field
  size_t chunk;                   // H * ld * sizeof(element)
};
__global__ void halo_pull_kernel(const PullArgs a) {
  const int side = blockIdx.y / 5, field = blockIdx.y % 5;
  if (!a.src[side]) return;
  const int4 *src = reinterpret_cast<const int4 *>(a.src[side] + field * a.src_stride[side] + a.src_off[side]);
  int4 *dst = reinterpret_cast<int4 *>(a.dst + field * a.dst_stride + a.dst_off[side]);
  const size_t n = a.chunk >> 4;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    dst[i] = src[i];  // loads cross NVLink when the neighbour lives on another GPU (peer access)
}
}  // namespace

void Model::pull_halos(const Model *lower, const Model *upper) {
  bind();
  const int H = std::max(halo_lo_, halo_hi_);
  if (H == 0) return;
  const Model *nb[2] = {halo_lo_ ? lower : nullptr, halo_hi_ ? upper : nullptr};
  bool direct = true;
  for (const Model *o : nb)
    if (o && o->device_ != device_) {
      int can = 0;
      PSV_CUDA(cudaDeviceCanAccessPeer(&can, device_, o->device_));
      if (!can) {
        direct = false;
        continue;
      }
      if (!slab_peer_checked_) {
        const cudaError_t e = cudaDeviceEnablePeerAccess(o->device_, 0);
        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) PSV_CUDA(e);
        cudaGetLastError();
      }
    }
  slab_peer_checked_ = true;
  // my rows [0, H) <- the lower neighbour's last H owned rows; my rows [nx-H, nx) <- the upper one's first H owned rows
  const size_t es = fp32_ ? sizeof(float) : sizeof(double);
  auto state = [es](const Model *m) {
    return m->fp32_ ? reinterpret_cast<char *>(m->field32_ptr(m->cur_, 0)) : reinterpret_cast<char *>(m->F(0));
  };
  for (const Model *o : nb)
    if (o && o->fp32_ != fp32_) throw state_error("The slabs of one grid must agree on the FP32 wavefield mode.");
  const size_t chunk = (size_t)H * ld_ * es;
  const size_t dst_off[2] = {0, (size_t)(cfg_.nx - H) * ld_ * es};
  size_t src_off[2] = {0, 0};
  if (nb[0]) src_off[0] = (size_t)(nb[0]->cfg_.nx - nb[0]->halo_hi_ - H) * ld_ * es;
  if (nb[1]) src_off[1] = (size_t)nb[1]->halo_lo_ * ld_ * es;
  if (direct) {
    PullArgs a{state(this), field_stride_ * es, {nb[0] ? state(nb[0]) : nullptr, nb[1] ? state(nb[1]) : nullptr},
               {nb[0] ? nb[0]->field_stride_ * es : 0, nb[1] ? nb[1]->field_stride_ * es : 0},
               {dst_off[0], dst_off[1]}, {src_off[0], src_off[1]}, chunk};
    halo_pull_kernel<<<dim3(32, 10), 256, 0, stream_>>>(a);
    PSV_CUDA(cudaGetLastError());
    ++stat_launches_;
    return;
  }
  for (int side = 0; side < 2; ++side)  // no peer mapping between the two GPUs: the copy engine routes it
    if (nb[side])
      for (int k = 0; k < 5; ++k)
        PSV_CUDA(cudaMemcpyPeerAsync(state(this) + k * field_stride_ * es + dst_off[side], device_,
                                     state(nb[side]) + k * nb[side]->field_stride_ * es + src_off[side],
                                     nb[side]->device_, chunk, stream_));
}

void Model::march_slabs(Model *const *slabs, int n, int i_shot, bool adjoint, bool store_fields) {
  if (n < 1 || !slabs) throw invalid_argument("psv_march_slabs needs the slabs of one grid.");
  const int H = n > 1 ? std::max(slabs[0]->halo_lo_, slabs[0]->halo_hi_) : 4;
  int next_row = 0;
  for (int i = 0; i < n; ++i) {
    Model *m = slabs[i];
    if (!m->slab_) throw state_error("psv_march_slabs: not a slab of a split grid.");
    const bool lo = i > 0, hi = i < n - 1;
    if ((m->halo_lo_ > 0) != lo || (m->halo_hi_ > 0) != hi || (lo && m->halo_lo_ != H) || (hi && m->halo_hi_ != H) ||
        m->nx_global_ != slabs[0]->nx_global_ || m->ld_ != slabs[0]->ld_ || m->x_off_ + m->halo_lo_ != next_row)
      throw invalid_argument("psv_march_slabs: pass ALL slabs of one grid, in slab order, with one halo width.");
    next_row = m->x_off_ + m->cfg_.nx - m->halo_hi_;
    for (int p = 0; p < 2; ++p) {
      m->bind();
      if (!m->slab_done_[p]) PSV_CUDA(cudaEventCreateWithFlags(&m->slab_done_[p], cudaEventDisableTiming));
      if (!m->slab_pulled_[p]) PSV_CUDA(cudaEventCreateWithFlags(&m->slab_pulled_[p], cudaEventDisableTiming));
    }
  }
  if (next_row != slabs[0]->nx_global_) throw invalid_argument("psv_march_slabs: the slabs do not cover the grid.");
  int begun = 0;
  try {
    for (; begun < n; ++begun) slabs[begun]->march_begin(i_shot, adjoint, store_fields);
    const int k = H / 4;  // one step consumes 4 halo rows
    int remaining = slabs[0]->cfg_.nt;
    for (int round = 0; remaining > 0; ++round) {
      const int steps = std::min(k, remaining), p = round & 1;
      for (int i = 0; i < n; ++i) {
        Model *m = slabs[i];
        m->bind();
        // The neighbours' pulls of the last exchange read my edge rows in the buffer my SECOND step of this round
        // overwrites (ping-pong state): the first step runs while their "pulled" events travel.  (A one-step round
        // needs no wait at all: my next "done" wait already orders me behind their pulls.)
        int done = m->march_steps(1);
        if (round > 0 && steps >= 2) {
          if (i > 0) PSV_CUDA(cudaStreamWaitEvent(m->stream_, slabs[i - 1]->slab_pulled_[p ^ 1], 0));
          if (i < n - 1) PSV_CUDA(cudaStreamWaitEvent(m->stream_, slabs[i + 1]->slab_pulled_[p ^ 1], 0));
        }
        if (steps >= 2) done += m->march_steps(steps - 1);
        if (done != steps) throw state_error("psv_march_slabs: the slabs march out of step.");
        PSV_CUDA(cudaEventRecord(m->slab_done_[p], m->stream_));
      }
      remaining -= steps;
      if (remaining <= 0 || n == 1) continue;
      for (int i = 0; i < n; ++i) {
        Model *m = slabs[i];
        m->bind();
        if (i > 0) PSV_CUDA(cudaStreamWaitEvent(m->stream_, slabs[i - 1]->slab_done_[p], 0));
        if (i < n - 1) PSV_CUDA(cudaStreamWaitEvent(m->stream_, slabs[i + 1]->slab_done_[p], 0));
        m->pull_halos(i > 0 ? slabs[i - 1] : nullptr, i < n - 1 ? slabs[i + 1] : nullptr);
        PSV_CUDA(cudaEventRecord(m->slab_pulled_[p], m->stream_));
      }
    }
  } catch (...) {
    for (int i = 0; i < begun; ++i) {  // leave no slab with a march in progress
      try {
        slabs[i]->march_.active = false;
        slabs[i]->synchronize();
      } catch (...) {
      }
    }
    throw;
  }
  for (int i = 0; i < n; ++i) slabs[i]->march_end();
}

void Model::halo_buffers(void **send_lo, void **send_hi, void **recv_lo, void **recv_hi, size_t *count) const {
  // one side = 5 fields x halo rows of the state's element type; `count` is its length in doubles
  const size_t side_bytes = 5 * (size_t)std::max(halo_lo_, halo_hi_) * ld_ * (fp32_ ? sizeof(float) : sizeof(double));
  *send_lo = halo_lo_ ? halo_send_.as<char>() : nullptr;
  *send_hi = halo_hi_ ? halo_send_.as<char>() + side_bytes : nullptr;
  *recv_lo = halo_lo_ ? halo_recv_.as<char>() : nullptr;
  *recv_hi = halo_hi_ ? halo_recv_.as<char>() + side_bytes : nullptr;
  *count = side_bytes / sizeof(double);
}

void Model::slab_info(long long out[7]) const {
  out[0] = nx_global_;
  out[1] = x_off_;
  out[2] = cfg_.nx;
  out[3] = halo_lo_;
  out[4] = halo_hi_;
  out[5] = ld_;
  out[6] = slab_ ? 1 : 0;
}

// ---- Monte-Carlo ensembles: n models of one experiment, one launch per time step ---------------
// The reference runs such ensembles as independent fdModel copies (copy ctor src/fdModel.cpp:64-97,
// SURVEY 8a20/8e); here they share the launches, nothing else.
void Model::check_batch(Model *const *models, int n, int i_shot) {
  if (!models || n < 1) throw invalid_argument("Batched simulate needs at least one model.");
  for (int i = 0; i < n; ++i)
    if (!models[i]) throw invalid_argument("Null model in the batch.");
  const Model &a = *models[0];
  a.check_shot(i_shot);
  if (a.variant_ < 2) throw invalid_argument("Batched simulate needs a fused variant (variant >= 2).");
  for (int i = 1; i < n; ++i) {
    const Model &b = *models[i];
    for (int k = 0; k < i; ++k)
      if (models[k] == models[i]) throw invalid_argument("The same model appears twice in the batch.");
    const Config &p = a.cfg_, &q = b.cfg_;
    const bool same = b.device_ == a.device_ && b.fp32_ == a.fp32_ && p.nx == q.nx && p.nz == q.nz && p.nt == q.nt && p.dx == q.dx &&
                      p.dz == q.dz && p.dt == q.dt && p.np_boundary == q.np_boundary && p.np_factor == q.np_factor &&
                      p.nx_inner_boundary == q.nx_inner_boundary && p.nz_inner_boundary == q.nz_inner_boundary &&
                      p.snapshot_interval == q.snapshot_interval && p.peak_frequency == q.peak_frequency &&
                      p.delay_cycles_per_shot == q.delay_cycles_per_shot && p.ix_sources == q.ix_sources &&
                      p.iz_sources == q.iz_sources && p.moment_angles == q.moment_angles &&
                      p.which_source_to_fire_in_which_shot == q.which_source_to_fire_in_which_shot &&
                      p.ix_receivers == q.ix_receivers && p.iz_receivers == q.iz_receivers;
    if (!same)
      throw invalid_argument("Batched simulate needs models of one experiment on one device: same grid, time axis, "
                             "boundary, sources, receivers and snapshot interval.");
  }
}

void Model::adopt_batch_state(const Model &leader) {
  cur_ = leader.cur_;
  stat_launches_ = leader.stat_launches_;
  stat_pair_launches_ = leader.stat_pair_launches_;
  stat_elapsed_ms_ = leader.stat_elapsed_ms_;
  stat_stress_ms_ = leader.stat_stress_ms_;
  stat_velocity_ms_ = leader.stat_velocity_ms_;
}

void Model::forward_batch(Model *const *models, int n, int i_shot, bool store_fields) {
  check_batch(models, n, i_shot);
  Model &lead = *models[0];
  lead.bind();
  for (int i = 1; i < n; ++i) {  // the followers' own streams hand over to the leader's
    Model &f = *models[i];
    if (store_fields) f.ensure_snapshots(i_shot);
    f.zero_dynamic_fields();
    PSV_CUDA(cudaStreamSynchronize(f.stream_));
  }
  lead.batch_followers_.assign(models + 1, models + n);
  try {
    lead.forward_simulate(i_shot, store_fields, false, false);
  } catch (...) {
    lead.batch_followers_.clear();
    throw;
  }
  lead.batch_followers_.clear();
  for (int i = 1; i < n; ++i) models[i]->adopt_batch_state(lead);
}

void Model::adjoint_batch(Model *const *models, int n, int i_shot) {
  check_batch(models, n, i_shot);
  Model &lead = *models[0];
  lead.bind();
  for (int i = 1; i < n; ++i) {
    Model &f = *models[i];
    if (f.snapshots_[i_shot].empty())
      throw state_error("adjoint_simulate needs the forward wavefield of this shot in every model of the batch: "
                        "run the stored forward simulation first.");
    f.zero_dynamic_fields();
    PSV_CUDA(cudaStreamSynchronize(f.stream_));
  }
  lead.batch_followers_.assign(models + 1, models + n);
  try {
    lead.adjoint_simulate(i_shot, false);
  } catch (...) {
    lead.batch_followers_.clear();
    throw;
  }
  lead.batch_followers_.clear();
  for (int i = 1; i < n; ++i) models[i]->adopt_batch_state(lead);
}

// fdModel::run_model (src/fdModel.cpp:1291-1308)
void Model::run_model(bool verbose, bool simulate_adjoint) {
  for (int s = 0; s < cfg_.n_shots; ++s) forward_simulate(s, true, verbose, false);
  calculate_l2_misfit();
  if (simulate_adjoint) {
    calculate_l2_adjoint_sources();
    reset_kernels();
    for (int s = 0; s < cfg_.n_shots; ++s) adjoint_simulate(s, verbose);
    map_kernels_to_velocity();
  }
}

// ---------------------------------------------------------------------------------------------
// point-wise stages
// ---------------------------------------------------------------------------------------------
namespace {
VectorArgs vector_args(const Config &c, int ld);
}  // namespace

// Per-shot gradient loop: forward(store) -> misfit of the shot -> adjoint sources ->
// adjoint -> (drop the shot's snapshots), then the map.  Same kernels and accumulation order across
// shots as run_model (src/fdModel.cpp:1291-1308), but only ONE shot's snapshots are resident at a
// time.  The misfit is the sum of the per-shot misfits (the reference sums shot-major as well).
double Model::run_gradient(bool release_each_shot) {
  reset_kernels();
  double total = 0.0;
  for (int s = 0; s < cfg_.n_shots; ++s) {
    forward_simulate(s, true, false, false);
    total += calculate_l2_misfit_shot(s);
    calculate_l2_adjoint_sources();
    adjoint_simulate(s, false);
    if (release_each_shot) release_snapshots(s);
  }
  map_kernels_to_velocity();
  misfit_ = total;
  return total;
}

// Device-resident steepest descent (tests/python/test_descent.py:95-129 of the
// reference): m = get_model_vector() once, then per round  g = get_gradient_vector();  m -= step * g;
// set_model_vector(m)  -- all three on the device, m and g never visit the host.
void Model::descent_load() {  // m <- block means of (vp, vs, rho)
  bind();
  if (slab_) throw state_error("Model vectors span the whole grid: not available on a slab.");
  const VectorArgs a = vector_args(cfg_, ld_);
  const size_t n = (size_t)cfg_.free_parameters;
  if (descent_vec_.bytes() < 2 * n * sizeof(double)) descent_vec_.allocate(2 * n * sizeof(double));
  vector_kernel<VEC_MEAN><<<blocks_for(a.per, 128), 128, 0, stream_>>>(a, vp_.data(), vs_.data(), rho_.data(),
                                                                       descent_vec_.as<>());
  PSV_CUDA(cudaGetLastError());
}

void Model::descent_apply(double step) {  // g <- block sums of the kernels; m -= step * g; fields <- m
  const VectorArgs a = vector_args(cfg_, ld_);
  const size_t n = (size_t)cfg_.free_parameters;
  double *m = descent_vec_.as<>(), *g = m + n;
  vector_kernel<VEC_SUM><<<blocks_for(a.per, 128), 128, 0, stream_>>>(a, vp_kernel_.data(), vs_kernel_.data(),
                                                                      density_v_kernel_.data(), g);
  descent_axpy_kernel<<<blocks_for(n, 256), 256, 0, stream_>>>(m, g, step, n);
  vector_kernel<VEC_FILL><<<blocks_for(a.per, 128), 128, 0, stream_>>>(a, vp_.data(), vs_.data(), rho_.data(), m);
  PSV_CUDA(cudaGetLastError());
  update_from_velocity();  // src/fdModel.cpp:1456
}

void Model::descent_step(double step) {
  descent_load();
  descent_apply(step);
}

// `iterations` rounds of { m -= step * g; set; gradient of all shots } without leaving the device;
// the kernels of the current model must exist (run_model / run_gradient) before the first round.
void Model::gradient_descent(int iterations, double step, bool release_each_shot, double *misfits) {
  if (iterations < 0) throw invalid_argument("iterations must be >= 0");
  descent_load();
  for (int i = 0; i < iterations; ++i) {
    descent_apply(step);
    const double m = run_gradient(release_each_shot);
    if (misfits) misfits[i] = m;
  }
}

// ---- device-side vectors, smoothing, L-BFGS ----
void Model::vector_from_fields(double *dev_m) {
  bind();
  if (slab_) throw state_error("Model vectors span the whole grid: not available on a slab.");
  const VectorArgs a = vector_args(cfg_, ld_);
  vector_kernel<VEC_MEAN><<<blocks_for(a.per, 128), 128, 0, stream_>>>(a, vp_.data(), vs_.data(), rho_.data(), dev_m);
  PSV_CUDA(cudaGetLastError());
}

void Model::fields_from_vector(const double *dev_m) {
  bind();
  if (slab_) throw state_error("Model vectors span the whole grid: not available on a slab.");
  const VectorArgs a = vector_args(cfg_, ld_);
  vector_kernel<VEC_FILL><<<blocks_for(a.per, 128), 128, 0, stream_>>>(a, vp_.data(), vs_.data(), rho_.data(),
                                                                       const_cast<double *>(dev_m));
  PSV_CUDA(cudaGetLastError());
  update_from_velocity();  // src/fdModel.cpp:1456
}

void Model::gradient_to_vector(double *dev_g, double sigma) {
  bind();
  if (slab_) throw state_error("Gradient vectors span the whole grid: not available on a slab.");
  const VectorArgs a = vector_args(cfg_, ld_);
  vector_kernel<VEC_SUM><<<blocks_for(a.per, 128), 128, 0, stream_>>>(a, vp_kernel_.data(), vs_kernel_.data(),
                                                                      density_v_kernel_.data(), dev_g);
  PSV_CUDA(cudaGetLastError());
  if (sigma > 0.0) {
    // the reference's _fwi class reshapes every third of the vector to (nz_free, nx_free) -- entries are x fastest
    // (src/fdModel.cpp:1501-1509) -- and filters it (docs-source/source/typical-workflow.rst:61-69)
    const size_t n = (size_t)cfg_.free_parameters;
    if (smooth_tmp_.bytes() < n * sizeof(double)) smooth_tmp_.allocate(n * sizeof(double));
    gaussian_smooth_panels(dev_g, smooth_tmp_.as<>(), 3, a.nbz, a.nbx, sigma, stream_);
  }
}

void Model::get_gradient_vector_smoothed(double *out, size_t n, double sigma) {
  bind();
  if (n != (size_t)cfg_.free_parameters) throw invalid_argument("Gradient vector has the wrong length.");
  if (sigma < 0.0) throw invalid_argument("Smoothing sigma must be >= 0.");
  if (scratch_.bytes() < n * sizeof(double)) scratch_.allocate(n * sizeof(double));
  gradient_to_vector(scratch_.as<>(), sigma);
  PSV_CUDA(cudaMemcpyAsync(out, scratch_.as<>(), n * sizeof(double), cudaMemcpyDeviceToHost, stream_));
  PSV_CUDA(cudaStreamSynchronize(stream_));
}

LbfgsResult Model::lbfgs(const LbfgsOptions &opt, double sigma, bool release_each_shot, double *misfits) {
  bind();
  if (sigma < 0.0) throw invalid_argument("Smoothing sigma must be >= 0.");
  const size_t n = (size_t)cfg_.free_parameters;
  DeviceBuffer x(n * sizeof(double));
  vector_from_fields(x.as<>());
  const LbfgsResult r = lbfgs_minimize(
      x.as<>(), n,
      [&](const double *m, double *g) {
        fields_from_vector(m);
        const double f = run_gradient(release_each_shot);
        gradient_to_vector(g, sigma);
        return f;
      },
      opt, misfits, stream_);
  fields_from_vector(x.as<>());  // the line search may have ended on a rejected trial point
  misfit_ = r.misfit;
  return r;
}

void Model::update_from_velocity() {
  bind();
  update_from_velocity_kernel<<<blocks_for(grid_elems_, 256), 256, 0, stream_>>>(
      vp_.data(), vs_.data(), rho_.data(), lm_.data(), mu_.data(), b_.data(), grid_elems_);
  PSV_CUDA(cudaGetLastError());
  PSV_CUDA(cudaStreamSynchronize(stream_));
  if (fp32_) refresh_medium32();
}

double Model::misfit_of_range(size_t first, size_t n) {
  bind();
  const size_t per_block = (size_t)kMisfitThreads * kMisfitPerThread;
  const unsigned nb = (unsigned)std::max<size_t>(1, (n + per_block - 1) / per_block);
  if (scratch_.bytes() < nb * sizeof(double)) scratch_.allocate(nb * sizeof(double));
  misfit_partial_kernel<<<nb, kMisfitThreads, 0, stream_>>>(rtf_ux_.as<>() + first, rtf_uz_.as<>() + first,
                                                            obs_ux_.as<>() + first, obs_uz_.as<>() + first, n,
                                                            0.5 * cfg_.dt, scratch_.as<>());
  PSV_CUDA(cudaGetLastError());
  std::vector<double> partial(nb);
  PSV_CUDA(cudaMemcpyAsync(partial.data(), scratch_.as<>(), nb * sizeof(double), cudaMemcpyDeviceToHost, stream_));
  PSV_CUDA(cudaStreamSynchronize(stream_));
  double acc = 0.0;
  for (double p : partial) acc += p;
  return acc;
}

double Model::calculate_l2_misfit() {
  misfit_ = misfit_of_range(0, (size_t)cfg_.n_shots * cfg_.nr * cfg_.nt);
  return misfit_;
}

double Model::calculate_l2_misfit_shot(int i_shot) {
  check_shot(i_shot);
  const size_t per_shot = (size_t)cfg_.nr * cfg_.nt;
  return misfit_of_range((size_t)i_shot * per_shot, per_shot);
}

void Model::calculate_l2_adjoint_sources() {
  bind();
  const size_t n = (size_t)cfg_.n_shots * cfg_.nr * cfg_.nt;
  if (n == 0) return;
  adjoint_sources_kernel<<<blocks_for(n, 256), 256, 0, stream_>>>(rtf_ux_.as<>(), rtf_uz_.as<>(), obs_ux_.as<>(),
                                                                  obs_uz_.as<>(), a_stf_ux_.as<>(),
                                                                  a_stf_uz_.as<>(), n);
  PSV_CUDA(cudaGetLastError());
  PSV_CUDA(cudaStreamSynchronize(stream_));
}

void Model::reset_kernels() {
  bind();
  PSV_CUDA(cudaMemsetAsync(lame_kernels_.as<char>(), 0, lame_kernels_.bytes(), stream_));
  PSV_CUDA(cudaStreamSynchronize(stream_));
}

void Model::map_kernels_to_velocity() {
  bind();
  double *k = lame();
  map_kernels_kernel<<<blocks_for(grid_elems_, 256), 256, 0, stream_>>>(
      vp_.data(), vs_.data(), b_.data(), k, k + grid_elems_, k + 2 * grid_elems_, vp_kernel_.data(),
      vs_kernel_.data(), density_v_kernel_.data(), grid_elems_);
  PSV_CUDA(cudaGetLastError());
  PSV_CUDA(cudaStreamSynchronize(stream_));
}

// ---------------------------------------------------------------------------------------------
// model / gradient vectors
// ---------------------------------------------------------------------------------------------
namespace {
VectorArgs vector_args(const Config &c, int ld) {
  VectorArgs a;
  a.npx = c.npx;
  a.npz = c.npz;
  a.nbx = c.nx_free / c.npx;
  a.nbz = c.nz_free / c.npz;
  a.x0 = c.nx_inner_boundary + c.np_boundary;
  a.z0 = c.nz_inner_boundary + c.np_boundary;
  a.ld = ld;
  a.per = (size_t)a.nbx * a.nbz;
  return a;
}
}  // namespace

void Model::get_model_vector(double *out, size_t n) {
  bind();
  if (slab_) throw state_error("Model and gradient vectors span the whole grid: assemble them from the slabs (psvwave_b200.SlabDomain does).");
  if (n != (size_t)cfg_.free_parameters) throw invalid_argument("Model vector has the wrong length.");
  const VectorArgs a = vector_args(cfg_, ld_);
  if (scratch_.bytes() < n * sizeof(double)) scratch_.allocate(n * sizeof(double));
  vector_kernel<VEC_MEAN><<<blocks_for(a.per, 128), 128, 0, stream_>>>(a, vp_.data(), vs_.data(), rho_.data(),
                                                                       scratch_.as<>());
  PSV_CUDA(cudaGetLastError());
  PSV_CUDA(cudaMemcpyAsync(out, scratch_.as<>(), n * sizeof(double), cudaMemcpyDeviceToHost, stream_));
  PSV_CUDA(cudaStreamSynchronize(stream_));
}

void Model::set_model_vector(const double *in, size_t n) {
  bind();
  if (slab_) throw state_error("Model and gradient vectors span the whole grid: assemble them from the slabs (psvwave_b200.SlabDomain does).");
  if (n != (size_t)cfg_.free_parameters) throw invalid_argument("Model vector has the wrong length.");
  const VectorArgs a = vector_args(cfg_, ld_);
  if (scratch_.bytes() < n * sizeof(double)) scratch_.allocate(n * sizeof(double));
  PSV_CUDA(cudaMemcpyAsync(scratch_.as<>(), in, n * sizeof(double), cudaMemcpyHostToDevice, stream_));
  vector_kernel<VEC_FILL><<<blocks_for(a.per, 128), 128, 0, stream_>>>(a, vp_.data(), vs_.data(), rho_.data(),
                                                                       scratch_.as<>());
  PSV_CUDA(cudaGetLastError());
  update_from_velocity();  // src/fdModel.cpp:1456
}

void Model::get_gradient_vector(double *out, size_t n) {
  bind();
  if (slab_) throw state_error("Model and gradient vectors span the whole grid: assemble them from the slabs (psvwave_b200.SlabDomain does).");
  if (n != (size_t)cfg_.free_parameters) throw invalid_argument("Gradient vector has the wrong length.");
  const VectorArgs a = vector_args(cfg_, ld_);
  if (scratch_.bytes() < n * sizeof(double)) scratch_.allocate(n * sizeof(double));
  vector_kernel<VEC_SUM><<<blocks_for(a.per, 128), 128, 0, stream_>>>(
      a, vp_kernel_.data(), vs_kernel_.data(), density_v_kernel_.data(), scratch_.as<>());
  PSV_CUDA(cudaGetLastError());
  PSV_CUDA(cudaMemcpyAsync(out, scratch_.as<>(), n * sizeof(double), cudaMemcpyDeviceToHost, stream_));
  PSV_CUDA(cudaStreamSynchronize(stream_));
}

// ---------------------------------------------------------------------------------------------
// arrays in / out
// ---------------------------------------------------------------------------------------------
size_t Model::array_size(psv_array which) const {
  const Config &c = cfg_;
  const size_t grid = (size_t)c.nx * c.nz;
  if (which >= PSV_VX && which <= PSV_TAPER) return grid;
  if (which == PSV_T) return c.nt;
  if (which == PSV_STF) return (size_t)c.n_sources * c.nt;
  if (which == PSV_MOMENT) return (size_t)c.n_sources * 4;
  if (which >= PSV_RTF_UX && which <= PSV_A_STF_UZ) return (size_t)c.n_shots * c.nr * c.nt;
  if (which >= PSV_ACCU_VX && which <= PSV_ACCU_TXZ) return (size_t)c.n_shots * c.snapshots * grid;
  throw invalid_argument("Unknown array id.");
}

void Model::get_array(psv_array which, double *out, size_t count) {
  bind();
  if (count != array_size(which)) throw invalid_argument("Array size mismatch.");
  PSV_CUDA(cudaStreamSynchronize(stream_));
  const Config &c = cfg_;
  const size_t grid = (size_t)c.nx * c.nz;
  double *k = lame();
  switch (which) {
    case PSV_VX: return grid_to_host(field_as_double(0), out);
    case PSV_VZ: return grid_to_host(field_as_double(1), out);
    case PSV_TXX: return grid_to_host(field_as_double(2), out);
    case PSV_TZZ: return grid_to_host(field_as_double(3), out);
    case PSV_TXZ: return grid_to_host(field_as_double(4), out);
    case PSV_LM: return grid_to_host(lm_.data(), out);
    case PSV_MU: return grid_to_host(mu_.data(), out);
    case PSV_B_VX:
    case PSV_B_VZ: return grid_to_host(b_.data(), out);
    case PSV_RHO: return grid_to_host(rho_.data(), out);
    case PSV_VP: return grid_to_host(vp_.data(), out);
    case PSV_VS: return grid_to_host(vs_.data(), out);
    case PSV_LAMBDA_KERNEL: return grid_to_host(k, out);
    case PSV_MU_KERNEL: return grid_to_host(k + grid_elems_, out);
    case PSV_DENSITY_L_KERNEL: return grid_to_host(k + 2 * grid_elems_, out);
    case PSV_VP_KERNEL: return grid_to_host(vp_kernel_.data(), out);
    case PSV_VS_KERNEL: return grid_to_host(vs_kernel_.data(), out);
    case PSV_DENSITY_V_KERNEL: return grid_to_host(density_v_kernel_.data(), out);
    case PSV_LA: {  // la = lm - 2*mu (src/fdModel.cpp:1046), not stored on the device
      std::vector<double> mu(grid);
      grid_to_host(lm_.data(), out);
      grid_to_host(mu_.data(), mu.data());
      for (size_t i = 0; i < grid; ++i) out[i] = out[i] - 2 * mu[i];
      return;
    }
    case PSV_TAPER: {
      for (int ix = 0; ix < c.nx; ++ix)
        for (int iz = 0; iz < c.nz; ++iz) {
          const int gx = ix + x_off_;
          const int d = std::min(std::min(gx, nx_global_ - 1 - gx), std::min(iz, c.nz - 1 - iz)) + 1;
          out[(size_t)ix * c.nz + iz] = taper_host_[std::min(d, c.np_boundary)];
        }
      return;
    }
    case PSV_T: std::memcpy(out, t_host_.data(), count * sizeof(double)); return;
    case PSV_STF: std::memcpy(out, stf_host_.data(), count * sizeof(double)); return;
    case PSV_MOMENT: std::memcpy(out, moment_host_.data(), count * sizeof(double)); return;
    case PSV_RTF_UX:
    case PSV_RTF_UZ:
    case PSV_RTF_UX_TRUE:
    case PSV_RTF_UZ_TRUE:
    case PSV_A_STF_UX:
    case PSV_A_STF_UZ: {
      const DeviceBuffer *b[] = {&rtf_ux_, &rtf_uz_, &obs_ux_, &obs_uz_, &a_stf_ux_, &a_stf_uz_};
      if (count) copy_sync(out, b[which - PSV_RTF_UX]->as<>(), count * sizeof(double), cudaMemcpyDeviceToHost);
      return;
    }
    case PSV_ACCU_VX:
    case PSV_ACCU_VZ:
    case PSV_ACCU_TXX:
    case PSV_ACCU_TZZ:
    case PSV_ACCU_TXZ: {
      const int field = which - PSV_ACCU_VX;
      const size_t per_shot = (size_t)c.snapshots * grid;
      for (int s = 0; s < c.n_shots; ++s) get_snapshots_shot(field, s, out + (size_t)s * per_shot);
      return;
    }
    default: throw invalid_argument("Unknown array id.");
  }
}

void Model::get_snapshots_shot(int field, int s, double *dst) {
  check_shot(s);
  bind();
  const Config &c = cfg_;
  const size_t grid = (size_t)c.nx * c.nz, per_shot = (size_t)c.snapshots * grid;
  if (snapshots_[s].empty() || per_shot == 0) {
    std::memset(dst, 0, per_shot * sizeof(double));
  } else if (fp32_) {  // single-precision snapshots: one converted grid at a time
    if (conv_.buf.empty()) conv_ = make_grid();
    for (int k = 0; k < c.snapshots; ++k) {
      const float *src = snapshots_[s].as<float>() + kBulkPad + ((size_t)field * n_snap_alloc_ + k) * grid_elems_;
      to_double_kernel<<<blocks_for(grid_elems_, 256), 256, 0, stream_>>>(src, conv_.data(), grid_elems_);
      grid_to_host(conv_.data(), dst + (size_t)k * grid);
    }
  } else {
    PSV_CUDA(cudaMemcpy2DAsync(dst, c.nz * sizeof(double), snapshot_ptr(s, field, 0), ld_ * sizeof(double),
                               c.nz * sizeof(double), (size_t)c.snapshots * c.nx, cudaMemcpyDeviceToHost, stream_));
    PSV_CUDA(cudaStreamSynchronize(stream_));
  }
}

void Model::set_array(psv_array which, const double *in, size_t count) {
  bind();
  if (count != array_size(which)) throw invalid_argument("Array size mismatch.");
  PSV_CUDA(cudaStreamSynchronize(stream_));
  double *k = lame();
  switch (which) {
    case PSV_VP: return grid_from_host(vp_.data(), in);
    case PSV_VS: return grid_from_host(vs_.data(), in);
    case PSV_RHO: return grid_from_host(rho_.data(), in);
    case PSV_LAMBDA_KERNEL: return grid_from_host(k, in);
    case PSV_MU_KERNEL: return grid_from_host(k + grid_elems_, in);
    case PSV_DENSITY_L_KERNEL: return grid_from_host(k + 2 * grid_elems_, in);
    case PSV_STF:
      stf_host_.assign(in, in + count);
      if (count) copy_sync(stf_.as<>(), in, count * sizeof(double), cudaMemcpyHostToDevice);
      return;
    case PSV_MOMENT:
      moment_host_.assign(in, in + count);
      build_plans();
      return;
    case PSV_RTF_UX:
    case PSV_RTF_UZ:
    case PSV_RTF_UX_TRUE:
    case PSV_RTF_UZ_TRUE:
    case PSV_A_STF_UX:
    case PSV_A_STF_UZ: {
      DeviceBuffer *b[] = {&rtf_ux_, &rtf_uz_, &obs_ux_, &obs_uz_, &a_stf_ux_, &a_stf_uz_};
      if (count) copy_sync(b[which - PSV_RTF_UX]->as<>(), in, count * sizeof(double), cudaMemcpyHostToDevice);
      return;
    }
    default: throw invalid_argument("This array cannot be set.");
  }
}

void Model::set_parameter_fields(const double *vp, const double *vs, const double *rho, size_t count) {
  if (count != (size_t)cfg_.nx * cfg_.nz)
    throw invalid_argument("The input ndarray does not have the right shape.");
  bind();
  PSV_CUDA(cudaStreamSynchronize(stream_));
  grid_from_host(vp_.data(), vp);
  grid_from_host(vs_.data(), vs);
  grid_from_host(rho_.data(), rho);
  update_from_velocity();
}

// ---------------------------------------------------------------------------------------------
// text I/O (same formats as the reference)
// ---------------------------------------------------------------------------------------------
void Model::write_receivers() {  // src/fdModel.cpp:971-1006
  const Config &c = cfg_;
  const size_t n = array_size(PSV_RTF_UX);
  std::vector<double> ux(n), uz(n);
  get_array(PSV_RTF_UX, ux.data(), n);
  get_array(PSV_RTF_UZ, uz.data(), n);
  for (int s = 0; s < c.n_shots; ++s) {
    std::ofstream fx(c.observed_data_folder + "/rtf_ux" + std::to_string(s) + ".txt");
    std::ofstream fz(c.observed_data_folder + "/rtf_uz" + std::to_string(s) + ".txt");
    if (!fx.good() || !fz.good()) throw io_error("Cannot open receiver files for writing in '" + c.observed_data_folder + "'.");
    fx.precision(std::numeric_limits<double>::digits10 + 10);
    fz.precision(std::numeric_limits<double>::digits10 + 10);
    for (int r = 0; r < c.nr; ++r) {
      fx << std::endl;
      fz << std::endl;
      const size_t row = ((size_t)s * c.nr + r) * c.nt;
      for (int it = 0; it < c.nt; ++it) {
        fx << ux[row + it] << " ";
        fz << uz[row + it] << " ";
      }
    }
  }
}

void Model::load_receivers(bool verbose) {  // src/fdModel.cpp:1053-1126
  const Config &c = cfg_;
  const size_t n = array_size(PSV_RTF_UX_TRUE);
  std::vector<double> ux(n), uz(n);
  for (int s = 0; s < c.n_shots; ++s) {
    std::ifstream fx(c.observed_data_folder + "/rtf_ux" + std::to_string(s) + ".txt");
    std::ifstream fz(c.observed_data_folder + "/rtf_uz" + std::to_string(s) + ".txt");
    if (verbose) {
      std::cout << "File for ux data at shot " << s << " is "
                << (fx.good() ? "good (exists at least)." : "ungood.") << std::endl;
      std::cout << "File for uz data at shot " << s << " is "
                << (fz.good() ? "good (exists at least)." : "ungood.") << std::endl;
    }
    if (!fx.good() || !fz.good()) throw invalid_argument("Not all data is present!");
    for (int r = 0; r < c.nr; ++r)
      for (int it = 0; it < c.nt; ++it) {
        const size_t q = ((size_t)s * c.nr + r) * c.nt + it;
        fx >> ux[q];
        fz >> uz[q];
      }
    if (!fx.good() || !fz.good()) {
      std::cout << "Received bad state of file at end of reading! Does the data match the set up?" << std::endl;
      throw invalid_argument("Not enough data is present!");
    }
    double extra;
    fx >> extra;
    fz >> extra;
    if (fx.good() || fz.good()) {
      std::cout << "Received good state of file past reading! Does the data match the set up?" << std::endl;
      throw invalid_argument("Too much data is present!");
    }
  }
  set_array(PSV_RTF_UX_TRUE, ux.data(), n);
  set_array(PSV_RTF_UZ_TRUE, uz.data(), n);
}

void Model::write_sources() {  // src/fdModel.cpp:1008-1032
  const Config &c = cfg_;
  for (int s = 0; s < c.n_shots; ++s) {
    std::ofstream f(c.stf_folder + "/sources_shot_" + std::to_string(s) + ".txt");
    if (!f.good()) throw io_error("Cannot open source file for writing in '" + c.stf_folder + "'.");
    f.precision(std::numeric_limits<double>::digits10 + 10);
    for (int src : c.which_source_to_fire_in_which_shot[s]) {
      f << std::endl;
      for (int it = 0; it < c.nt; ++it) f << stf_host_[(size_t)src * c.nt + it] << " ";
    }
  }
}

void Model::write_kernels() {  // src/fdModel.cpp:1326-1356
  const Config &c = cfg_;
  const size_t n = (size_t)c.nx * c.nz;
  const psv_array ids[] = {PSV_VP_KERNEL, PSV_VS_KERNEL, PSV_DENSITY_V_KERNEL};
  const char *names[] = {"kernel_vp.txt", "kernel_vs.txt", "kernel_density.txt"};
  std::vector<double> k(n);
  for (int j = 0; j < 3; ++j) {
    get_array(ids[j], k.data(), n);
    std::ofstream f(names[j]);
    if (!f.good()) throw io_error(std::string("Cannot open ") + names[j] + " for writing.");
    for (int ix = 0; ix < c.nx; ++ix) {
      for (int iz = 0; iz < c.nz; ++iz) f << k[(size_t)ix * c.nz + iz] << " ";
      f << std::endl;
    }
  }
}

void Model::write_wavefield_files(int it) {  // src/fdModel.cpp:753-776
  const Config &c = cfg_;
  const size_t n = (size_t)c.nx * c.nz;
  std::vector<double> a(n), b(n);
  get_array(PSV_VX, a.data(), n);
  get_array(PSV_VZ, b.data(), n);
  std::ofstream fx("snapshots/vx" + zero_pad(it, 5) + ".txt"), fz("snapshots/vz" + zero_pad(it, 5) + ".txt");
  for (int ix = 0; ix < c.nx; ++ix) {
    for (int iz = 0; iz < c.nz; ++iz) {
      fx << a[(size_t)ix * c.nz + iz] << " ";
      fz << b[(size_t)ix * c.nz + iz] << " ";
    }
    fx << std::endl;
    fz << std::endl;
  }
}

void Model::load_model(const std::string &de_path, const std::string &vp_path, const std::string &vs_path,
                       bool verbose) {  // src/fdModel.cpp:1185-1289
  const Config &c = cfg_;
  std::ifstream de(de_path), vp(vp_path), vs(vs_path);
  if (verbose) {
    std::cout << "Loading models." << std::endl;
    std::cout << "File: " << de_path << std::endl;
    std::cout << "File for density is " << (de.good() ? "good (exists at least)." : "ungood.") << std::endl;
    std::cout << "File: " << vp_path << std::endl;
    std::cout << "File for P-wave velocity is " << (vp.good() ? "good (exists at least)." : "ungood.") << std::endl;
    std::cout << "File: " << vs_path << std::endl;
    std::cout << "File for S-wave velocity is " << (vs.good() ? "good (exists at least)." : "ungood.") << std::endl;
  }
  if (!de.good() || !vp.good() || !vs.good())
    throw invalid_argument("The files for target models don't seem to exist.\r\nPaths:\r\n\tDensity target: " +
                           de_path + "\r\n\tP-wave target : " + vp_path + "\r\n\tS-wave target : " + vs_path + "\r\n");
  const size_t n = (size_t)c.nx * c.nz;
  std::vector<double> r(n), p(n), s(n);
  for (size_t i = 0; i < n; ++i) {  // ix-major, iz fastest; (the reference also prints one line per point)
    de >> r[i];
    vp >> p[i];
    vs >> s[i];
    if (!de.good() || !vp.good() || !vs.good())
      throw invalid_argument("Received bad state of one of the files at end of reading. Does the data match the domain?");
  }
  double extra;
  de >> extra;
  vp >> extra;
  vs >> extra;
  if (de.good() || vp.good() || vs.good()) {
    std::cout << "Received good state of file past reading. Does the data match the domain?" << std::endl;
    throw invalid_argument("Too much data is present!");
  }
  set_parameter_fields(p.data(), s.data(), r.data(), n);
  if (verbose) std::cout << std::endl;
}

void Model::load_vector(const std::string &path, bool verbose, double *out, size_t n) {  // :1517-1572
  const Config &c = cfg_;
  if (n != (size_t)3 * c.nx_free * c.nz_free) throw invalid_argument("load_vector: output has the wrong length.");
  std::ifstream f(path);
  if (verbose) {
    std::cout << "Loading model vector." << std::endl;
    std::cout << "File: " << path << std::endl;
    std::cout << "File is " << (f.good() ? "good (exists at least)." : "ungood.") << std::endl;
  }
  if (!f.good())
    throw invalid_argument("The files for target models don't seem to exist.\r\nPaths:\r\n\tDensity target: " + path + "\r\n");
  for (size_t i = 0; i < n; ++i) f >> out[i];
  if (!f.good()) {
    std::cout << "Received bad state of one of the files at end of reading. Does the data match the domain?" << std::endl;
    throw invalid_argument("Not enough data is present!");
  }
  double extra;
  f >> extra;
  if (f.good()) {
    std::cout << "Received good state of file past reading. Does the data match the domain?" << std::endl;
    throw invalid_argument("Too much data is present!");
  }
}

// ---------------------------------------------------------------------------------------------
// options / statistics
// ---------------------------------------------------------------------------------------------
void Model::set_option(const std::string &name, long long v) {
  static const char *const kNeutral[] = {"profile", "graph", "resident", "resident_rows", "resident_timing",
                                         "poison_guards"};  // do not change what a step launch runs or costs
  bool neutral = false;
  for (const char *n : kNeutral) neutral = neutral || name == n;
  if (!neutral) {
    tuned_warps_ = tuned_strip_ = tuned_batch_ = 0;  // measured for the options that were in force
    pipe_resident_ = 0;
  }
  if (name == "poison_guards") {
    poison_ = v != 0;
    if (poison_) {
      poison_arrays(false);
      PSV_CUDA(cudaStreamSynchronize(stream_));
    }
  } else if (name == "tune") {
    tune_ = (int)std::max(-1ll, std::min(1ll, v));
  } else if (name == "variant") {
    if (v < 0 || v >= kVariantCount) throw invalid_argument("variant out of range");
    variant_ = (int)v;
  } else if (name == "exact") {
    exact_ = v != 0;
  } else if (name == "profile") {
    profile_ = v != 0;
  } else if (name == "strip") {
    if (v < 1) throw invalid_argument("strip must be >= 1");
    strip_ = (int)v;
  } else if (name == "fused_strip") {
    if (v < 0) throw invalid_argument("fused_strip must be >= 0 (0 = whole waves)");
    fused_strip_ = (int)v;
  } else if (name == "fp32") {
    bind();
    if ((v != 0) != fp32_) {
      PSV_CUDA(cudaStreamSynchronize(stream_));
      for (DeviceBuffer &sn : snapshots_) sn.release();  // the element size of the snapshots changes
      snapshot_pool_.clear();
      fp32_ = v != 0;
      if (fp32_) ensure_fp32();
      zero_dynamic_fields();
    }
  } else if (name == "l2_persist_mb") {
    if (v < 0 || v > 120) throw invalid_argument("l2_persist_mb must be 0..120");
    l2_persist_mb_ = (int)v;
    apply_l2_window();
  } else if (name == "pdl") {
    pdl_ = v != 0;
  } else if (name == "resident") {
    if (v < -1 || v > 1) throw invalid_argument("resident must be -1 (auto), 0 or 1");
    resident_ = (int)v;
  } else if (name == "resident_timing") {
    resident_timing_ = v != 0;


  } else if (name == "resident_rows") {
    if (v < 0 || v == 1) throw invalid_argument("resident_rows must be 0 (auto) or >= 2");
    resident_rows_ = (int)v;
  } else if (name == "graph") {
    if (v < -1 || v > 1) throw invalid_argument("graph must be -1 (auto), 0 or 1");
    graph_ = (int)v;
    if (graph_ == 0) drop_graphs();
  } else if (name == "l2_mode") {
    if (v < 0 || v > 3) throw invalid_argument("l2_mode must be 0..3");
    l2_mode_ = (int)v;
  } else if (name == "pipe_stages") {
    if (v < 2 || v > 16) throw invalid_argument("pipe_stages must be in 2..16");
    pipe_stages_ = (int)v;
    pipe_resident_ = 0;
  } else if (name == "pipe_warps") {
    if (v != 0 && v != 3 && v != 4 && v != 8) throw invalid_argument("pipe_warps must be 0 (auto), 3, 4 or 8");
    pipe_warps_ = (int)v;
    pipe_resident_ = 0;
  } else if (name == "pair_warps") {
    if (v != 4 && v != 6) throw invalid_argument("pair_warps must be 4 or 6");
    pair_warps_ = (int)v;
    pair_stages_ = 0;
  } else if (name == "pair_stages") {
    if (v < 0 || v > 8) throw invalid_argument("pair_stages must be in 0..8 (0 = as many as fit)");
    pair_stages_ = (int)v;
  } else if (name == "pair_strip") {
    if (v < 0) throw invalid_argument("pair_strip must be >= 0 (0 = whole waves)");
    pair_strip_ = (int)v;
  } else if (name == "fused_block") {
    if (v != 32 && v != 64 && v != 128) throw invalid_argument("fused_block must be 32, 64 or 128");
    fused_block_ = (int)v;
    fused_resident_blocks_ = 0;
  } else if (name == "block") {
    if (v != 32 && v != 64 && v != 96 && v != 128) throw invalid_argument("block must be 32, 64, 96 or 128");
    block_ = (int)v;
  } else {
    throw invalid_argument("Unknown option '" + name + "'.");
  }
}

long long Model::get_option(const std::string &name) const {
  if (name == "variant") return variant_;
  if (name == "exact") return exact_;
  if (name == "profile") return profile_;
  if (name == "strip") return strip_;
  if (name == "block") return block_;
  if (name == "fused_strip") return fused_strip_;
  if (name == "fused_block") return fused_block_;
  if (name == "pipe_stages") return pipe_stages_;
  if (name == "pipe_warps") return pipe_warps();
  if (name == "l2_mode") return l2_mode_;
  if (name == "pdl") return pdl_ ? 1 : 0;
  if (name == "resident") return resident_;
  if (name == "resident_rows") return resident_rows_;
  if (name == "tune") return tune_;
  if (name == "graph") return graph_;
  if (name == "graphs_cached") return (long long)graphs_.size();
  if (name == "l2_persist_mb") return l2_persist_mb_;
  if (name == "fp32") return fp32_ ? 1 : 0;
  if (name == "pair_warps") return pair_warps_;
  if (name == "pair_stages") return pair_stages();
  if (name == "pair_strip") return pair_strip_;
  throw invalid_argument("Unknown option '" + name + "'.");
}

double Model::get_stat(const std::string &name) const {
  if (name == "elapsed_ms") return stat_elapsed_ms_;
  if (name == "launches") return (double)stat_launches_;
  if (name == "stress_ms") return stat_stress_ms_;
  if (name == "velocity_ms") return stat_velocity_ms_;
  if (name == "pair_launches") return (double)stat_pair_launches_;
  if (name == "resident_ctas") return (double)stat_resident_;
  if (name == "guard_violations") return (double)const_cast<Model *>(this)->count_guard_violations();
  if (name == "tuned_warps") return tuned_warps_;
  if (name == "tuned_strip") return tuned_strip_;
  if (name == "tune_ms") return stat_tune_ms_;
  if (name.rfind("resident_phase_max", 0) == 0 && name.size() == 19 && name[18] >= '0' && name[18] <= '7')
    return stat_resident_phase_max_[name[18] - '0'];
  if (name.rfind("resident_phase", 0) == 0 && name.size() == 15 && name[14] >= '0' && name[14] <= '7')
    return stat_resident_phase_[name[14] - '0'];
  if (name == "pitch") return ld_;
  if (name == "snapshot_bytes_per_shot") return 5.0 * n_snap_alloc_ * (double)grid_elems_ * sizeof(double);
  throw invalid_argument("Unknown statistic '" + name + "'.");
}

}  // namespace psv
