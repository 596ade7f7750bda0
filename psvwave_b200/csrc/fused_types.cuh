// Plain-data arguments of the fused step kernels (no device code: host launchers and kernels share it).
#pragma once

#include "kernels.cuh"

namespace psv {

constexpr int kFusedOwnedPairs = 28;  // lanes 2..29 of a warp

// Injection operation bucketed by row (sorted by target row, original order kept).
struct RowOp {
  int z;          // target column
  int field;      // 0 = vx, 1 = vz
  int source;     // row of stf
  int x;          // target row (local)
  double weight;  // moment entry (x0.25 where the reference does)
  double sign;    // +1 / -1
  double d4;      // dx^4 or dz^4
  double rd4;     // 1 / d4 (default arithmetic multiplies, "exact" divides like the reference)
};

enum PointMode { POINT_NONE = 0, POINT_FORWARD = 1, POINT_ADJOINT = 2 };

struct FusedModel {  // per-model pointers (constant during one simulate call)
  double *f[2][5];   // ping-pong state: [parity][vx vz txx tzz txz]
  const double *lm, *mu, *b;
  double *snap;      // this shot's snapshots [5][n_snap_alloc][nx*ld]
  double *k_lam, *k_mu, *k_rho;
  double *ux, *uz;        // synthetics of this shot [nr][nt]      (POINT_FORWARD)
  const double *ax, *az;  // adjoint sources of this shot [nr][nt] (POINT_ADJOINT)
};

struct FusedCommon {
  GridShape g;
  SweepConsts k;
  const double *taper_table;
  int parity;       // old state = f[parity], new state = f[parity ^ 1]
  int n_zwarps;     // ceil(ld/2 / 28)
  int strip;        // rows per warp
  int tile_pairs;   // variant 4: z-pairs per CTA tile
  int pdl;             // variant 3: launched with programmatic stream serialization (see fused_pipe_kernel)
  int rec_increments;  // 1: receiver samples are stored as increments, the running sum follows the time loop
  int l2_mode;      // variant 3: bit 0 = medium rows evict_last in L2, bit 1 = state rows evict_first
  int it_step;      // variant 4: time index of the second step = it + it_step (+1 forward, -1 adjoint)
  size_t field_stride;  // doubles between consecutive state arrays: f[p][k] = f[0][0] + (5p+k)*field_stride
  // snapshots / kernel accumulation
  size_t snap_field_stride;  // n_snap_alloc * nx*ld
  size_t snap_offset;        // snapshot index * nx*ld
  int wx0, wx1, wz0, wz1;
  double sidt;
  // point operations
  int point_mode;
  int nt, it;
  double dt, dxdz, rdxdz;  // rdxdz = 1 / (dx*dz), default arithmetic only
  const int *rec_row_start;  // [nx+1] CSR over rows of the receivers sorted by (ix, r)
  const int *rec_z;          // column of sorted receiver q
  const int *rec_x;          // row of sorted receiver q
  const int *rec_id;         // original receiver number of sorted receiver q
  const int *op_row_start;   // [nx+1] CSR over rows of this shot's injection operations
  const RowOp *ops;
  const double *stf;         // [n_sources][nt]
};

struct FusedModel32 {  // per-model pointers of the FP32 mode
  float *f[2][5];      // ping-pong state: [parity][vx vz txx tzz txz]
  const float *lm, *mu, *b;
  float *snap;         // this shot's snapshots [5][n_snap_alloc][nx*ld], single precision
  double *k_lam, *k_mu, *k_rho;
  double *ux, *uz;
  const double *ax, *az;
};

// ---- resident whole-loop kernel (resident.cuh) ----
constexpr int kResidentThreads = 512;
constexpr int kResidentMaxOps = 128;
constexpr size_t kResidentSmemBudget = 232448 - 8192;  // dynamic shared memory of the loop kernel (its static arrays aside)  // injection operations (forward) / receivers (adjoint) of one strip

struct ResidentArgs {
  FusedCommon c;  // geometry, constants (k.sdt carries the sign of the time direction), point lists, window
  const FusedModel *models;
  // Velocity rows in flight between neighbouring strips: [2 parities][vx vz][nx][ld] packets of 16 bytes, each
  // {low word, flag, high word, flag} of one double with flag = 1 + the time steps its producer has finished
  // (zero at launch).  A reader polls the packet itself: no separate counters, no fences.
  uint4 *packets;
  unsigned flag_base;  // flags of this launch are flag_base + 1 .. flag_base + nsteps
  unsigned *abort;   // set to 1 when a wait timed out (host turns it into an error)
  int rows;          // rows per CTA
  int nsteps;        // nt
  int direction;     // +1 forward, -1 adjoint
  int store;         // forward: store snapshots every si steps
  int si;            // snapshot interval
  size_t grid_elems; // nx * ld (one snapshot of one field)
  long long *timing;          // optional [CTAs][8]: clock cycles thread 0 spent per phase, summed over the loop
};

// Shared-memory tiles of one CTA (offsets in doubles).  Every tile is preceded by 16 doubles of slack, so the
// z-neighbour reads of a row's first / last pair stay inside the allocation (they only feed frame points, whose
// results are discarded).
struct ResidentLayout {
  int pitch, rows;
  int vx, vz, xx, zz, xz, lm, mu, b, tz, tr, head, items, end;
  bool cached;  // the per-thread item records (resident.cuh) fit next to the tiles
  __host__ __device__ ResidentLayout(int ld, int r) : pitch(ld), rows(r) {
    int off = 16;  // slack
    vx = off; off += (r + 8) * ld + 16;
    vz = off; off += (r + 8) * ld + 16;
    xx = off; off += (r + 4) * ld + 16;
    zz = off; off += (r + 4) * ld + 16;
    xz = off; off += (r + 4) * ld + 16;
    lm = off; off += (r + 4) * ld + 16;
    mu = off; off += (r + 4) * ld + 16;
    b = off; off += r * ld + 16;
    tz = off; off += ld;                      // Cerjan taper by column
    tr = off; off += ((r + 4 + 15) / 16) * 16;  // Cerjan taper by row, rows xs-2 .. xe+1
    head = off; off += ((r * (ld / 2) + 1) / 2 + 15) / 16 * 16;  // ints: first injection of each own pair
    items = off;  // [threads] stress records of 3 doubles, then [threads] velocity records of 4 doubles
    cached = (size_t)(off + 7 * kResidentThreads) * sizeof(double) <= kResidentSmemBudget;
    if (cached) off += 7 * kResidentThreads;
    end = off;
  }
  __host__ __device__ size_t bytes() const { return (size_t)end * sizeof(double); }
};

}  // namespace psv
