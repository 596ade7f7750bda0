// Variant 2: ONE launch per time step.
//
// The stress sweep (src/fdModel.cpp:599-644 / :864-911), the velocity sweep (:646-678 /
// :913-948), the snapshot store (:555-572), the sensitivity-kernel accumulation (:817-861),
// receiver sampling (:576-596), moment-tensor injection (:681-751) and adjoint-source
// injection (:951-958) of one time step in a single pass over the grid:
//     reads  vx vz txx tzz txz (old state) + lm mu b            8 words / point
//     writes vx vz txx tzz txz (new state, ping-pong buffers)   5 words / point
// = 104 B per grid-point update against 144 B for the two-sweep variant (176 B algorithmic
// with la, b_vz and the taper array the reference streams).
//
// Work decomposition: a WARP owns 28 z-pairs (56 grid columns) of a strip of rows and marches
// along x.  Lanes 0,1 and 30,31 carry the z-halo: lane l holds pair 28*zw - 2 + l.  The old
// velocities of rows j-2..j+2 and the NEW stresses of rows j-4..j ride in rolling register
// windows, so each row costs exactly one coalesced 16-byte load per array and lane;
// z-neighbours (old velocities for the stress update, new stresses for the velocity update)
// come from the adjacent lanes by warp shuffle.  Row j's stresses are updated first, the
// velocities of row j-2 -- whose x-operators need new stresses up to row j -- right after.
// No shared memory, no block barrier; warps are fully independent.  The state is
// double-buffered (read `old`, write `new`) because neighbouring warps/strips recompute the
// halo from the old state: strips overlap by 4 stress rows, warps by 2 pairs on each side.
//
// Point operations ride along: receivers and injection operations are bucketed by grid row
// (CSR over ix, original order kept inside a row), and the lane that owns the target point
// applies them -- sampling from the old velocities it holds, injections on the freshly
// updated velocities before they are stored -- in the reference's sequence, with the same
// explicitly rounded operations as the stand-alone point kernels (pointwise.cuh).
//
// Batched launches (Monte-Carlo ensembles): blockIdx.z selects one FusedModel (the per-model
// pointers); geometry, point-operation lists and the time index are shared by the batch.
#pragma once

#include <cstdio>

#include "fused_types.cuh"

namespace psv {

__device__ __forceinline__ double shfl_up1(double v) { return __shfl_up_sync(0xffffffffu, v, 1); }
__device__ __forceinline__ double shfl_dn1(double v) { return __shfl_down_sync(0xffffffffu, v, 1); }
__device__ __forceinline__ double2 ld2s(const double *p, long long pair) {
  return __ldg(reinterpret_cast<const double2 *>(p) + pair);
}
__device__ __forceinline__ void st2s(double *p, long long pair, double2 v) {
  *(reinterpret_cast<double2 *>(p) + pair) = v;
}

// min of two taper values (no NaNs here, so a compare-select is exact and cheaper than fmin)
__device__ __forceinline__ double dmin2(double a, double b) { return a < b ? a : b; }
// Taper table entry for distance-to-edge d (clamped into the table).  The table is monotone,
// so taper(ix, iz) = min(table[row distance], table[column distance]) (src/fdModel.cpp:263-298).
__device__ __forceinline__ double taper_axis(const double *__restrict__ table, int np, int i, int n) {
  int d = min(i, n - 1 - i) + 1;
  d = max(min(d, np), 0);
  return __ldg(table + d);
}

// The 8 double2 a lane needs for one iteration of the march (row j; velocities of row j-2).
struct RowData {
  double2 vx_p2, vz_p1;   // old vx row j+2, old vz row j+1
  double2 txx, tzz, txz;  // old stresses row j
  double2 lm, mu;         // medium row j
  double2 b;              // 1/rho row j-2
};

// Row source of variant 2: straight from global memory (read-only path).
// Rolling register windows of the march at its first row j = xs - 2.
struct MarchWindow {
  double2 vx_m2, vx_m1, vx_0, vx_p1;  // old vx rows j-2 .. j+1
  double2 vz_m2, vz_m1, vz_0;         // old vz rows j-2 .. j
};

// Window fill straight from the old state in global memory (variants 2 and 3, stage A of 4).
__device__ __forceinline__ void window_from_global(const double *ovx, const double *ovz, long long r,
                                                   long long ld2n, MarchWindow &w) {
  w.vx_m2 = __ldg(reinterpret_cast<const double2 *>(ovx) + r - 2 * ld2n);
  w.vx_m1 = __ldg(reinterpret_cast<const double2 *>(ovx) + r - ld2n);
  w.vx_0 = __ldg(reinterpret_cast<const double2 *>(ovx) + r);
  w.vx_p1 = __ldg(reinterpret_cast<const double2 *>(ovx) + r + ld2n);
  w.vz_m2 = __ldg(reinterpret_cast<const double2 *>(ovz) + r - 2 * ld2n);
  w.vz_m1 = __ldg(reinterpret_cast<const double2 *>(ovz) + r - ld2n);
  w.vz_0 = __ldg(reinterpret_cast<const double2 *>(ovz) + r);
}

// What the correlation of one row needs besides the march's own data: the forward snapshot
// (5 fields) and the three Lame kernels it adds to.
struct CorrData {
  double2 fvx, fvz, fxx, fzz, fxz, kr, kl, km;
};
__device__ __forceinline__ void corr_from_global(const FusedCommon &c, const FusedModel &m, long long r, CorrData &q) {
  const double *s = m.snap + c.snap_offset;
  q.fvx = ld2s(s, r);
  q.fvz = ld2s(s + c.snap_field_stride, r);
  q.fxx = ld2s(s + 2 * c.snap_field_stride, r);
  q.fzz = ld2s(s + 3 * c.snap_field_stride, r);
  q.fxz = ld2s(s + 4 * c.snap_field_stride, r);
  q.kr = *(reinterpret_cast<const double2 *>(m.k_rho) + r);
  q.kl = *(reinterpret_cast<const double2 *>(m.k_lam) + r);
  q.km = *(reinterpret_cast<const double2 *>(m.k_mu) + r);
}

struct GlobalRows {
  const double *ovx, *ovz, *oxx, *ozz, *oxz, *lm, *mu, *b;
  long long ld2n;
  __device__ __forceinline__ void fetch_corr(const FusedCommon &c, const FusedModel &m, long long r, bool lane_ok,
                                             CorrData &q) const {
    if (lane_ok) corr_from_global(c, m, r, q);
  }
  __device__ __forceinline__ void release() const {}
  __device__ __forceinline__ void prologue(int, long long r, MarchWindow &w) const {
    window_from_global(ovx, ovz, r, ld2n, w);
  }
  __device__ __forceinline__ void fetch(int, long long r, RowData &d) const {
    d.vx_p2 = ld2s(ovx, r + 2 * ld2n);
    d.vz_p1 = ld2s(ovz, r + ld2n);
    d.txx = ld2s(oxx, r);
    d.tzz = ld2s(ozz, r);
    d.txz = ld2s(oxz, r);
    d.lm = ld2s(lm, r);
    d.mu = ld2s(mu, r);
    d.b = ld2s(b, r - 2 * ld2n);
  }
};

// What one warp of one stage marches over.
struct MarchRange {
  int xs, xe;    // velocities of rows [xs, xe) and stresses of rows [xs-2, xe+2) are computed
  int mx0, mx1;  // rows whose global side effects (snapshot, correlation, receiver samples) are this CTA's
  bool own;      // lane computes a pair whose results go to the sink
  bool mine;     // lane's pair belongs to this CTA's tile (global side effects)
  int it;        // time index of this stage's step (receiver sample / stf / adjoint-source column)
};

// Result sink of variants 2 and 3 (and of stage B of variant 4): the new state in global memory.
struct GlobalSink {
  double *nvx, *nvz, *nxx, *nzz, *nxz;
  int xs, xe;
  bool own;
  __device__ __forceinline__ void stress(int j, long long r, double2 txx, double2 tzz, double2 txz, double2,
                                         double2) const {
    if (own && j >= xs && j < xe) {
      st2s(nxx, r, txx);
      st2s(nzz, r, tzz);
      st2s(nxz, r, txz);
    }
  }
  __device__ __forceinline__ void velocity(int, long long rv, double2 vx, double2 vz, double2) const {
    if (own) {
      st2s(nvx, rv, vx);
      st2s(nvz, rv, vz);
    }
  }
  __device__ __forceinline__ void commit() const {}
};

// One warp marches rows [xs, xe) for the z-pairs {kp0 + lane}; lanes 2..29 own their pair.
template <class M, int MODE, class Rows, class Sink>
__device__ __forceinline__ void fused_march(const FusedCommon &c, const FusedModel &m, Rows &rows, Sink &sink,
                                            const int lane, const int kp, const MarchRange rg) {
  const int nz = c.g.nz;
  const int gx0 = c.g.x_off, gnx = c.g.nx_global;  // rows of a slab count from gx0 in the whole grid
  const long long ld2n = c.g.ld >> 1;
  const int z0 = kp * 2;
  const int xs = rg.xs, xe = rg.xe;
  const bool own = rg.own, mine = rg.mine;
  const bool za = z0 >= 2 && z0 < nz - 2;
  const bool zb = z0 + 1 >= 2 && z0 + 1 < nz - 2;
  const double tza = taper_axis(c.taper_table, c.g.np, z0, nz), tzb = taper_axis(c.taper_table, c.g.np, z0 + 1, nz);

  // Cursors into the row-sorted point lists.  The row of the next receiver / injection operation sits in a
  // register, so a row without point operations costs one compare instead of two dependent loads.
  int rq = 0, rq_end = 0, oq = 0, oq_end = 0;
  if (c.point_mode == POINT_FORWARD) {  // samples on the rows [mx0, mx1), injections on [xs, xe)
    rq = c.rec_row_start[rg.mx0];
    rq_end = c.rec_row_start[rg.mx1];
    oq = c.op_row_start[xs];
    oq_end = c.op_row_start[xe];
  } else if (c.point_mode == POINT_ADJOINT) {  // adjoint sources on [xs, xe)
    rq = c.rec_row_start[xs];
    rq_end = c.rec_row_start[xe];
  }
  constexpr int kNoRow = 0x7fffffff;
  int op_next = oq < oq_end ? c.ops[oq].x : kNoRow;
  // The receivers of the strip ride in registers, one list entry per lane (row, column, trace and -- adjoint runs --
  // this step's adjoint-source samples), fetched with all loads in flight at once; a receiver row then costs a few
  // shuffles instead of three dependent global loads in EVERY warp of the launch (cold L1: ~0.5 us per receiver row,
  // 5 % of a step with 300 receivers on the 3100 x 1320 grid).  The window moves on after 32 entries.
  int win0 = rq, lx = kNoRow, lz = 0, lid = 0;
  double lax = 0.0, laz = 0.0;
  auto load_window = [&]() {
    const int q = win0 + lane;
    lx = kNoRow;
    if (q < rq_end) {
      lx = c.rec_x[q];
      lz = c.rec_z[q];
      lid = c.rec_id[q];
      if (c.point_mode == POINT_ADJOINT) {
        const size_t t = (size_t)lid * c.nt + rg.it;
        lax = m.ax[t];
        laz = m.az[t];
      }
    }
  };
  if (rq < rq_end) load_window();
  int rec_next = __shfl_sync(0xffffffffu, lx, 0);
  auto next_receiver = [&]() {  // advance the cursor; returns the row of the next receiver
    ++rq;
    if (rq >= rq_end) return kNoRow;
    if (rq - win0 == 32) {
      win0 = rq;
      load_window();
    }
    return __shfl_sync(0xffffffffu, lx, rq - win0);
  };

  // j = row whose stresses are updated in this iteration; velocities of row j-2 follow.
  int j = xs - 2;
  long long r = (long long)j * ld2n + kp;
  // old velocities: vx rows j-2..j+1 (+ j+2 loaded per iteration), vz rows j-2..j (+ j+1)
  MarchWindow w0;
  rows.prologue(j, r, w0);
  double2 vx_m2 = w0.vx_m2, vx_m1 = w0.vx_m1, vx_0 = w0.vx_0, vx_p1 = w0.vx_p1;
  double2 vz_m2 = w0.vz_m2, vz_m1 = w0.vz_m1, vz_0 = w0.vz_0;
  // new stresses: txx rows j-4..j-1, txz rows j-3..j-1, tzz rows j-2..j-1
  const double2 zero2 = make_double2(0.0, 0.0);
  double2 xx_m4 = zero2, xx_m3 = zero2, xx_m2 = zero2, xx_m1 = zero2;
  double2 xz_m3 = zero2, xz_m2 = zero2, xz_m1 = zero2;
  double2 zz_m2 = zero2, zz_m1 = zero2;

#ifndef PSV_MARCH_UNROLL
#define PSV_MARCH_UNROLL 1
#endif
#define PSV_PRAGMA_(x) _Pragma(#x)
#define PSV_UNROLL_(n) PSV_PRAGMA_(unroll n)
  PSV_UNROLL_(PSV_MARCH_UNROLL)
  for (; j < xe + 2; ++j, r += ld2n) {
    RowData d;
    rows.fetch(j, r, d);
    const double2 vx_p2 = d.vx_p2, vz_p1 = d.vz_p1;
    double2 txx = d.txx, tzz = d.tzz, txz = d.txz;
    const double2 lm = d.lm, mu = d.mu;
    const bool row_mine = j >= rg.mx0 && j < rg.mx1;  // warp-uniform
    const double tr_j = taper_axis(c.taper_table, c.g.np, gx0 + j, gnx),
                 tr_x = taper_axis(c.taper_table, c.g.np, gx0 + j - 2, gnx);

    if (row_mine) {
      if (MODE == MODE_SNAPSHOT && mine) {
        double *s = m.snap + c.snap_offset;
        st2s(s, r, vx_0);
        st2s(s + c.snap_field_stride, r, vz_0);
        st2s(s + 2 * c.snap_field_stride, r, txx);
        st2s(s + 3 * c.snap_field_stride, r, tzz);
        st2s(s + 4 * c.snap_field_stride, r, txz);
      }
      if (MODE == MODE_ACCUMULATE) {
        if (j >= c.wx0 && j < c.wx1) {  // warp-uniform: the row source stages the extra arrays for exactly these rows
          const bool lane_ok = mine && z0 + 1 >= c.wz0 && z0 < c.wz1;
          CorrData q;
          q.kr = q.kl = q.km = q.fvx = q.fvz = q.fxx = q.fzz = q.fxz = make_double2(0.0, 0.0);
          rows.fetch_corr(c, m, r, lane_ok, q);
          if (lane_ok) {
            if (z0 >= c.wz0 && z0 < c.wz1)
              kernel_accumulate<M>(c.sidt, lm.x, mu.x, q.fvx.x, q.fvz.x, q.fxx.x, q.fzz.x, q.fxz.x, vx_0.x, vz_0.x,
                                   txx.x, tzz.x, txz.x, q.kr.x, q.kl.x, q.km.x);
            if (z0 + 1 >= c.wz0 && z0 + 1 < c.wz1)
              kernel_accumulate<M>(c.sidt, lm.y, mu.y, q.fvx.y, q.fvz.y, q.fxx.y, q.fzz.y, q.fxz.y, vx_0.y, vz_0.y,
                                   txx.y, tzz.y, txz.y, q.kr.y, q.kl.y, q.km.y);
            st2s(m.k_rho, r, q.kr);
            st2s(m.k_lam, r, q.kl);
            st2s(m.k_mu, r, q.km);
          }
        }
      }
      // receiver sampling from the old velocities (src/fdModel.cpp:576-596)
      if (c.point_mode == POINT_FORWARD && j == rec_next) {
        do {
          const int z = __shfl_sync(0xffffffffu, lz, rq - win0), id = __shfl_sync(0xffffffffu, lid, rq - win0);
          if (mine && (z >> 1) == kp) {
            const size_t t = (size_t)id * c.nt + rg.it;
            const double pvx = (z & 1) ? vx_0.y : vx_0.x, pvz = (z & 1) ? vz_0.y : vz_0.x;
            const double sx = M::kExact ? __ddiv_rn(__dmul_rn(c.dt, pvx), c.dxdz) : (c.dt * pvx) * c.rdxdz;
            const double sz = M::kExact ? __ddiv_rn(__dmul_rn(c.dt, pvz), c.dxdz) : (c.dt * pvz) * c.rdxdz;
            if (c.rec_increments || rg.it == 0) {  // (no read-modify-write of global memory inside the march)
              m.ux[t] = sx;
              m.uz[t] = sz;
            } else {
              m.ux[t] = __dadd_rn(m.ux[t - 1], sx);
              m.uz[t] = __dadd_rn(m.uz[t - 1], sz);
            }
          }
          rec_next = next_receiver();
        } while (rec_next == j);
      }
    }

    rows.release();  // the row's slot of a shared-memory ring can be refilled from here on

    // ---- stress update of row j (all lanes: halo lanes feed their neighbours; masked points
    //      keep their value through selects, not branches) ----
    {
      const double vxL_y = shfl_up1(vx_0.y), vxR_x = shfl_dn1(vx_0.x), vxR_y = shfl_dn1(vx_0.y);
      const double vzL_x = shfl_up1(vz_0.x), vzL_y = shfl_up1(vz_0.y), vzR_x = shfl_dn1(vz_0.x);
      const bool rowok = gx0 + j >= 2 && gx0 + j < gnx - 2;
      double axx = txx.x, azz = tzz.x, axz = txz.x, bxx = txx.y, bzz = tzz.y, bxz = txz.y;
      {  // point (j, z0)
        const double dxp_vx = diff4<M>(c.k, vx_p1.x, vx_0.x, vx_m1.x, vx_p2.x);
        const double dzm_vz = diff4<M>(c.k, vz_0.x, vzL_y, vzL_x, vz_0.y);
        const double dzp_vx = diff4<M>(c.k, vx_0.y, vx_0.x, vxL_y, vxR_x);
        const double dxm_vz = diff4<M>(c.k, vz_0.x, vz_m1.x, vz_m2.x, vz_p1.x);
        stress_update<M>(c.k, dxp_vx, dzm_vz, dzp_vx, dxm_vz, lm.x, mu.x, dmin2(tr_j, tza), axx, azz, axz);
      }
      {  // point (j, z0+1)
        const double dxp_vx = diff4<M>(c.k, vx_p1.y, vx_0.y, vx_m1.y, vx_p2.y);
        const double dzm_vz = diff4<M>(c.k, vz_0.y, vz_0.x, vzL_y, vzR_x);
        const double dzp_vx = diff4<M>(c.k, vxR_x, vx_0.y, vx_0.x, vxR_y);
        const double dxm_vz = diff4<M>(c.k, vz_0.y, vz_m1.y, vz_m2.y, vz_p1.y);
        stress_update<M>(c.k, dxp_vx, dzm_vz, dzp_vx, dxm_vz, lm.y, mu.y, dmin2(tr_j, tzb), bxx, bzz, bxz);
      }
      const bool ua = rowok && za, ub = rowok && zb;
      txx.x = ua ? axx : txx.x;
      tzz.x = ua ? azz : tzz.x;
      txz.x = ua ? axz : txz.x;
      txx.y = ub ? bxx : txx.y;
      tzz.y = ub ? bzz : tzz.y;
      txz.y = ub ? bxz : txz.y;
    }
    sink.stress(j, r, txx, tzz, txz, lm, mu);

    // ---- velocity update of row x = j-2 (new stresses: txx x-2..x+1, txz x-1..x+2, tzz x) ----
    const int x = j - 2;
    if (x >= xs) {
      const long long rv = r - 2 * ld2n;
      const double2 b = d.b;
      double2 vx = vx_m2, vz = vz_m2;
      {
        const double xzL_x = shfl_up1(xz_m2.x), xzL_y = shfl_up1(xz_m2.y), xzR_x = shfl_dn1(xz_m2.x);
        const double zzL_y = shfl_up1(zz_m2.y), zzR_x = shfl_dn1(zz_m2.x), zzR_y = shfl_dn1(zz_m2.y);
        const bool rowok = gx0 + x >= 2 && gx0 + x < gnx - 2;
        double avx = vx.x, avz = vz.x, bvx = vx.y, bvz = vz.y;
        {
          const double dxm_txx = diff4<M>(c.k, xx_m2.x, xx_m3.x, xx_m4.x, xx_m1.x);
          const double dzm_txz = diff4<M>(c.k, xz_m2.x, xzL_y, xzL_x, xz_m2.y);
          const double dxp_txz = diff4<M>(c.k, xz_m1.x, xz_m2.x, xz_m3.x, txz.x);
          const double dzp_tzz = diff4<M>(c.k, zz_m2.y, zz_m2.x, zzL_y, zzR_x);
          velocity_update<M>(c.k, dxm_txx, dzm_txz, dxp_txz, dzp_tzz, b.x, dmin2(tr_x, tza), avx, avz);
        }
        {
          const double dxm_txx = diff4<M>(c.k, xx_m2.y, xx_m3.y, xx_m4.y, xx_m1.y);
          const double dzm_txz = diff4<M>(c.k, xz_m2.y, xz_m2.x, xzL_y, xzR_x);
          const double dxp_txz = diff4<M>(c.k, xz_m1.y, xz_m2.y, xz_m3.y, txz.y);
          const double dzp_tzz = diff4<M>(c.k, zzR_x, zz_m2.y, zz_m2.x, zzR_y);
          velocity_update<M>(c.k, dxm_txx, dzm_txz, dxp_txz, dzp_tzz, b.y, dmin2(tr_x, tzb), bvx, bvz);
        }
        const bool ua = rowok && za, ub = rowok && zb;
        vx.x = ua ? avx : vx.x;
        vz.x = ua ? avz : vz.x;
        vx.y = ub ? bvx : vx.y;
        vz.y = ub ? bvz : vz.y;
      }
      // moment-tensor injection on the updated velocities (src/fdModel.cpp:681-751)
      if (x == op_next) {
        // The operations of this row are fetched by the lanes in parallel (lane l takes operation oq + l and its
        // source-time-function sample: two global round trips for the whole row instead of two per operation),
        // then applied one after the other in list order by the lane that owns each target point.  (Keeping the
        // strip's operations in registers for the whole march, like the receivers, costs 8 registers and measured
        // 2 % slower on the bench workload: sources touch 2 strips of 37, receivers every tenth row.)
        do {
          const int n = min(32, oq_end - oq);
          int oz = 0, ofs = 0;
          double base = 0.0, dd = 1.0;
          bool valid = false;
          if (lane < n) {
            const RowOp op = c.ops[oq + lane];
            valid = op.x == x;
            if (valid) {
              oz = op.z;
              ofs = op.field | (op.sign > 0 ? 0 : 2);
              base = __dmul_rn(__dmul_rn(op.weight, c.stf[(size_t)op.source * c.nt + rg.it]), c.dt);
              dd = M::kExact ? op.d4 : op.rd4;
            }
          }
          const int cnt = __popc(__ballot_sync(0xffffffffu, valid));  // the row's operations are a prefix
          for (int l = 0; l < cnt; ++l) {
            const int z = __shfl_sync(0xffffffffu, oz, l), fs = __shfl_sync(0xffffffffu, ofs, l);
            const double ba = __shfl_sync(0xffffffffu, base, l), d = __shfl_sync(0xffffffffu, dd, l);
            if (own && (z >> 1) == kp) {
              const bool odd = z & 1;
              const double bb = odd ? b.y : b.x;
              const double amp = M::kExact ? __ddiv_rn(__dmul_rn(ba, bb), d) : (ba * bb) * d;
              double t = (fs & 1) == 0 ? (odd ? vx.y : vx.x) : (odd ? vz.y : vz.x);
              t = (fs & 2) == 0 ? __dadd_rn(t, amp) : __dsub_rn(t, amp);
              if ((fs & 1) == 0) {
                if (odd) vx.y = t; else vx.x = t;
              } else {
                if (odd) vz.y = t; else vz.x = t;
              }
            }
          }
          oq += cnt;
          op_next = oq < oq_end ? c.ops[oq].x : kNoRow;
        } while (op_next == x);
      }
      // adjoint-source injection at the receivers (src/fdModel.cpp:951-958)
      if (c.point_mode == POINT_ADJOINT && x == rec_next) {
        do {
          const int z = __shfl_sync(0xffffffffu, lz, rq - win0);
          const double nax = __shfl_sync(0xffffffffu, lax, rq - win0), naz = __shfl_sync(0xffffffffu, laz, rq - win0);
          if (own && (z >> 1) == kp) {
            const double bb = (z & 1) ? b.y : b.x;
            const double ax = M::kExact ? __ddiv_rn(__dmul_rn(__dmul_rn(c.dt, bb), nax), c.dxdz)
                                        : ((c.dt * bb) * nax) * c.rdxdz;
            const double az = M::kExact ? __ddiv_rn(__dmul_rn(__dmul_rn(c.dt, bb), naz), c.dxdz)
                                        : ((c.dt * bb) * naz) * c.rdxdz;
            if (z & 1) {
              vx.y = __dadd_rn(vx.y, ax);
              vz.y = __dadd_rn(vz.y, az);
            } else {
              vx.x = __dadd_rn(vx.x, ax);
              vz.x = __dadd_rn(vz.x, az);
            }
          }
          rec_next = next_receiver();
        } while (rec_next == x);
      }
      sink.velocity(x, rv, vx, vz, b);
    }
    sink.commit();

    // rotate the windows
    vx_m2 = vx_m1; vx_m1 = vx_0; vx_0 = vx_p1; vx_p1 = vx_p2;
    vz_m2 = vz_m1; vz_m1 = vz_0; vz_0 = vz_p1;
    xx_m4 = xx_m3; xx_m3 = xx_m2; xx_m2 = xx_m1; xx_m1 = txx;
    xz_m3 = xz_m2; xz_m2 = xz_m1; xz_m1 = txz;
    zz_m2 = zz_m1; zz_m1 = tzz;
  }
}

template <class M, int MODE>
__global__ void __launch_bounds__(128) fused_step_kernel(const FusedCommon c,
                                                         const FusedModel *__restrict__ models) {
  const int lane = threadIdx.x & 31;
  const int zw = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (zw >= c.n_zwarps) return;  // warp-uniform
  const FusedModel &m = models[blockIdx.z];
  const int kp = zw * kFusedOwnedPairs - 2 + lane;  // my z-pair (may lie outside the row: halo lanes)
  const int xs = blockIdx.y * c.strip;
  const int xe = min(xs + c.strip, c.g.nx);
  GlobalRows rows{m.f[c.parity][0], m.f[c.parity][1], m.f[c.parity][2], m.f[c.parity][3], m.f[c.parity][4],
                  m.lm, m.mu, m.b, (long long)(c.g.ld >> 1)};
  const bool own = lane >= 2 && lane < 30 && kp < (c.g.ld >> 1);
  GlobalSink sink{m.f[c.parity ^ 1][0], m.f[c.parity ^ 1][1], m.f[c.parity ^ 1][2], m.f[c.parity ^ 1][3],
                  m.f[c.parity ^ 1][4], xs, xe, own};
  fused_march<M, MODE>(c, m, rows, sink, lane, kp, MarchRange{xs, xe, xs, xe, own, own, c.it});
}

// ------------------------------------------------------------------------------------------
// Variant 3: the same march fed by a TMA pipeline.
// A CTA = NW consumer warps (adjacent z-warps of one strip) + one producer warp.  For every
// row of the march the producer's elected lane issues 8 bulk-asynchronous copies
// (cp.async.bulk global -> shared, one contiguous row segment of NW*28+4 pairs per array)
// into a ring of kPipeStages row slots, each guarded by a full/empty mbarrier pair.  The
// copies run up to kPipeStages rows ahead of the consumers, so the bytes in flight per SM are
// set by the ring depth (~100 KB per CTA), not by registers x occupancy as in variant 2.
// Consumers wait on the row's `full` barrier, lift their 8 double2 out of shared memory,
// release the slot at once (`empty` arrive, one lane per warp) and run the identical
// arithmetic.  Results go straight from registers to global memory (coalesced 16 B / lane).
// ------------------------------------------------------------------------------------------
constexpr int kPipeArrays = 8;

__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
#ifdef PSV_DEBUG_WAIT
// Debug build: bounded spin, then report who waits for what and carry on (results are garbage).
__device__ __noinline__ void mbar_wait(unsigned bar, unsigned parity) {
  for (int spin = 0; spin < 400000; ++spin) {
    unsigned ok;
    asm volatile(
        "{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    if (ok) return;
  }
  printf("mbar timeout: block (%d,%d) warp %d lane %d bar-index %u parity %u\n", blockIdx.x, blockIdx.y,
         threadIdx.x >> 5, threadIdx.x & 31, (bar & 0xffffu) / 8, parity);
}
#else
__device__ __forceinline__ void mbar_wait(unsigned bar, unsigned parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "PSV_WAIT:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra PSV_DONE;\n"
      "bra PSV_WAIT;\n"
      "PSV_DONE:\n"
      "}\n" ::"r"(bar),
      "r"(parity)
      : "memory");
}
#endif
__device__ __forceinline__ void bulk_g2s(unsigned dst, const void *src, unsigned bytes, unsigned bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}

// The same copy with an L2 eviction-priority hint (createpolicy): the medium arrays are re-read by
// every time step and fit (partly) in the 126 MB L2, the state arrays stream through.
__device__ __forceinline__ void bulk_g2s_hint(unsigned dst, const void *src, unsigned bytes, unsigned bar,
                                              unsigned long long policy) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(dst),
      "l"(src), "r"(bytes), "r"(bar), "l"(policy)
      : "memory");
}
__device__ __forceinline__ unsigned long long l2_policy_evict_last() {
  unsigned long long p;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ unsigned long long l2_policy_evict_first() {
  unsigned long long p;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
  return p;
}

template <int NW>
struct PipeGeom {
  static constexpr int kOwned = NW * kFusedOwnedPairs;  // pairs owned by the CTA
  static constexpr int kPairs = kOwned + 4;             // + 2 halo pairs on each side
  static constexpr int kSeg = kPairs * 16;              // bytes of one row segment of one array
  static constexpr int kStage = kSeg * kPipeArrays;     // bytes of one ring slot
};

// Segments of one ring slot: the 8 arrays of the march, plus -- in correlation launches, for the rows
// of the free window -- the 5 snapshot fields and the 3 Lame kernels (src/fdModel.cpp:817-861), so
// that the correlation reads ride the same TMA pipeline instead of stalling the march on global loads.
constexpr int kPipeArraysCorr = 16;

// Row source of variant 3: the ring in shared memory.
struct PipeRows {
  const unsigned char *ring;  // first slot
  unsigned full0, empty0;     // shared addresses of the barrier arrays
  int stages, seg, stage_bytes;
  int lane_off;               // byte offset of my pair inside a segment
  int slot;
  unsigned phase;
  int lane;
  const double *ovx, *ovz;    // old velocities in global memory (window fill)
  long long ld2n;
  bool has_corr;              // correlation rows carry segments 8..15
  __device__ __forceinline__ void prologue(int, long long r, MarchWindow &w) const {
    window_from_global(ovx, ovz, r, ld2n, w);
  }
  __device__ __forceinline__ void fetch(int, long long, RowData &d) {
    mbar_wait(full0 + 8 * slot, phase);
    const unsigned char *s = ring + (size_t)slot * stage_bytes + lane_off;
    d.vx_p2 = *reinterpret_cast<const double2 *>(s);
    d.vz_p1 = *reinterpret_cast<const double2 *>(s + seg);
    d.txx = *reinterpret_cast<const double2 *>(s + 2 * seg);
    d.tzz = *reinterpret_cast<const double2 *>(s + 3 * seg);
    d.txz = *reinterpret_cast<const double2 *>(s + 4 * seg);
    d.lm = *reinterpret_cast<const double2 *>(s + 5 * seg);
    d.mu = *reinterpret_cast<const double2 *>(s + 6 * seg);
    d.b = *reinterpret_cast<const double2 *>(s + 7 * seg);
  }
  __device__ __forceinline__ void fetch_corr(const FusedCommon &c, const FusedModel &m, long long r, bool lane_ok,
                                             CorrData &q) const {
    if (has_corr) {
      const unsigned char *s = ring + (size_t)slot * stage_bytes + lane_off;
      q.fvx = *reinterpret_cast<const double2 *>(s + 8 * seg);
      q.fvz = *reinterpret_cast<const double2 *>(s + 9 * seg);
      q.fxx = *reinterpret_cast<const double2 *>(s + 10 * seg);
      q.fzz = *reinterpret_cast<const double2 *>(s + 11 * seg);
      q.fxz = *reinterpret_cast<const double2 *>(s + 12 * seg);
      q.kr = *reinterpret_cast<const double2 *>(s + 13 * seg);
      q.kl = *reinterpret_cast<const double2 *>(s + 14 * seg);
      q.km = *reinterpret_cast<const double2 *>(s + 15 * seg);
    } else if (lane_ok) {
      corr_from_global(c, m, r, q);
    }
  }
  // hand the slot back (everything this warp wants from it sits in registers)
  __device__ __forceinline__ void release() {
    __syncwarp();
    if (lane == 0) mbar_arrive(empty0 + 8 * slot);
    if (++slot == stages) {
      slot = 0;
      phase ^= 1;
    }
  }
};

template <class M, int MODE, int NW>
__global__ void __launch_bounds__((NW + 1) * 32) fused_pipe_kernel(const FusedCommon c,
                                                                   const FusedModel *__restrict__ models,
                                                                   const int stages) {
  using G = PipeGeom<NW>;
  extern __shared__ __align__(128) unsigned char psv_smem[];
  unsigned long long *bars = reinterpret_cast<unsigned long long *>(psv_smem);  // [stages] full, [stages] empty
  unsigned char *ring = psv_smem + 128 * ((2 * stages * 8 + 127) / 128);
  const unsigned full0 = smem_u32(bars), empty0 = full0 + 8 * stages;
  constexpr int kSlotBytes = G::kSeg * (MODE == MODE_ACCUMULATE ? kPipeArraysCorr : kPipeArrays);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < stages; ++s) {
      mbar_init(full0 + 8 * s, 1);
      mbar_init(empty0 + 8 * s, NW);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (c.pdl) {
    // Programmatic dependent launch: the next time step's CTAs may be scheduled (and set up their barriers) on
    // SMs this launch has already left; everything that touches the wavefield -- this launch reads the state
    // the previous one wrote and overwrites the state it read -- comes after the wait.
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
  }

  const FusedModel &m = models[blockIdx.z];
  const long long ld2n = c.g.ld >> 1;
  const int xs = blockIdx.y * c.strip;
  const int xe = min(xs + c.strip, c.g.nx);
  const int kp_tile = blockIdx.x * G::kOwned - 2;  // first pair of the CTA's row segment

  if (warp == NW) {
    // ---- producer: one lane streams the rows of the strip into the ring ----
    if (lane == 0) {
      const double *src[kPipeArrays] = {m.f[c.parity][0], m.f[c.parity][1], m.f[c.parity][2], m.f[c.parity][3],
                                        m.f[c.parity][4], m.lm,           m.mu,             m.b};
      const long long row_off[kPipeArrays] = {2, 1, 0, 0, 0, 0, 0, -2};  // row of each array relative to j
      int slot = 0;
      unsigned phase = 0;
      const unsigned ring_u32 = smem_u32(ring);
      // correlation rows (the consumers use the same predicate: row of this strip inside the free window)
      const int cx0 = max(xs, c.wx0), cx1 = min(xe, c.wx1);
      const unsigned long long keep = l2_policy_evict_last(), stream = l2_policy_evict_first();
      const double *snap0 = MODE == MODE_ACCUMULATE ? m.snap + c.snap_offset : nullptr;
      for (int j = xs - 2; j < xe + 2; ++j) {
        mbar_wait(empty0 + 8 * slot, phase ^ 1);
        const unsigned bar = full0 + 8 * slot;
        const bool corr = MODE == MODE_ACCUMULATE && j >= cx0 && j < cx1;
        mbar_expect_tx(bar, corr ? G::kSeg * kPipeArraysCorr : G::kStage);
        const unsigned dst = ring_u32 + slot * kSlotBytes;
#pragma unroll
        for (int a = 0; a < kPipeArrays; ++a) {
          const double2 *from = reinterpret_cast<const double2 *>(src[a]) + ((long long)j + row_off[a]) * ld2n + kp_tile;
          const bool medium = a >= 5;
          if (medium && (c.l2_mode & 1)) bulk_g2s_hint(dst + a * G::kSeg, from, G::kSeg, bar, keep);
          else if (!medium && (c.l2_mode & 2)) bulk_g2s_hint(dst + a * G::kSeg, from, G::kSeg, bar, stream);
          else bulk_g2s(dst + a * G::kSeg, from, G::kSeg, bar);
        }
        if (corr) {
          const long long off = (long long)j * ld2n + kp_tile;
#pragma unroll
          for (int f = 0; f < 5; ++f)
            bulk_g2s(dst + (8 + f) * G::kSeg, reinterpret_cast<const double2 *>(snap0 + f * c.snap_field_stride) + off,
                     G::kSeg, bar);
          bulk_g2s(dst + 13 * G::kSeg, reinterpret_cast<const double2 *>(m.k_rho) + off, G::kSeg, bar);
          bulk_g2s(dst + 14 * G::kSeg, reinterpret_cast<const double2 *>(m.k_lam) + off, G::kSeg, bar);
          bulk_g2s(dst + 15 * G::kSeg, reinterpret_cast<const double2 *>(m.k_mu) + off, G::kSeg, bar);
        }
        if (++slot == stages) {
          slot = 0;
          phase ^= 1;
        }
      }
    }
    return;
  }

  // ---- consumers ----
  PipeRows rows{ring, full0, empty0, stages, G::kSeg, kSlotBytes, (warp * kFusedOwnedPairs + lane) * 16, 0, 0u, lane,
                m.f[c.parity][0], m.f[c.parity][1], ld2n, MODE == MODE_ACCUMULATE};
  const int kp = kp_tile + warp * kFusedOwnedPairs + lane;
  const bool own = lane >= 2 && lane < 30 && kp < (int)ld2n;
  GlobalSink sink{m.f[c.parity ^ 1][0], m.f[c.parity ^ 1][1], m.f[c.parity ^ 1][2], m.f[c.parity ^ 1][3],
                  m.f[c.parity ^ 1][4], xs, xe, own};
  fused_march<M, MODE>(c, m, rows, sink, lane, kp, MarchRange{xs, xe, xs, xe, own, own, c.it});
}

}  // namespace psv
