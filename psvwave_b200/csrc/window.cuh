// Variant 4 (default): one launch per time step, shared-memory row WINDOW fed by TMA.
//
// Same single-pass formulation as fused.cuh (8 words read + 5 written per grid point and step,
// ping-pong state), re-cut so that the instruction stream per point is short and many warps
// fit on an SM:
//   * A CTA = NW consumer warps + 1 producer warp marches a strip of rows for a z-segment of
//     NW*30 + 4 pairs.  The producer's elected lane streams two rings of row packets into
//     shared memory with bulk-asynchronous copies (cp.async.bulk + mbarrier complete_tx):
//       ring V  {vx, vz}(row)                 -- a row stays resident for 5 march steps
//       ring T  {txx, tzz, txz, lm, mu}(j), b(j-2) -- consumed (lifted to registers) at once
//   * The OLD velocities are never held in registers: the stress update of row j reads rows
//     j-2..j+2 and the z-neighbours straight from the V window (conflict-free 16-byte LDS),
//     so there is no register rotation and no shuffle for them.
//   * The NEW stresses ride in register rings of period 4 (txx rows j-4..j-1, txz j-3..j,
//     tzz j-2..j); the march is unrolled by 4 so ring rotation is pure renaming.  Their
//     z-neighbours for the velocity update of row j-2 come from the adjacent lanes by shuffle;
//     lanes 0 and 31 of a warp are the stress halo (30 of 32 lanes own their pair).
//   * Masks (2-cell frame, pad columns, strip ends) are selects, not branches; the Cerjan taper
//     of a point is min(table[row distance], table[column distance]) (the table is monotone),
//     one uniform load per row plus two per-lane constants.
// Receiver sampling / injections sit behind one warp-uniform flag per strip and are out of
// line.  Arithmetic and evaluation order are those of stencil_math.cuh: bit-identical to the
// reference in ExactMath.
#pragma once

#include "fused.cuh"

namespace psv {

constexpr int kWinOwned = 30;   // pairs owned per consumer warp (lanes 1..30)
constexpr int kWinVArrays = 2;  // vx, vz
constexpr int kWinTArrays = 6;  // txx, tzz, txz, lm, mu (row j), b (row j-2)

// Scalars of the point operations, passed BY VALUE to the out-of-line helpers (taking the
// address of the kernel parameter block would force a local-memory copy of it).
struct PointArgs {
  const int *rec_row_start, *rec_z, *rec_id, *op_row_start;
  const RowOp *ops;
  const double *stf;
  double *ux, *uz;
  const double *ax, *az;
  int nt, it, point_mode;
  double dt, dxdz;
};

__device__ __forceinline__ PointArgs make_point_args(const FusedCommon &c, const FusedModel &m) {
  return PointArgs{c.rec_row_start, c.rec_z, c.rec_id, c.op_row_start, c.ops, c.stf, m.ux, m.uz, m.ax, m.az,
                   c.nt, c.it, c.point_mode, c.dt, c.dxdz};
}

// receiver sampling of row j from the old velocities (src/fdModel.cpp:576-596)
__device__ __noinline__ void win_record_row(const PointArgs p, int j, int kp, bool own, double2 vx, double2 vz) {
  const int q1 = p.rec_row_start[j + 1];
  for (int q = p.rec_row_start[j]; q < q1; ++q) {
    const int z = p.rec_z[q];
    if (own && (z >> 1) == kp) {
      const size_t t = (size_t)p.rec_id[q] * p.nt + p.it;
      const double sx = __ddiv_rn(__dmul_rn(p.dt, (z & 1) ? vx.y : vx.x), p.dxdz);
      const double sz = __ddiv_rn(__dmul_rn(p.dt, (z & 1) ? vz.y : vz.x), p.dxdz);
      if (p.it == 0) {
        p.ux[t] = sx;
        p.uz[t] = sz;
      } else {
        p.ux[t] = __dadd_rn(p.ux[t - 1], sx);
        p.uz[t] = __dadd_rn(p.uz[t - 1], sz);
      }
    }
  }
}

struct V2 {
  double2 vx, vz;
};

// moment-tensor injection (src/fdModel.cpp:681-751) / adjoint-source injection (:951-958) on
// the freshly updated velocities of row x, in the reference's order
__device__ __noinline__ V2 win_inject_row(const PointArgs p, int x, int kp, bool own, double2 b, double2 vx,
                                          double2 vz) {
  if (p.point_mode == POINT_FORWARD) {
    const int q1 = p.op_row_start[x + 1];
    for (int q = p.op_row_start[x]; q < q1; ++q) {
      const RowOp op = p.ops[q];
      if (own && (op.z >> 1) == kp) {
        const double w = p.stf[(size_t)op.source * p.nt + p.it];
        const bool odd = op.z & 1;
        const double bb = odd ? b.y : b.x;
        const double amp = __ddiv_rn(__dmul_rn(__dmul_rn(__dmul_rn(op.weight, w), p.dt), bb), op.d4);
        double t = op.field == 0 ? (odd ? vx.y : vx.x) : (odd ? vz.y : vz.x);
        t = op.sign > 0 ? __dadd_rn(t, amp) : __dsub_rn(t, amp);
        if (op.field == 0) {
          if (odd) vx.y = t; else vx.x = t;
        } else {
          if (odd) vz.y = t; else vz.x = t;
        }
      }
    }
  } else {
    const int q1 = p.rec_row_start[x + 1];
    for (int q = p.rec_row_start[x]; q < q1; ++q) {
      const int z = p.rec_z[q];
      if (own && (z >> 1) == kp) {
        const size_t t = (size_t)p.rec_id[q] * p.nt + p.it;
        const bool odd = z & 1;
        const double bb = odd ? b.y : b.x;
        const double ax = __ddiv_rn(__dmul_rn(__dmul_rn(p.dt, bb), p.ax[t]), p.dxdz);
        const double az = __ddiv_rn(__dmul_rn(__dmul_rn(p.dt, bb), p.az[t]), p.dxdz);
        if (odd) {
          vx.y = __dadd_rn(vx.y, ax);
          vz.y = __dadd_rn(vz.y, az);
        } else {
          vx.x = __dadd_rn(vx.x, ax);
          vz.x = __dadd_rn(vz.x, az);
        }
      }
    }
  }
  return V2{vx, vz};
}

// Shared-memory layout of one CTA (compile-time): ring V has 8 slots (a row lives for 5 march
// steps + 3 of prefetch), ring T has ST slots; the march is unrolled by 4, so every slot
// address is `half base + constant` and ring rotation costs nothing.
template <int NW, int ST>
struct WinLayout {
  static_assert(ST == 2 || ST == 4, "ring T depth must divide the unroll factor");
  static constexpr int kSV = 8;
  static constexpr int kST = ST;
  static constexpr int kOwned = NW * kWinOwned;
  static constexpr int kPairs = kOwned + 4;
  static constexpr int kSeg = kPairs * 16;
  static constexpr int kVSlot = kSeg * kWinVArrays;
  static constexpr int kTSlot = kSeg * kWinTArrays;
  static constexpr int kBarFullV = 0, kBarEmptyV = 64, kBarFullT = 128, kBarEmptyT = 128 + 8 * ST;
  static constexpr int kRingV = 256;
  static constexpr int kRingT = kRingV + kSV * kVSlot;
  static constexpr int kTotal = kRingT + ST * kTSlot;
};

extern __shared__ __align__(128) unsigned char psv_win_smem[];

struct WinCtx {
  unsigned vA, vB;   // byte offsets into psv_win_smem of the V half holding rows (U+i < 4) / the other half
  unsigned tB;       // ring T, at my pair
  unsigned bar0;     // shared-space address of the barrier block
  unsigned hA;       // 0 / 32: byte offset of half A inside the V barrier arrays
  unsigned parVw;    // parity to wait for on the V packet fetched in this loop iteration
  unsigned parT;     // parity of ring T (ST == 4)
  int j;             // row whose stresses are updated in the current step
  long long r;       // pair index of (j, my pair) in the pitched grid
  int lane, kp;
  bool own, za, zb;
  double tza, tzb;   // column taper of my two points
  long long ld2n;
  int xs, xe, nx;
  bool strip_rec, strip_inj;
  double *nf;        // new state, field 0 (vx) ; field f is nf + f*fs
  size_t fs;
};

// Keeps the compiler from hoisting every shared-memory load of a step to its top (which costs
// ~70 live registers): loads stay inside their phase.
#define PSV_SCHED_FENCE() asm volatile("" ::: "memory")

__device__ __forceinline__ double2 wlds2(unsigned off) { return *reinterpret_cast<const double2 *>(psv_win_smem + off); }
__device__ __forceinline__ double wlds1(unsigned off) { return *reinterpret_cast<const double *>(psv_win_smem + off); }

// One march step (U = step index mod 4): stresses of row j, then velocities of row j-2.
// Ring arguments arrive rotated: *_d is the slot that is dead at entry and receives row j.
template <class M, int MODE, class L, int U>
__device__ __forceinline__ void win_step(const FusedCommon &c, const FusedModel &m, WinCtx &w,
                                         double2 &xx_m4, double2 &xx_m3, double2 &xx_m2, double2 &xx_m1,
                                         double2 &xz_d, double2 &xz_m3, double2 &xz_m2, double2 &xz_m1,
                                         double2 &zz_d, double2 &zz_m3, double2 &zz_m2, double2 &zz_m1) {
  (void)zz_m3; (void)zz_m1;  // period-4 ring: rows j-3 / j-1 just ride along
  const int j = w.j;
  const int nx = w.nx;
  // row tapers of rows j and j-2 (uniform addresses, L1-resident table)
  double tr_j, tr_x;
  {
    int d = min(j, nx - 1 - j) + 1;
    d = max(min(d, c.g.np), 0);
    tr_j = __ldg(c.taper_table + d);
    d = min(j - 2, nx + 1 - j) + 1;
    d = max(min(d, c.g.np), 0);
    tr_x = __ldg(c.taper_table + d);
  }
  // V slots of rows j-2+i: index U+i in the 8-slot ring, split in two halves of 4
#define PSV_VOFF(i) ((((U + (i)) < 4) ? w.vA : w.vB) + (unsigned)(((U + (i)) & 3) * L::kVSlot))
  // ---- wait for the packets of this step: V row j+2, T row j ----
  mbar_wait(w.bar0 + L::kBarFullV + (w.hA ^ 32u) + 8 * U, w.parVw);
  constexpr int ts = U % L::kST;
  mbar_wait(w.bar0 + L::kBarFullT + 8 * ts, L::kST == 2 ? (unsigned)((U >> 1) & 1) : w.parT);
  const unsigned T = w.tB + ts * L::kTSlot;

  const unsigned Vm2 = PSV_VOFF(0), Vm1 = PSV_VOFF(1), V0 = PSV_VOFF(2), Vp1 = PSV_VOFF(3), Vp2 = PSV_VOFF(4);
  const double2 vx_m1 = wlds2(Vm1), vx_0 = wlds2(V0), vx_p1 = wlds2(Vp1), vx_p2 = wlds2(Vp2);
  const double vxL_y = wlds1(V0 - 8);
  const double2 vxR = wlds2(V0 + 16);
  const bool row_mine = j >= w.xs && j < w.xe;  // warp-uniform
  const long long r = w.r;
  // x- and z-derivatives of vx at row j (the vx rows are dead afterwards, except vx_0 below)
  const double dxp_vx_a = diff4<M>(c.k, vx_p1.x, vx_0.x, vx_m1.x, vx_p2.x);
  const double dxp_vx_b = diff4<M>(c.k, vx_p1.y, vx_0.y, vx_m1.y, vx_p2.y);
  const double dzp_vx_a = diff4<M>(c.k, vx_0.y, vx_0.x, vxL_y, vxR.x);
  const double dzp_vx_b = diff4<M>(c.k, vxR.x, vx_0.y, vx_0.x, vxR.y);
  PSV_SCHED_FENCE();
  const double2 vz_m2 = wlds2(Vm2 + L::kSeg), vz_m1 = wlds2(Vm1 + L::kSeg), vz_0 = wlds2(V0 + L::kSeg),
                vz_p1 = wlds2(Vp1 + L::kSeg);
  const double2 vzL = wlds2(V0 + L::kSeg - 16);
  const double vzR_x = wlds1(V0 + L::kSeg + 16);
  const double dzm_vz_a = diff4<M>(c.k, vz_0.x, vzL.y, vzL.x, vz_0.y);
  const double dzm_vz_b = diff4<M>(c.k, vz_0.y, vz_0.x, vzL.y, vzR_x);
  const double dxm_vz_a = diff4<M>(c.k, vz_0.x, vz_m1.x, vz_m2.x, vz_p1.x);
  const double dxm_vz_b = diff4<M>(c.k, vz_0.y, vz_m1.y, vz_m2.y, vz_p1.y);
  PSV_SCHED_FENCE();
  double2 txx = wlds2(T), tzz = wlds2(T + L::kSeg), txz = wlds2(T + 2 * L::kSeg);
  const double2 lm = wlds2(T + 3 * L::kSeg), mu = wlds2(T + 4 * L::kSeg), b = wlds2(T + 5 * L::kSeg);
  __syncwarp();
  if (w.lane == 0) mbar_arrive(w.bar0 + L::kBarEmptyT + 8 * ts);

  if (MODE == MODE_SNAPSHOT) {
    if (row_mine && w.own) {
      double *s = m.snap + c.snap_offset;
      st2s(s, r, vx_0);
      st2s(s + c.snap_field_stride, r, vz_0);
      st2s(s + 2 * c.snap_field_stride, r, txx);
      st2s(s + 3 * c.snap_field_stride, r, tzz);
      st2s(s + 4 * c.snap_field_stride, r, txz);
    }
  }
  if (MODE == MODE_ACCUMULATE) {
    const int z0 = w.kp * 2;
    if (row_mine && w.own && j >= c.wx0 && j < c.wx1 && z0 + 1 >= c.wz0 && z0 < c.wz1) {
      const double *s = m.snap + c.snap_offset;
      const double2 fvx = ld2s(s, r), fvz = ld2s(s + c.snap_field_stride, r),
                    fxx = ld2s(s + 2 * c.snap_field_stride, r), fzz = ld2s(s + 3 * c.snap_field_stride, r),
                    fxz = ld2s(s + 4 * c.snap_field_stride, r);
      double2 kr = *(reinterpret_cast<const double2 *>(m.k_rho) + r),
              kl = *(reinterpret_cast<const double2 *>(m.k_lam) + r),
              km = *(reinterpret_cast<const double2 *>(m.k_mu) + r);
      if (z0 >= c.wz0 && z0 < c.wz1)
        kernel_accumulate<M>(c.sidt, lm.x, mu.x, fvx.x, fvz.x, fxx.x, fzz.x, fxz.x, vx_0.x, vz_0.x, txx.x, tzz.x,
                             txz.x, kr.x, kl.x, km.x);
      if (z0 + 1 >= c.wz0 && z0 + 1 < c.wz1)
        kernel_accumulate<M>(c.sidt, lm.y, mu.y, fvx.y, fvz.y, fxx.y, fzz.y, fxz.y, vx_0.y, vz_0.y, txx.y, tzz.y,
                             txz.y, kr.y, kl.y, km.y);
      st2s(m.k_rho, r, kr);
      st2s(m.k_lam, r, kl);
      st2s(m.k_mu, r, km);
    }
  }
  if (w.strip_rec && row_mine) win_record_row(make_point_args(c, m), j, w.kp, w.own, vx_0, vz_0);

  // ---- stress update of row j (all 32 lanes; masked points keep their value) ----
  {
    const bool rowok = j >= 2 && j < nx - 2;
    double nxx_x = txx.x, nzz_x = tzz.x, nxz_x = txz.x, nxx_y = txx.y, nzz_y = tzz.y, nxz_y = txz.y;
    stress_update<M>(c.k, dxp_vx_a, dzm_vz_a, dzp_vx_a, dxm_vz_a, lm.x, mu.x, dmin2(tr_j, w.tza), nxx_x, nzz_x, nxz_x);
    stress_update<M>(c.k, dxp_vx_b, dzm_vz_b, dzp_vx_b, dxm_vz_b, lm.y, mu.y, dmin2(tr_j, w.tzb), nxx_y, nzz_y, nxz_y);
    const bool ua = rowok && w.za, ub = rowok && w.zb;
    txx.x = ua ? nxx_x : txx.x;
    tzz.x = ua ? nzz_x : tzz.x;
    txz.x = ua ? nxz_x : txz.x;
    txx.y = ub ? nxx_y : txx.y;
    tzz.y = ub ? nzz_y : tzz.y;
    txz.y = ub ? nxz_y : txz.y;
  }
  if (row_mine && w.own) {
    double *p = w.nf + 2 * w.fs;
    st2s(p, r, txx);
    st2s(p + w.fs, r, tzz);
    st2s(p + 2 * w.fs, r, txz);
  }
  xz_d = txz;  // row j joins the txz / tzz rings (their dead slots); txx joins after the velocity stage
  zz_d = tzz;

  PSV_SCHED_FENCE();
  // ---- velocity update of row x = j-2 ----
  const int x = j - 2;
  if (x >= w.xs) {  // warp-uniform (false only during the first 4 steps of a strip)
    double2 vx = wlds2(Vm2), vz = vz_m2;
    const double xzL_x = shfl_up1(xz_m2.x), xzL_y = shfl_up1(xz_m2.y), xzR_x = shfl_dn1(xz_m2.x);
    const double zzL_y = shfl_up1(zz_m2.y), zzR_x = shfl_dn1(zz_m2.x), zzR_y = shfl_dn1(zz_m2.y);
    const bool rowok = x >= 2 && x < nx - 2;
    double nvx_x = vx.x, nvz_x = vz.x, nvx_y = vx.y, nvz_y = vz.y;
    {
      const double dxm_txx = diff4<M>(c.k, xx_m2.x, xx_m3.x, xx_m4.x, xx_m1.x);
      const double dzm_txz = diff4<M>(c.k, xz_m2.x, xzL_y, xzL_x, xz_m2.y);
      const double dxp_txz = diff4<M>(c.k, xz_m1.x, xz_m2.x, xz_m3.x, txz.x);
      const double dzp_tzz = diff4<M>(c.k, zz_m2.y, zz_m2.x, zzL_y, zzR_x);
      velocity_update<M>(c.k, dxm_txx, dzm_txz, dxp_txz, dzp_tzz, b.x, dmin2(tr_x, w.tza), nvx_x, nvz_x);
    }
    {
      const double dxm_txx = diff4<M>(c.k, xx_m2.y, xx_m3.y, xx_m4.y, xx_m1.y);
      const double dzm_txz = diff4<M>(c.k, xz_m2.y, xz_m2.x, xzL_y, xzR_x);
      const double dxp_txz = diff4<M>(c.k, xz_m1.y, xz_m2.y, xz_m3.y, txz.y);
      const double dzp_tzz = diff4<M>(c.k, zzR_x, zz_m2.y, zz_m2.x, zzR_y);
      velocity_update<M>(c.k, dxm_txx, dzm_txz, dxp_txz, dzp_tzz, b.y, dmin2(tr_x, w.tzb), nvx_y, nvz_y);
    }
    const bool ua = rowok && w.za, ub = rowok && w.zb;
    vx.x = ua ? nvx_x : vx.x;
    vz.x = ua ? nvz_x : vz.x;
    vx.y = ub ? nvx_y : vx.y;
    vz.y = ub ? nvz_y : vz.y;
    if (w.strip_inj) {
      const V2 v = win_inject_row(make_point_args(c, m), x, w.kp, w.own, b, vx, vz);
      vx = v.vx;
      vz = v.vz;
    }
    if (w.own) {
      const long long rv = r - 2 * w.ld2n;
      st2s(w.nf, rv, vx);
      st2s(w.nf + w.fs, rv, vz);
    }
  }
  xx_m4 = txx;  // dead now: becomes row j of the txx ring
  // release the V slot of row j-2 (packet k, slot 4*(L&1)+U)
  __syncwarp();
  if (w.lane == 0) mbar_arrive(w.bar0 + L::kBarEmptyV + w.hA + 8 * U);
  w.j = j + 1;
  w.r = r + w.ld2n;
#undef PSV_VOFF
}

template <class M, int MODE, int NW, int ST, int MINB>
__global__ void __launch_bounds__((NW + 1) * 32, MINB)
    window_step_kernel(const FusedCommon c, const FusedModel *__restrict__ models) {
  using L = WinLayout<NW, ST>;
  const unsigned bar0 = smem_u32(psv_win_smem);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < L::kSV; ++s) {
      mbar_init(bar0 + L::kBarFullV + 8 * s, 1);
      mbar_init(bar0 + L::kBarEmptyV + 8 * s, NW);
    }
    for (int s = 0; s < ST; ++s) {
      mbar_init(bar0 + L::kBarFullT + 8 * s, 1);
      mbar_init(bar0 + L::kBarEmptyT + 8 * s, NW);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  const FusedModel &m = models[blockIdx.z];
  const long long ld2n = c.g.ld >> 1;
  const int nx = c.g.nx, nz = c.g.nz;
  const int xs = blockIdx.y * c.strip;
  const int xe = min(xs + c.strip, nx);
  const int steps = xe - xs + 4;                  // march steps k = 0 .. steps-1 (rows xs-2 .. xe+1)
  const int kp_seg = blockIdx.x * L::kOwned - 2;  // first pair of the CTA's row segment

  if (warp == NW) {
    // ---- producer: V packet t = row xs-4+t (t < steps+4), T packet k = t-4 = row xs-2+k ----
    if (lane == 0) {
      const double2 *old = reinterpret_cast<const double2 *>(m.f[c.parity][0]);  // field f at old + f*fs2
      const long long fs2 = (long long)(c.field_stride >> 1);
      const double2 *lm = reinterpret_cast<const double2 *>(m.lm), *mu = reinterpret_cast<const double2 *>(m.mu),
                    *bb = reinterpret_cast<const double2 *>(m.b);
      const unsigned ringV = bar0 + L::kRingV, ringT = bar0 + L::kRingT;
      long long off = (long long)(xs - 4) * ld2n + kp_seg;  // row of the V packet
      int vs = 0, tsl = 0;
      unsigned vph = 1, tph = 1;  // parity to wait for on the empty barriers
      const int nV = steps + 4;
      for (int t = 0; t < nV; ++t, off += ld2n) {
        {
          mbar_wait(bar0 + L::kBarEmptyV + 8 * vs, vph);
          const unsigned bar = bar0 + L::kBarFullV + 8 * vs;
          mbar_expect_tx(bar, L::kVSlot);
          bulk_g2s(ringV + vs * L::kVSlot, old + off, L::kSeg, bar);
          bulk_g2s(ringV + vs * L::kVSlot + L::kSeg, old + fs2 + off, L::kSeg, bar);
          if (++vs == L::kSV) {
            vs = 0;
            vph ^= 1;
          }
        }
        if (t >= 4) {
          mbar_wait(bar0 + L::kBarEmptyT + 8 * tsl, tph);
          const unsigned bar = bar0 + L::kBarFullT + 8 * tsl;
          mbar_expect_tx(bar, L::kTSlot);
          const long long toff = off - 2 * ld2n;  // row xs-2+k
          const unsigned dst = ringT + tsl * L::kTSlot;
          bulk_g2s(dst, old + 2 * fs2 + toff, L::kSeg, bar);
          bulk_g2s(dst + L::kSeg, old + 3 * fs2 + toff, L::kSeg, bar);
          bulk_g2s(dst + 2 * L::kSeg, old + 4 * fs2 + toff, L::kSeg, bar);
          bulk_g2s(dst + 3 * L::kSeg, lm + toff, L::kSeg, bar);
          bulk_g2s(dst + 4 * L::kSeg, mu + toff, L::kSeg, bar);
          bulk_g2s(dst + 5 * L::kSeg, bb + toff - 2 * ld2n, L::kSeg, bar);
          if (++tsl == ST) {
            tsl = 0;
            tph ^= 1;
          }
        }
      }
    }
    return;
  }

  // ---- consumers ----
  WinCtx w;
  const int seg_idx = warp * kWinOwned + lane + 1;  // my pair inside the segment
  w.vA = L::kRingV + seg_idx * 16;
  w.vB = w.vA + 4 * L::kVSlot;
  w.tB = L::kRingT + seg_idx * 16;
  w.bar0 = bar0;
  w.hA = 0;
  w.parVw = 0;  // packet 4: ((0+1)>>1)&1
  w.parT = 0;
  w.j = xs - 2;
  w.lane = lane;
  w.kp = kp_seg + seg_idx;
  w.r = (long long)w.j * ld2n + w.kp;
  const int z0 = w.kp * 2;
  w.own = lane >= 1 && lane <= 30 && w.kp < (int)ld2n;
  w.za = z0 >= 2 && z0 < nz - 2;
  w.zb = z0 + 1 >= 2 && z0 + 1 < nz - 2;
  {
    int da = min(z0, nz - 1 - z0) + 1, db = min(z0 + 1, nz - 2 - z0) + 1;
    da = max(min(da, c.g.np), 0);
    db = max(min(db, c.g.np), 0);
    w.tza = __ldg(c.taper_table + da);
    w.tzb = __ldg(c.taper_table + db);
  }
  w.ld2n = ld2n;
  w.xs = xs;
  w.xe = xe;
  w.nx = nx;
  w.strip_rec = false;
  w.strip_inj = false;
  if (c.point_mode != POINT_NONE) {
    const bool rec = c.rec_row_start[xe] > c.rec_row_start[xs];
    if (c.point_mode == POINT_FORWARD) {
      w.strip_rec = rec;
      w.strip_inj = c.op_row_start[xe] > c.op_row_start[xs];
    } else {
      w.strip_inj = rec;
    }
  }
  w.nf = m.f[c.parity ^ 1][0];
  w.fs = c.field_stride;

  const double2 zero2 = make_double2(0.0, 0.0);
  double2 xxa = zero2, xxb = zero2, xxc = zero2, xxd = zero2;  // txx ring
  double2 xza = zero2, xzb = zero2, xzc = zero2, xzd = zero2;  // txz ring
  double2 zza = zero2, zzb = zero2, zzc = zero2, zzd = zero2;  // tzz ring

  for (int s = 0; s < 4; ++s) mbar_wait(bar0 + L::kBarFullV + 8 * s, 0);  // rows xs-4 .. xs-1

  int left = steps;
#pragma unroll 1
  for (unsigned loop = 0;; ++loop) {
    // ring roles per sub-step: (dead -> receives row j | j-3 | j-2 | j-1), rotating by one per step
    win_step<M, MODE, L, 0>(c, m, w, xxa, xxb, xxc, xxd, xza, xzb, xzc, xzd, zza, zzb, zzc, zzd);
    if (left <= 1) break;
    win_step<M, MODE, L, 1>(c, m, w, xxb, xxc, xxd, xxa, xzb, xzc, xzd, xza, zzb, zzc, zzd, zza);
    if (left <= 2) break;
    win_step<M, MODE, L, 2>(c, m, w, xxc, xxd, xxa, xxb, xzc, xzd, xza, xzb, zzc, zzd, zza, zzb);
    if (left <= 3) break;
    win_step<M, MODE, L, 3>(c, m, w, xxd, xxa, xxb, xxc, xzd, xza, xzb, xzc, zzd, zza, zzb, zzc);
    left -= 4;
    if (left <= 0) break;
    // next group of 4 steps: the V halves swap roles
    const unsigned t = w.vA;
    w.vA = w.vB;
    w.vB = t;
    w.hA ^= 32u;
    w.parVw = ((loop + 2) >> 1) & 1u;
    w.parT = (loop + 1) & 1u;
  }
}

}  // namespace psv
