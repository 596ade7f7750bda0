// sm_100a kernels of the P-SV time loop.
//
// Device layout: every grid array is [nx][ld] doubles, iz fastest, with row pitch
// ld = round_up(nz, 16) (rows start on 128-byte lines, ld is even so rows split into
// aligned double2 "z-pairs") and a guard band of 2 rows + 16 doubles in front of and behind
// the array, so the stencil's +-2 row / +-1 pair neighbour loads never need a bounds test.
// Pad columns (nz <= iz < ld) and the guard band hold zeros and are never written with
// anything else.
//
// Sweeps (one launch each per time step):
//   stress sweep   = reference loop src/fdModel.cpp:599-644 (forward) / :864-911 (adjoint)
//                    fused with the snapshot store :555-572 (MODE_SNAPSHOT) or the
//                    sensitivity-kernel accumulation :817-861 (MODE_ACCUMULATE); both act on
//                    the state *before* the update, which this sweep has in registers.
//   velocity sweep = reference loop src/fdModel.cpp:646-678 / :913-948
// Point operations (receiver sampling, source / adjoint-source injection) are tiny launches
// ordered on the same stream (sampling.cuh).
#pragma once

#include "stencil_math.cuh"

namespace psv {

enum SweepMode { MODE_PLAIN = 0, MODE_SNAPSHOT = 1, MODE_ACCUMULATE = 2 };

struct GridShape {
  int nx, nz, ld;  // ld = row pitch in doubles (even, multiple of 16)
  int np;          // taper width
  // x-slab of a larger grid (domain split, fused variants only): local row j is row j + x_off of
  // a grid of nx_global rows.  A whole grid has x_off = 0, nx_global = nx.
  int x_off, nx_global;
};

struct StressArgs {
  GridShape g;
  SweepConsts k;
  const double *vx, *vz;    // read only in this sweep
  double *txx, *tzz, *txz;  // updated in place (point-wise dependence only)
  const double *lm, *mu;
  const double *taper_table;
  // MODE_SNAPSHOT: destinations; MODE_ACCUMULATE: sources (the stored forward state)
  double *s_vx, *s_vz, *s_txx, *s_tzz, *s_txz;
  // MODE_ACCUMULATE
  double *k_rho, *k_lam, *k_mu;
  int wx0, wx1, wz0, wz1;  // free-parameter window [wx0,wx1) x [wz0,wz1)
  double sidt;             // snapshot_interval * dt
  int strip;               // rows per thread in the marching variant
};

struct VelocityArgs {
  GridShape g;
  SweepConsts k;
  const double *txx, *tzz, *txz;  // read only
  double *vx, *vz;                // updated in place
  const double *b;                // 1/rho
  const double *taper_table;
  int strip;
};

// ------------------------------------------------------------------------------------------
// Variant 0: one thread per grid point, every neighbour fetched straight from global/L1.
// The plainest possible statement of the sweeps; kept as the in-tree cross-check for the
// optimised variants.
// ------------------------------------------------------------------------------------------
template <class M, int MODE>
__global__ void __launch_bounds__(256) stress_point_kernel(const StressArgs a) {
  const int iz = blockIdx.x * 32 + threadIdx.x;
  const int ix = blockIdx.y * 8 + threadIdx.y;
  const int nx = a.g.nx, nz = a.g.nz, ld = a.g.ld;
  if (iz >= nz || ix >= nx) return;
  const size_t i = (size_t)ix * ld + iz;
  double txx = a.txx[i], tzz = a.tzz[i], txz = a.txz[i];
  const double vx0 = a.vx[i], vz0 = a.vz[i];
  if (MODE == MODE_SNAPSHOT) {
    a.s_vx[i] = vx0;
    a.s_vz[i] = vz0;
    a.s_txx[i] = txx;
    a.s_tzz[i] = tzz;
    a.s_txz[i] = txz;
  }
  const bool interior = ix >= 2 && ix < nx - 2 && iz >= 2 && iz < nz - 2;
  double lm = 0, mu = 0;
  if (interior || MODE == MODE_ACCUMULATE) {
    lm = a.lm[i];
    mu = a.mu[i];
  }
  if (MODE == MODE_ACCUMULATE) {
    if (ix >= a.wx0 && ix < a.wx1 && iz >= a.wz0 && iz < a.wz1) {
      double kr = a.k_rho[i], kl = a.k_lam[i], km = a.k_mu[i];
      kernel_accumulate<M>(a.sidt, lm, mu, a.s_vx[i], a.s_vz[i], a.s_txx[i], a.s_tzz[i], a.s_txz[i],
                           vx0, vz0, txx, tzz, txz, kr, kl, km);
      a.k_rho[i] = kr;
      a.k_lam[i] = kl;
      a.k_mu[i] = km;
    }
  }
  if (!interior) return;
  const double *vx = a.vx, *vz = a.vz;
  const double dxp_vx = diff4<M>(a.k, vx[i + ld], vx0, vx[i - ld], vx[i + 2 * (size_t)ld]);
  const double dzm_vz = diff4<M>(a.k, vz0, vz[i - 1], vz[i - 2], vz[i + 1]);
  const double dzp_vx = diff4<M>(a.k, vx[i + 1], vx0, vx[i - 1], vx[i + 2]);
  const double dxm_vz = diff4<M>(a.k, vz0, vz[i - ld], vz[i - 2 * (size_t)ld], vz[i + ld]);
  const double tp = taper_at(a.taper_table, a.g.np, ix, iz, nx, nz);
  stress_update<M>(a.k, dxp_vx, dzm_vz, dzp_vx, dxm_vz, lm, mu, tp, txx, tzz, txz);
  a.txx[i] = txx;
  a.tzz[i] = tzz;
  a.txz[i] = txz;
}

template <class M>
__global__ void __launch_bounds__(256) velocity_point_kernel(const VelocityArgs a) {
  const int iz = blockIdx.x * 32 + threadIdx.x;
  const int ix = blockIdx.y * 8 + threadIdx.y;
  const int nx = a.g.nx, nz = a.g.nz, ld = a.g.ld;
  if (!(ix >= 2 && ix < nx - 2 && iz >= 2 && iz < nz - 2)) return;
  const size_t i = (size_t)ix * ld + iz;
  const double *txx = a.txx, *tzz = a.tzz, *txz = a.txz;
  const double txx0 = txx[i], tzz0 = tzz[i], txz0 = txz[i];
  const double dxm_txx = diff4<M>(a.k, txx0, txx[i - ld], txx[i - 2 * (size_t)ld], txx[i + ld]);
  const double dzm_txz = diff4<M>(a.k, txz0, txz[i - 1], txz[i - 2], txz[i + 1]);
  const double dxp_txz = diff4<M>(a.k, txz[i + ld], txz0, txz[i - ld], txz[i + 2 * (size_t)ld]);
  const double dzp_tzz = diff4<M>(a.k, tzz[i + 1], tzz0, tzz[i - 1], tzz[i + 2]);
  const double tp = taper_at(a.taper_table, a.g.np, ix, iz, nx, nz);
  double vx = a.vx[i], vz = a.vz[i];
  velocity_update<M>(a.k, dxm_txx, dzm_txz, dxp_txz, dzp_tzz, a.b[i], tp, vx, vz);
  a.vx[i] = vx;
  a.vz[i] = vz;
}

// ------------------------------------------------------------------------------------------
// Variant 1: vectorised x-marching sweeps.
// A thread owns one z-pair (two adjacent iz, one aligned 16-byte double2) and walks `strip`
// consecutive rows.  Everything the x-direction operators need travels with it in a rolling
// register window (4 rows of each differentiated field), so each row costs one new
// coalesced 128-bit load per x-differentiated field plus the point-wise arrays; the z
// operators read the two neighbouring pairs of the current row (L1 hits: the same lines were
// streamed in as the row-ahead loads of the neighbouring threads).  No shared memory, no
// block barrier: the sweep is a pure stream, all 32 lanes of every warp issue the same
// 16-byte-per-lane loads.
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ double2 ld2(const double *p, size_t pair) {
  return __ldg(reinterpret_cast<const double2 *>(p) + pair);
}
__device__ __forceinline__ double2 ld2_rw(const double *p, size_t pair) {
  return *(reinterpret_cast<const double2 *>(p) + pair);
}
__device__ __forceinline__ void st2(double *p, size_t pair, double2 v) {
  *(reinterpret_cast<double2 *>(p) + pair) = v;
}

template <class M, int MODE>
__global__ void __launch_bounds__(128) stress_march_kernel(const StressArgs a) {
  const int nx = a.g.nx, nz = a.g.nz, ld = a.g.ld;
  const int ld2n = ld >> 1;
  const int kp = blockIdx.x * blockDim.x + threadIdx.x;  // pair index inside the row
  if (kp >= ld2n) return;
  const int z0 = kp * 2;
  const int xs = blockIdx.y * a.strip;
  const int xe = min(xs + a.strip, nx);
  // element validity in z (rows are tested per iteration)
  const bool za = z0 >= 2 && z0 < nz - 2;
  const bool zb = z0 + 1 >= 2 && z0 + 1 < nz - 2;

  size_t r = (size_t)xs * ld2n + kp;
  // rolling windows: vx rows x-1..x+2, vz rows x-2..x+1 (guard rows cover x<0 and x>=nx)
  double2 vx_m1 = ld2(a.vx, r - ld2n), vx_0 = ld2(a.vx, r), vx_p1 = ld2(a.vx, r + ld2n);
  double2 vz_m2 = ld2(a.vz, r - 2 * (size_t)ld2n), vz_m1 = ld2(a.vz, r - ld2n), vz_0 = ld2(a.vz, r);

#pragma unroll 2
  for (int x = xs; x < xe; ++x, r += ld2n) {
    const double2 vx_p2 = ld2(a.vx, r + 2 * (size_t)ld2n);
    const double2 vz_p1 = ld2(a.vz, r + ld2n);
    double2 txx = ld2_rw(a.txx, r), tzz = ld2_rw(a.tzz, r), txz = ld2_rw(a.txz, r);
    const double2 lm = ld2(a.lm, r), mu = ld2(a.mu, r);

    if (MODE == MODE_SNAPSHOT) {
      st2(a.s_vx, r, vx_0);
      st2(a.s_vz, r, vz_0);
      st2(a.s_txx, r, txx);
      st2(a.s_tzz, r, tzz);
      st2(a.s_txz, r, txz);
    }
    if (MODE == MODE_ACCUMULATE) {
      if (x >= a.wx0 && x < a.wx1 && z0 + 1 >= a.wz0 && z0 < a.wz1) {
        const double2 fvx = ld2(a.s_vx, r), fvz = ld2(a.s_vz, r), fxx = ld2(a.s_txx, r),
                      fzz = ld2(a.s_tzz, r), fxz = ld2(a.s_txz, r);
        double2 kr = ld2_rw(a.k_rho, r), kl = ld2_rw(a.k_lam, r), km = ld2_rw(a.k_mu, r);
        if (z0 >= a.wz0 && z0 < a.wz1)
          kernel_accumulate<M>(a.sidt, lm.x, mu.x, fvx.x, fvz.x, fxx.x, fzz.x, fxz.x, vx_0.x, vz_0.x,
                               txx.x, tzz.x, txz.x, kr.x, kl.x, km.x);
        if (z0 + 1 >= a.wz0 && z0 + 1 < a.wz1)
          kernel_accumulate<M>(a.sidt, lm.y, mu.y, fvx.y, fvz.y, fxx.y, fzz.y, fxz.y, vx_0.y, vz_0.y,
                               txx.y, tzz.y, txz.y, kr.y, kl.y, km.y);
        st2(a.k_rho, r, kr);
        st2(a.k_lam, r, kl);
        st2(a.k_mu, r, km);
      }
    }

    if (x >= 2 && x < nx - 2 && (za || zb)) {
      // z-neighbours of the current row
      const double2 vxL = ld2(a.vx, r - 1), vxR = ld2(a.vx, r + 1);
      const double2 vzL = ld2(a.vz, r - 1), vzR = ld2(a.vz, r + 1);
      if (za) {  // point (x, z0)
        const double dxp_vx = diff4<M>(a.k, vx_p1.x, vx_0.x, vx_m1.x, vx_p2.x);
        const double dzm_vz = diff4<M>(a.k, vz_0.x, vzL.y, vzL.x, vz_0.y);
        const double dzp_vx = diff4<M>(a.k, vx_0.y, vx_0.x, vxL.y, vxR.x);
        const double dxm_vz = diff4<M>(a.k, vz_0.x, vz_m1.x, vz_m2.x, vz_p1.x);
        const double tp = taper_at(a.taper_table, a.g.np, x, z0, nx, nz);
        stress_update<M>(a.k, dxp_vx, dzm_vz, dzp_vx, dxm_vz, lm.x, mu.x, tp, txx.x, tzz.x, txz.x);
      }
      if (zb) {  // point (x, z0+1)
        const double dxp_vx = diff4<M>(a.k, vx_p1.y, vx_0.y, vx_m1.y, vx_p2.y);
        const double dzm_vz = diff4<M>(a.k, vz_0.y, vz_0.x, vzL.y, vzR.x);
        const double dzp_vx = diff4<M>(a.k, vxR.x, vx_0.y, vx_0.x, vxR.y);
        const double dxm_vz = diff4<M>(a.k, vz_0.y, vz_m1.y, vz_m2.y, vz_p1.y);
        const double tp = taper_at(a.taper_table, a.g.np, x, z0 + 1, nx, nz);
        stress_update<M>(a.k, dxp_vx, dzm_vz, dzp_vx, dxm_vz, lm.y, mu.y, tp, txx.y, tzz.y, txz.y);
      }
      st2(a.txx, r, txx);
      st2(a.tzz, r, tzz);
      st2(a.txz, r, txz);
    }
    vx_m1 = vx_0;
    vx_0 = vx_p1;
    vx_p1 = vx_p2;
    vz_m2 = vz_m1;
    vz_m1 = vz_0;
    vz_0 = vz_p1;
  }
}

template <class M>
__global__ void __launch_bounds__(128) velocity_march_kernel(const VelocityArgs a) {
  const int nx = a.g.nx, nz = a.g.nz, ld = a.g.ld;
  const int ld2n = ld >> 1;
  const int kp = blockIdx.x * blockDim.x + threadIdx.x;
  if (kp >= ld2n) return;
  const int z0 = kp * 2;
  const bool za = z0 >= 2 && z0 < nz - 2;
  const bool zb = z0 + 1 >= 2 && z0 + 1 < nz - 2;
  if (!(za || zb)) return;
  const int xs = max(blockIdx.y * a.strip, 2);
  const int xe = min((int)(blockIdx.y + 1) * a.strip, nx - 2);
  if (xs >= xe) return;

  size_t r = (size_t)xs * ld2n + kp;
  // rolling windows: txx rows x-2..x+1, txz rows x-1..x+2
  double2 txx_m2 = ld2(a.txx, r - 2 * (size_t)ld2n), txx_m1 = ld2(a.txx, r - ld2n), txx_0 = ld2(a.txx, r);
  double2 txz_m1 = ld2(a.txz, r - ld2n), txz_0 = ld2(a.txz, r), txz_p1 = ld2(a.txz, r + ld2n);

#pragma unroll 2
  for (int x = xs; x < xe; ++x, r += ld2n) {
    const double2 txx_p1 = ld2(a.txx, r + ld2n);
    const double2 txz_p2 = ld2(a.txz, r + 2 * (size_t)ld2n);
    const double2 tzz_0 = ld2(a.tzz, r), tzzL = ld2(a.tzz, r - 1), tzzR = ld2(a.tzz, r + 1);
    const double2 txzL = ld2(a.txz, r - 1), txzR = ld2(a.txz, r + 1);
    const double2 b = ld2(a.b, r);
    double2 vx = ld2_rw(a.vx, r), vz = ld2_rw(a.vz, r);
    if (za) {
      const double dxm_txx = diff4<M>(a.k, txx_0.x, txx_m1.x, txx_m2.x, txx_p1.x);
      const double dzm_txz = diff4<M>(a.k, txz_0.x, txzL.y, txzL.x, txz_0.y);
      const double dxp_txz = diff4<M>(a.k, txz_p1.x, txz_0.x, txz_m1.x, txz_p2.x);
      const double dzp_tzz = diff4<M>(a.k, tzz_0.y, tzz_0.x, tzzL.y, tzzR.x);
      const double tp = taper_at(a.taper_table, a.g.np, x, z0, nx, nz);
      velocity_update<M>(a.k, dxm_txx, dzm_txz, dxp_txz, dzp_tzz, b.x, tp, vx.x, vz.x);
    }
    if (zb) {
      const double dxm_txx = diff4<M>(a.k, txx_0.y, txx_m1.y, txx_m2.y, txx_p1.y);
      const double dzm_txz = diff4<M>(a.k, txz_0.y, txz_0.x, txzL.y, txzR.x);
      const double dxp_txz = diff4<M>(a.k, txz_p1.y, txz_0.y, txz_m1.y, txz_p2.y);
      const double dzp_tzz = diff4<M>(a.k, tzzR.x, tzz_0.y, tzz_0.x, tzzR.y);
      const double tp = taper_at(a.taper_table, a.g.np, x, z0 + 1, nx, nz);
      velocity_update<M>(a.k, dxm_txx, dzm_txz, dxp_txz, dzp_tzz, b.y, tp, vx.y, vz.y);
    }
    st2(a.vx, r, vx);
    st2(a.vz, r, vz);
    txx_m2 = txx_m1;
    txx_m1 = txx_0;
    txx_0 = txx_p1;
    txz_m1 = txz_0;
    txz_0 = txz_p1;
    txz_p1 = txz_p2;
  }
}

}  // namespace psv
