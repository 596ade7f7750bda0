// Point-wise arithmetic of the P-SV velocity-stress scheme, shared by every kernel variant.
//
// Two arithmetic policies:
//   ExactMath  every operation is an explicitly rounded IEEE-754 double operation
//              (__dmul_rn/__dadd_rn/__ddiv_rn are never contracted into FMAs) issued in the
//              reference's literal evaluation order (src/fdModel.cpp:615-642, 661-676,
//              832-858), so the result is BIT-IDENTICAL to the reference's x86-64 build.
//   FastMath   the same formulas with divisions by dx/dz folded into precomputed
//              reciprocals and FMA contraction allowed; differs from the reference at the
//              1e-15 relative level (budget: 1e-10), and keeps the kernels HBM-bound.
#pragma once

#include <cuda_runtime.h>

namespace psv {

struct ExactMath {
  static constexpr bool kExact = true;
  __device__ __forceinline__ static double mul(double a, double b) { return __dmul_rn(a, b); }
  __device__ __forceinline__ static double add(double a, double b) { return __dadd_rn(a, b); }
  __device__ __forceinline__ static double sub(double a, double b) { return __dsub_rn(a, b); }
  __device__ __forceinline__ static double div(double a, double b) { return __ddiv_rn(a, b); }
};

struct FastMath {
  static constexpr bool kExact = false;
  __device__ __forceinline__ static double mul(double a, double b) { return a * b; }
  __device__ __forceinline__ static double add(double a, double b) { return a + b; }
  __device__ __forceinline__ static double sub(double a, double b) { return a - b; }
  __device__ __forceinline__ static double div(double a, double b) { return a / b; }
};

// Scalars of one sweep.  sdt = +dt (forward) or -dt (adjoint): f - dt*X == f + (-dt)*X
// bit for bit, so one body serves both time directions (src/fdModel.cpp:882-946).
struct SweepConsts {
  double sdt;       // signed time step
  double dx, dz;    // grid spacing (ExactMath divides by these)
  double rdx, rdz;  // reciprocals (FastMath)
  double c1, c2;    // 9/8, 1/24 (src/fdModel.h:201-202)
};

// Staggered 4th-order difference: c1*(a - b) + c2*(c - d).  The default arithmetic spells its FMAs out, so that
// every kernel (and every call site inside one) rounds the same way whatever the compiler would have contracted.
template <class M>
__device__ __forceinline__ double diff4(const SweepConsts &k, double a, double b, double c, double d) {
  if (M::kExact) return M::add(M::mul(k.c1, M::sub(a, b)), M::mul(k.c2, M::sub(c, d)));
  return __fma_rn(k.c1, __dsub_rn(a, b), __dmul_rn(k.c2, __dsub_rn(c, d)));
}

// Cerjan taper value of grid point (ix, iz): table[min(np, 1 + distance to nearest edge)]
// (src/fdModel.cpp:263-298 builds exactly this as a full array; the table holds the np+1
// distinct values, computed on the host with the reference's expression).
__device__ __forceinline__ double taper_at(const double *__restrict__ table, int np, int ix, int iz,
                                           int nx, int nz) {
  int d = min(min(ix, nx - 1 - ix), min(iz, nz - 1 - iz)) + 1;
  d = min(d, np);
  return __ldg(table + d);
}

// Stress update of one point (src/fdModel.cpp:615-642), one component at a time (the resident loop kernel updates
// only the components a halo row is read for).  la = lm - 2*mu is recomputed with the operation
// update_from_velocity uses (src/fdModel.cpp:1046), so it equals the reference's stored `la` bit for bit and the
// array need not be read.
//   dxp_vx = D+x vx, dzm_vz = D-z vz, dzp_vx = D+z vx, dxm_vz = D-x vz
template <class M>
__device__ __forceinline__ double stress_txx(const SweepConsts &k, double dxp_vx, double dzm_vz, double lm, double mu,
                                             double tp, double txx) {
  const double la = M::sub(lm, M::mul(2.0, mu));
  if (M::kExact)
    return M::mul(tp, M::add(txx, M::mul(k.sdt, M::add(M::div(M::mul(lm, dxp_vx), k.dx),
                                                        M::div(M::mul(la, dzm_vz), k.dz)))));
  const double ex = __dmul_rn(dxp_vx, k.rdx), ez = __dmul_rn(dzm_vz, k.rdz);
  return __dmul_rn(tp, __fma_rn(k.sdt, __fma_rn(lm, ex, __dmul_rn(la, ez)), txx));
}
template <class M>
__device__ __forceinline__ double stress_tzz(const SweepConsts &k, double dxp_vx, double dzm_vz, double lm, double mu,
                                             double tp, double tzz) {
  const double la = M::sub(lm, M::mul(2.0, mu));
  if (M::kExact)
    return M::mul(tp, M::add(tzz, M::mul(k.sdt, M::add(M::div(M::mul(la, dxp_vx), k.dx),
                                                        M::div(M::mul(lm, dzm_vz), k.dz)))));
  const double ex = __dmul_rn(dxp_vx, k.rdx), ez = __dmul_rn(dzm_vz, k.rdz);
  return __dmul_rn(tp, __fma_rn(k.sdt, __fma_rn(la, ex, __dmul_rn(lm, ez)), tzz));
}
template <class M>
__device__ __forceinline__ double stress_txz(const SweepConsts &k, double dzp_vx, double dxm_vz, double mu, double tp,
                                             double txz) {
  if (M::kExact)
    return M::mul(tp, M::add(txz, M::mul(M::mul(k.sdt, mu), M::add(M::div(dzp_vx, k.dz), M::div(dxm_vz, k.dx)))));
  return __dmul_rn(tp, __fma_rn(__dmul_rn(k.sdt, mu), __fma_rn(dzp_vx, k.rdz, __dmul_rn(dxm_vz, k.rdx)), txz));
}
template <class M>
__device__ __forceinline__ void stress_update(const SweepConsts &k, double dxp_vx, double dzm_vz,
                                              double dzp_vx, double dxm_vz, double lm, double mu,
                                              double tp, double &txx, double &tzz, double &txz) {
  txx = stress_txx<M>(k, dxp_vx, dzm_vz, lm, mu, tp, txx);
  tzz = stress_tzz<M>(k, dxp_vx, dzm_vz, lm, mu, tp, tzz);
  txz = stress_txz<M>(k, dzp_vx, dxm_vz, mu, tp, txz);
}

// Velocity update of one point (src/fdModel.cpp:661-676); b = 1/rho (b_vx == b_vz).  Both components have the form
// v = tp * (v + b*sdt * (dx_term/dx + dz_term/dz)):  vx: dx_term = D-x txx, dz_term = D-z txz;
// vz: dx_term = D+x txz, dz_term = D+z tzz.
template <class M>
__device__ __forceinline__ double velocity_component(const SweepConsts &k, double dx_term, double dz_term, double b,
                                                     double tp, double v) {
  if (M::kExact) {
    const double bs = M::mul(b, k.sdt);
    return M::mul(tp, M::add(v, M::mul(bs, M::add(M::div(dx_term, k.dx), M::div(dz_term, k.dz)))));
  }
  const double bs = __dmul_rn(b, k.sdt);
  return __dmul_rn(tp, __fma_rn(bs, __fma_rn(dx_term, k.rdx, __dmul_rn(dz_term, k.rdz)), v));
}
//   dxm_txx = D-x txx, dzm_txz = D-z txz, dxp_txz = D+x txz, dzp_tzz = D+z tzz
template <class M>
__device__ __forceinline__ void velocity_update(const SweepConsts &k, double dxm_txx, double dzm_txz,
                                                double dxp_txz, double dzp_tzz, double b, double tp,
                                                double &vx, double &vz) {
  vx = velocity_component<M>(k, dxm_txx, dzm_txz, b, tp, vx);
  vz = velocity_component<M>(k, dxp_txz, dzp_tzz, b, tp, vz);
}

// Sensitivity-kernel accumulation of one point (src/fdModel.cpp:832-858): forward snapshot
// (f*) against the adjoint state *before* this step's update.  sidt = snapshot_interval*dt.
template <class M>
__device__ __forceinline__ void kernel_accumulate(double sidt, double lm, double mu, double fvx,
                                                  double fvz, double fxx, double fzz, double fxz,
                                                  double vx, double vz, double txx, double tzz,
                                                  double txz, double &k_rho, double &k_lam,
                                                  double &k_mu) {
  const double la = M::sub(lm, M::mul(2.0, mu));
  k_rho = M::sub(k_rho, M::mul(sidt, M::add(M::mul(fvx, vx), M::mul(fvz, vz))));
  if (!M::kExact) {
    // the same formulas with the 8 divisions folded into 3 reciprocals of medium-only terms
    const double r = la * (1.0 / lm);
    const double F1 = fxx - fzz * r, F2 = fzz - fxx * r, A1 = txx - tzz * r, A2 = tzz - txx * r;
    const double d0 = lm - la * r;
    const double rden = 1.0 / (d0 * d0);
    k_lam = k_lam + (sidt * ((F1 + F2) * (A1 + A2))) * rden;
    const double t1 = (A1 * F1 + A2 * F2) * rden;
    const double t2 = (txz * fxz) * (0.5 / (mu * mu));
    k_mu = k_mu + (sidt * 2.0) * (t1 + t2);
    return;
  }
  // q(a, b) = a - (b*la)/lm
  const double F1 = M::sub(fxx, M::div(M::mul(fzz, la), lm));
  const double F2 = M::sub(fzz, M::div(M::mul(fxx, la), lm));
  const double A1 = M::sub(txx, M::div(M::mul(tzz, la), lm));
  const double A2 = M::sub(tzz, M::div(M::mul(txx, la), lm));
  const double d0 = M::sub(lm, M::div(M::mul(la, la), lm));
  const double den = M::mul(d0, d0);
  k_lam = M::add(k_lam, M::div(M::mul(sidt, M::mul(M::add(F1, F2), M::add(A1, A2))), den));
  const double t1 = M::div(M::add(M::mul(A1, F1), M::mul(A2, F2)), den);
  const double t2 = M::mul(2.0, M::div(M::mul(txz, fxz), M::mul(M::mul(4.0, mu), mu)));
  k_mu = M::add(k_mu, M::mul(M::mul(sidt, 2.0), M::add(t1, t2)));
}

}  // namespace psv
