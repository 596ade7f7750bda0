// Point operations and point-wise array kernels around the sweeps.  All of them use
// explicitly rounded IEEE operations in the reference's evaluation order (they touch a
// handful of values per step, or run once per model update), so they are bit-identical to
// the reference regardless of the arithmetic policy of the sweeps.
#pragma once

#include <cuda_runtime.h>

#include "fused_types.cuh"

namespace psv {

// ---- receiver sampling (src/fdModel.cpp:576-596) -------------------------------------------
// u[it] = u[it-1] + dt*v/(dx*dz) for every receiver of the shot; one thread per receiver.
struct RecordArgs {
  const double *vx, *vz;
  const int *rec_index;  // linear (pitched) grid index per receiver
  double *ux, *uz;       // rows of this shot: [nr][nt]
  int nr, nt, it;
  double dt, dxdz;
};

__global__ void record_receivers_kernel(const RecordArgs a) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= a.nr) return;
  const int p = a.rec_index[r];
  const size_t q = (size_t)r * a.nt + a.it;
  const double sx = __ddiv_rn(__dmul_rn(a.dt, a.vx[p]), a.dxdz);
  const double sz = __ddiv_rn(__dmul_rn(a.dt, a.vz[p]), a.dxdz);
  if (a.it == 0) {
    a.ux[q] = sx;
    a.uz[q] = sz;
  } else {
    a.ux[q] = __dadd_rn(a.ux[q - 1], sx);
    a.uz[q] = __dadd_rn(a.uz[q - 1], sz);
  }
}

// Running sum of src/fdModel.cpp:584-594 over the increments dt*v/(dx*dz) the fused step kernels
// stored: u[it] = u[it-1] + s[it], one thread per trace, the reference's sequential order and
// rounding (bit-identical to summing inside the time loop).
__global__ void receiver_prefix_kernel(double *ux, double *uz, int nr, int nt) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= 2 * nr) return;
  double *u = (r < nr ? ux : uz) + (size_t)(r % nr) * nt;
  double acc = u[0];
  for (int it = 1; it < nt; ++it) {
    acc = __dadd_rn(acc, u[it]);
    u[it] = acc;
  }
}

// ---- moment-tensor source injection (src/fdModel.cpp:681-751) ------------------------------
// The host flattens the shot's sources into point operations in the reference's order
// (source list order, then the 12 updates of a source in the order they are written) and
// chains the operations that hit the same (field, point): one thread walks one chain, so
// coinciding updates are applied in exactly the reference's sequence without atomics.
struct InjectOp {
  int point;      // linear (pitched) index of the target point
  int field;      // 0 = vx, 1 = vz
  int source;     // row of stf
  int next;       // next op of the same chain, -1 = end
  double weight;  // (+-)moment entry, already multiplied by 0.25 where the reference does
  double sign;    // +1 / -1  (reference uses += / -=)
  double d4;      // dx^4 or dz^4
};

struct InjectArgs {
  double *vx, *vz;
  const double *b;      // b_vz at the target point (the reference uses b_vz for both fields)
  const double *stf;    // [n_sources][nt]
  const InjectOp *ops;
  const int *heads;     // first op of every chain
  int n_heads, nt, it;
  double dt;
};

__global__ void inject_sources_kernel(const InjectArgs a) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= a.n_heads) return;
  int o = a.heads[c];
  const InjectOp first = a.ops[o];
  double *f = first.field == 0 ? a.vx : a.vz;
  double v = f[first.point];
  const double b = a.b[first.point];
  while (o >= 0) {
    const InjectOp op = a.ops[o];
    const double w = a.stf[(size_t)op.source * a.nt + a.it];
    // moment*stf*dt*b/d4, left to right
    const double amp = __ddiv_rn(__dmul_rn(__dmul_rn(__dmul_rn(op.weight, w), a.dt), b), op.d4);
    v = op.sign > 0 ? __dadd_rn(v, amp) : __dsub_rn(v, amp);
    o = op.next;
  }
  f[first.point] = v;
}

// ---- adjoint source injection (src/fdModel.cpp:951-958) ------------------------------------
// v[rec] += dt*b*a_stf[it]/(dx*dz); receivers sharing a grid point are chained in receiver
// order (one thread per chain).
struct AdjointInjectArgs {
  double *vx, *vz;
  const double *b;
  const double *ax, *az;  // adjoint sources of this shot: [nr][nt]
  const int *rec_index, *rec_next, *rec_heads;
  int n_heads, nt, it;
  double dt, dxdz;
};

__global__ void inject_adjoint_kernel(const AdjointInjectArgs a) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= a.n_heads) return;
  int r = a.rec_heads[c];
  const int p = a.rec_index[r];
  const double b = a.b[p];
  double vx = a.vx[p], vz = a.vz[p];
  while (r >= 0) {
    const size_t q = (size_t)r * a.nt + a.it;
    vx = __dadd_rn(vx, __ddiv_rn(__dmul_rn(__dmul_rn(a.dt, b), a.ax[q]), a.dxdz));
    vz = __dadd_rn(vz, __ddiv_rn(__dmul_rn(__dmul_rn(a.dt, b), a.az[q]), a.dxdz));
    r = a.rec_next[r];
  }
  a.vx[p] = vx;
  a.vz[p] = vz;
}

// ---- medium (src/fdModel.cpp:1034-1051) ----------------------------------------------------
// lm = vp^2*rho, mu = vs^2*rho, b = 1/rho.  (la = lm - 2mu and b_vz = b_vx are derived.)
__global__ void update_from_velocity_kernel(const double *vp, const double *vs, const double *rho,
                                            double *lm, double *mu, double *b, size_t n) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double r = rho[i], p = vp[i], s = vs[i];
  mu[i] = __dmul_rn(__dmul_rn(s, s), r);
  lm[i] = __dmul_rn(__dmul_rn(p, p), r);
  b[i] = __ddiv_rn(1.0, r);
}

// ---- kernels (lambda, mu, rho_l) -> (vp, vs, rho) (src/fdModel.cpp:1164-1183) ---------------
__global__ void map_kernels_kernel(const double *vp, const double *vs, const double *b,
                                   const double *k_lam, const double *k_mu, const double *k_rho,
                                   double *k_vp, double *k_vs, double *k_rhov, size_t n) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double p = vp[i], s = vs[i], bb = b[i], kl = k_lam[i], km = k_mu[i];
  k_vp[i] = __ddiv_rn(__dmul_rn(__dmul_rn(2.0, p), kl), bb);
  k_vs[i] = __ddiv_rn(__dsub_rn(__dmul_rn(__dmul_rn(2.0, s), km), __dmul_rn(__dmul_rn(4.0, s), kl)), bb);
  const double pp = __dmul_rn(p, p), ss2 = __dmul_rn(__dmul_rn(2.0, s), s), ss = __dmul_rn(s, s);
  k_rhov[i] = __dadd_rn(__dadd_rn(k_rho[i], __dmul_rn(__dsub_rn(pp, ss2), kl)), __dmul_rn(ss, km));
}

// ---- L2 adjoint sources (src/fdModel.cpp:1147-1162) ----------------------------------------
__global__ void adjoint_sources_kernel(const double *ux, const double *uz, const double *ox,
                                       const double *oz, double *ax, double *az, size_t n) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  ax[i] = __dsub_rn(ux[i], ox[i]);
  az[i] = __dsub_rn(uz[i], oz[i]);
}

// ---- L2 misfit (src/fdModel.cpp:1128-1145) ---------------------------------------------------
// Deterministic two-stage sum: every block adds a fixed contiguous chunk (per-thread serial
// partials, then a fixed-shape tree), the host adds the block partials in block order.  The
// reference adds serially; the association differs, the terms (0.5*dt*d^2) do not.
constexpr int kMisfitThreads = 256;
constexpr int kMisfitPerThread = 16;

__global__ void misfit_partial_kernel(const double *ux, const double *uz, const double *ox,
                                      const double *oz, size_t n, double half_dt, double *partial) {
  __shared__ double sh[kMisfitThreads];
  const size_t base = ((size_t)blockIdx.x * kMisfitThreads + threadIdx.x) * kMisfitPerThread;
  double acc = 0.0;
  for (int j = 0; j < kMisfitPerThread; ++j) {
    const size_t i = base + j;
    if (i < n) {
      const double dx = __dsub_rn(ox[i], ux[i]), dz = __dsub_rn(oz[i], uz[i]);
      acc = __dadd_rn(acc, __dmul_rn(half_dt, __dmul_rn(dx, dx)));
      acc = __dadd_rn(acc, __dmul_rn(half_dt, __dmul_rn(dz, dz)));
    }
  }
  sh[threadIdx.x] = acc;
  __syncthreads();
  for (int s = kMisfitThreads / 2; s > 0; s >>= 1) {
    if (threadIdx.x < s) sh[threadIdx.x] = __dadd_rn(sh[threadIdx.x], sh[threadIdx.x + s]);
    __syncthreads();
  }
  if (threadIdx.x == 0) partial[blockIdx.x] = sh[0];
}

// ---- model / gradient vectors (src/fdModel.cpp:1358-1515) -----------------------------------
// One thread per basis block (bx, bz); entry index bx + bz*nbx (x fastest).  The block's
// points are visited sub_ix outer / sub_iz inner exactly like the reference, so block means
// and block sums are bit-identical.
struct VectorArgs {
  int nbx, nbz, npx, npz;  // blocks per direction, points per block
  int x0, z0;              // first grid point of the free window
  int ld;
  size_t per;              // entries per parameter = nbx*nbz
};

enum VectorMode { VEC_MEAN = 0, VEC_SUM = 1, VEC_FILL = 2 };

template <int MODE>
__global__ void vector_kernel(const VectorArgs a, double *f0, double *f1, double *f2, double *v) {
  const size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= a.per) return;
  const int bx = (int)(k % a.nbx), bz = (int)(k / a.nbx);
  const int gx = bx * a.npx + a.x0, gz = bz * a.npz + a.z0;
  const double cnt = (double)(a.npx * a.npz);
  double s0 = 0.0, s1 = 0.0, s2 = 0.0;
  if (MODE == VEC_FILL) {
    s0 = v[k];
    s1 = v[k + a.per];
    s2 = v[k + 2 * a.per];
  }
  for (int sx = 0; sx < a.npx; ++sx) {
    for (int sz = 0; sz < a.npz; ++sz) {
      const size_t i = (size_t)(gx + sx) * a.ld + (gz + sz);
      if (MODE == VEC_MEAN) {
        s0 = __dadd_rn(s0, __ddiv_rn(f0[i], cnt));
        s1 = __dadd_rn(s1, __ddiv_rn(f1[i], cnt));
        s2 = __dadd_rn(s2, __ddiv_rn(f2[i], cnt));
      } else if (MODE == VEC_SUM) {
        s0 = __dadd_rn(s0, f0[i]);
        s1 = __dadd_rn(s1, f1[i]);
        s2 = __dadd_rn(s2, f2[i]);
      } else {
        f0[i] = s0;
        f1[i] = s1;
        f2[i] = s2;
      }
    }
  }
  if (MODE != VEC_FILL) {
    v[k] = s0;
    v[k + a.per] = s1;
    v[k + 2 * a.per] = s2;
  }
}

// Steepest-descent update of a device-resident model vector: m[k] <- m[k] - step * g[k], the
// `m -= step * g` of the reference's tests/python/test_descent.py:110 with explicitly rounded
// operations (no FMA), so the device loop is bit-identical to the host loop.
__global__ void descent_axpy_kernel(double *m, const double *g, const double step, const size_t n) {
  const size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k < n) m[k] = __dsub_rn(m[k], __dmul_rn(step, g[k]));
}

// Running sum of the receiver increments the marches stored (src/fdModel.cpp:584-594): u[it] = u[it-1] + s[it],
// one thread per trace and model (blockIdx.y), the reference's sequential order and rounding.
__global__ void receiver_prefix_models_kernel(const FusedModel *__restrict__ models, int nr, int nt) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= 2 * nr) return;
  const FusedModel &m = models[blockIdx.y];
  double *u = (r < nr ? m.ux : m.uz) + (size_t)(r % nr) * nt;
  double acc = u[0];
  for (int it = 1; it < nt; ++it) {
    acc = __dadd_rn(acc, u[it]);
    u[it] = acc;
  }
}

// double <-> float images of grid arrays (medium refresh, getters)
__global__ void to_float_kernel(const double *__restrict__ in, float *__restrict__ out, size_t n) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = (float)in[i];
}
__global__ void to_double_kernel(const float *__restrict__ in, double *__restrict__ out, size_t n) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = (double)in[i];
}

}  // namespace psv
