// Optional FP32 wavefield mode ("fp32" = 1): ONE launch per time step, like variant 2/3, with the
// wavefield (vx vz txx tzz txz), the medium the stencil reads (lm mu b) and the stored snapshots
// in single precision -- 52 B per grid-point update instead of 104 B.  Everything that is summed
// over many steps stays in double: seismograms u[t] (src/fdModel.cpp:576-596), the lambda/mu/rho
// sensitivity kernels (:817-861, correlated in double from the float states), and everything
// outside the time loop (model, misfit, adjoint sources, map, gradient vector).
// The reference offers the same switch as a compile-time typedef (`real_simulation`,
// src/fdModel.h:17-21).  Stated tolerance against the FP64 reference: 1e-3 relative L2 on
// synthetics, kernels and gradient (tests/test_gpu_fp32.py; measured ~1e-5..1e-4).
//
// Same march as fused.cuh (a warp owns 28 z-pairs of a strip of rows, rolling register windows
// along x, z-neighbours by warp shuffle, new stresses consumed from registers), restated on
// float2 pairs with loads straight from global memory; see fused.cuh for the geometry.
#pragma once

#include "fused.cuh"

namespace psv {

struct Consts32 {
  float sdt, rdx, rdz, c1, c2;
};

__device__ __forceinline__ float2 ld2f(const float *p, long long pair) {
  return __ldg(reinterpret_cast<const float2 *>(p) + pair);
}
__device__ __forceinline__ void st2f(float *p, long long pair, float2 v) {
  *(reinterpret_cast<float2 *>(p) + pair) = v;
}
__device__ __forceinline__ float shfl_up1f(float v) { return __shfl_up_sync(0xffffffffu, v, 1); }
__device__ __forceinline__ float shfl_dn1f(float v) { return __shfl_down_sync(0xffffffffu, v, 1); }
__device__ __forceinline__ float diff4f(const Consts32 &k, float a, float b, float c, float d) {
  return k.c1 * (a - b) + k.c2 * (c - d);
}
__device__ __forceinline__ void stress_update_f(const Consts32 &k, float dxp_vx, float dzm_vz, float dzp_vx,
                                                float dxm_vz, float lm, float mu, float tp, float &txx, float &tzz,
                                                float &txz) {
  const float la = lm - 2.0f * mu;
  const float ex = dxp_vx * k.rdx, ez = dzm_vz * k.rdz;
  txx = tp * (txx + k.sdt * (lm * ex + la * ez));
  tzz = tp * (tzz + k.sdt * (la * ex + lm * ez));
  txz = tp * (txz + (k.sdt * mu) * (dzp_vx * k.rdz + dxm_vz * k.rdx));
}
__device__ __forceinline__ void velocity_update_f(const Consts32 &k, float dxm_txx, float dzm_txz, float dxp_txz,
                                                  float dzp_tzz, float b, float tp, float &vx, float &vz) {
  const float bs = b * k.sdt;
  vx = tp * (vx + bs * (dxm_txx * k.rdx + dzm_txz * k.rdz));
  vz = tp * (vz + bs * (dxp_txz * k.rdx + dzp_tzz * k.rdz));
}
__device__ __forceinline__ float taper_axis_f(const double *__restrict__ table, int np, int i, int n) {
  return (float)taper_axis(table, np, i, n);
}

template <int MODE>
__global__ void __launch_bounds__(128) fused32_step_kernel(const FusedCommon c, const FusedModel32 *__restrict__ models) {
  const int lane = threadIdx.x & 31;
  const int zw = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (zw >= c.n_zwarps) return;  // warp-uniform
  const FusedModel32 &m = models[blockIdx.z];
  const int kp = zw * kFusedOwnedPairs - 2 + lane;
  const int xs = blockIdx.y * c.strip;
  const int xe = min(xs + c.strip, c.g.nx);
  const long long ld2n = c.g.ld >> 1;
  const bool own = lane >= 2 && lane < 30 && kp < (int)ld2n;
  const float *ovx = m.f[c.parity][0], *ovz = m.f[c.parity][1], *oxx = m.f[c.parity][2], *ozz = m.f[c.parity][3],
              *oxz = m.f[c.parity][4];
  float *nvx = m.f[c.parity ^ 1][0], *nvz = m.f[c.parity ^ 1][1], *nxx = m.f[c.parity ^ 1][2],
        *nzz = m.f[c.parity ^ 1][3], *nxz = m.f[c.parity ^ 1][4];
  const Consts32 k{(float)c.k.sdt, (float)c.k.rdx, (float)c.k.rdz, (float)c.k.c1, (float)c.k.c2};

  const int nz = c.g.nz;
  const int gx0 = c.g.x_off, gnx = c.g.nx_global;
  const int z0 = kp * 2;
  const bool za = z0 >= 2 && z0 < nz - 2;
  const bool zb = z0 + 1 >= 2 && z0 + 1 < nz - 2;
  const float tza = taper_axis_f(c.taper_table, c.g.np, z0, nz), tzb = taper_axis_f(c.taper_table, c.g.np, z0 + 1, nz);

  bool strip_rec = false, strip_ops = false;
  if (c.point_mode != POINT_NONE) {
    strip_rec = c.rec_row_start[xe] > c.rec_row_start[xs];
    if (c.point_mode == POINT_FORWARD) strip_ops = c.op_row_start[xe] > c.op_row_start[xs];
  }

  int j = xs - 2;
  long long r = (long long)j * ld2n + kp;
  float2 vx_m2 = ld2f(ovx, r - 2 * ld2n), vx_m1 = ld2f(ovx, r - ld2n), vx_0 = ld2f(ovx, r), vx_p1 = ld2f(ovx, r + ld2n);
  float2 vz_m2 = ld2f(ovz, r - 2 * ld2n), vz_m1 = ld2f(ovz, r - ld2n), vz_0 = ld2f(ovz, r);
  const float2 zero2 = make_float2(0.f, 0.f);
  float2 xx_m4 = zero2, xx_m3 = zero2, xx_m2 = zero2, xx_m1 = zero2;
  float2 xz_m3 = zero2, xz_m2 = zero2, xz_m1 = zero2;
  float2 zz_m2 = zero2, zz_m1 = zero2;

  // software pipeline: the 8 loads of row j+1 are issued before row j is computed, so a warp always
  // has one row of memory latency in flight (the FP64 default gets the same from its TMA ring)
  float2 n_vx = ld2f(ovx, r + 2 * ld2n), n_vz = ld2f(ovz, r + ld2n), n_xx = ld2f(oxx, r), n_zz = ld2f(ozz, r),
         n_xz = ld2f(oxz, r), n_lm = ld2f(m.lm, r), n_mu = ld2f(m.mu, r), n_b = ld2f(m.b, r - 2 * ld2n);
#pragma unroll 1
  for (; j < xe + 2; ++j, r += ld2n) {
    const float2 vx_p2 = n_vx, vz_p1 = n_vz;
    float2 txx = n_xx, tzz = n_zz, txz = n_xz;
    const float2 lm = n_lm, mu = n_mu;
    const float2 b = n_b;
    {  // rows past the strip end fall into the guard band / the next strip's rows: harmless reads
      const long long rn = r + ld2n;
      n_vx = ld2f(ovx, rn + 2 * ld2n);
      n_vz = ld2f(ovz, rn + ld2n);
      n_xx = ld2f(oxx, rn);
      n_zz = ld2f(ozz, rn);
      n_xz = ld2f(oxz, rn);
      n_lm = ld2f(m.lm, rn);
      n_mu = ld2f(m.mu, rn);
      n_b = ld2f(m.b, rn - 2 * ld2n);
    }
    const bool row_mine = j >= xs && j < xe;  // warp-uniform
    const float tr_j = taper_axis_f(c.taper_table, c.g.np, gx0 + j, gnx),
                tr_x = taper_axis_f(c.taper_table, c.g.np, gx0 + j - 2, gnx);

    if (row_mine) {
      if (MODE == MODE_SNAPSHOT && own) {
        float *s = m.snap + c.snap_offset;
        st2f(s, r, vx_0);
        st2f(s + c.snap_field_stride, r, vz_0);
        st2f(s + 2 * c.snap_field_stride, r, txx);
        st2f(s + 3 * c.snap_field_stride, r, tzz);
        st2f(s + 4 * c.snap_field_stride, r, txz);
      }
      if (MODE == MODE_ACCUMULATE) {
        if (own && j >= c.wx0 && j < c.wx1 && z0 + 1 >= c.wz0 && z0 < c.wz1) {
          const float *s = m.snap + c.snap_offset;
          const float2 fvx = ld2f(s, r), fvz = ld2f(s + c.snap_field_stride, r), fxx = ld2f(s + 2 * c.snap_field_stride, r),
                       fzz = ld2f(s + 3 * c.snap_field_stride, r), fxz = ld2f(s + 4 * c.snap_field_stride, r);
          double2 kr = *(reinterpret_cast<const double2 *>(m.k_rho) + r), kl = *(reinterpret_cast<const double2 *>(m.k_lam) + r),
                  km = *(reinterpret_cast<const double2 *>(m.k_mu) + r);
          if (z0 >= c.wz0 && z0 < c.wz1)
            kernel_accumulate<FastMath>(c.sidt, lm.x, mu.x, fvx.x, fvz.x, fxx.x, fzz.x, fxz.x, vx_0.x, vz_0.x, txx.x, tzz.x,
                                        txz.x, kr.x, kl.x, km.x);
          if (z0 + 1 >= c.wz0 && z0 + 1 < c.wz1)
            kernel_accumulate<FastMath>(c.sidt, lm.y, mu.y, fvx.y, fvz.y, fxx.y, fzz.y, fxz.y, vx_0.y, vz_0.y, txx.y, tzz.y,
                                        txz.y, kr.y, kl.y, km.y);
          st2s(m.k_rho, r, kr);
          st2s(m.k_lam, r, kl);
          st2s(m.k_mu, r, km);
        }
      }
      if (c.point_mode == POINT_FORWARD && strip_rec) {  // receiver sampling, running sum in double
        const int q1 = c.rec_row_start[j + 1];
        for (int q = c.rec_row_start[j]; q < q1; ++q) {
          const int z = c.rec_z[q];
          if (own && (z >> 1) == kp) {
            const size_t t = (size_t)c.rec_id[q] * c.nt + c.it;
            const double sx = __ddiv_rn(__dmul_rn(c.dt, (double)((z & 1) ? vx_0.y : vx_0.x)), c.dxdz);
            const double sz = __ddiv_rn(__dmul_rn(c.dt, (double)((z & 1) ? vz_0.y : vz_0.x)), c.dxdz);
            if (c.it == 0) {
              m.ux[t] = sx;
              m.uz[t] = sz;
            } else {
              m.ux[t] = __dadd_rn(m.ux[t - 1], sx);
              m.uz[t] = __dadd_rn(m.uz[t - 1], sz);
            }
          }
        }
      }
    }

    {  // ---- stress update of row j ----
      const float vxL_y = shfl_up1f(vx_0.y), vxR_x = shfl_dn1f(vx_0.x), vxR_y = shfl_dn1f(vx_0.y);
      const float vzL_x = shfl_up1f(vz_0.x), vzL_y = shfl_up1f(vz_0.y), vzR_x = shfl_dn1f(vz_0.x);
      const bool rowok = gx0 + j >= 2 && gx0 + j < gnx - 2;
      float axx = txx.x, azz = tzz.x, axz = txz.x, bxx = txx.y, bzz = tzz.y, bxz = txz.y;
      {
        const float dxp_vx = diff4f(k, vx_p1.x, vx_0.x, vx_m1.x, vx_p2.x);
        const float dzm_vz = diff4f(k, vz_0.x, vzL_y, vzL_x, vz_0.y);
        const float dzp_vx = diff4f(k, vx_0.y, vx_0.x, vxL_y, vxR_x);
        const float dxm_vz = diff4f(k, vz_0.x, vz_m1.x, vz_m2.x, vz_p1.x);
        stress_update_f(k, dxp_vx, dzm_vz, dzp_vx, dxm_vz, lm.x, mu.x, fminf(tr_j, tza), axx, azz, axz);
      }
      {
        const float dxp_vx = diff4f(k, vx_p1.y, vx_0.y, vx_m1.y, vx_p2.y);
        const float dzm_vz = diff4f(k, vz_0.y, vz_0.x, vzL_y, vzR_x);
        const float dzp_vx = diff4f(k, vxR_x, vx_0.y, vx_0.x, vxR_y);
        const float dxm_vz = diff4f(k, vz_0.y, vz_m1.y, vz_m2.y, vz_p1.y);
        stress_update_f(k, dxp_vx, dzm_vz, dzp_vx, dxm_vz, lm.y, mu.y, fminf(tr_j, tzb), bxx, bzz, bxz);
      }
      const bool ua = rowok && za, ub = rowok && zb;
      txx.x = ua ? axx : txx.x;
      tzz.x = ua ? azz : tzz.x;
      txz.x = ua ? axz : txz.x;
      txx.y = ub ? bxx : txx.y;
      tzz.y = ub ? bzz : tzz.y;
      txz.y = ub ? bxz : txz.y;
    }
    if (own && row_mine) {
      st2f(nxx, r, txx);
      st2f(nzz, r, tzz);
      st2f(nxz, r, txz);
    }

    const int x = j - 2;  // ---- velocity update of row x ----
    if (x >= xs) {
      const long long rv = r - 2 * ld2n;
      float2 vx = vx_m2, vz = vz_m2;
      {
        const float xzL_x = shfl_up1f(xz_m2.x), xzL_y = shfl_up1f(xz_m2.y), xzR_x = shfl_dn1f(xz_m2.x);
        const float zzL_y = shfl_up1f(zz_m2.y), zzR_x = shfl_dn1f(zz_m2.x), zzR_y = shfl_dn1f(zz_m2.y);
        const bool rowok = gx0 + x >= 2 && gx0 + x < gnx - 2;
        float avx = vx.x, avz = vz.x, bvx = vx.y, bvz = vz.y;
        {
          const float dxm_txx = diff4f(k, xx_m2.x, xx_m3.x, xx_m4.x, xx_m1.x);
          const float dzm_txz = diff4f(k, xz_m2.x, xzL_y, xzL_x, xz_m2.y);
          const float dxp_txz = diff4f(k, xz_m1.x, xz_m2.x, xz_m3.x, txz.x);
          const float dzp_tzz = diff4f(k, zz_m2.y, zz_m2.x, zzL_y, zzR_x);
          velocity_update_f(k, dxm_txx, dzm_txz, dxp_txz, dzp_tzz, b.x, fminf(tr_x, tza), avx, avz);
        }
        {
          const float dxm_txx = diff4f(k, xx_m2.y, xx_m3.y, xx_m4.y, xx_m1.y);
          const float dzm_txz = diff4f(k, xz_m2.y, xz_m2.x, xzL_y, xzR_x);
          const float dxp_txz = diff4f(k, xz_m1.y, xz_m2.y, xz_m3.y, txz.y);
          const float dzp_tzz = diff4f(k, zzR_x, zz_m2.y, zz_m2.x, zzR_y);
          velocity_update_f(k, dxm_txx, dzm_txz, dxp_txz, dzp_tzz, b.y, fminf(tr_x, tzb), bvx, bvz);
        }
        const bool ua = rowok && za, ub = rowok && zb;
        vx.x = ua ? avx : vx.x;
        vz.x = ua ? avz : vz.x;
        vx.y = ub ? bvx : vx.y;
        vz.y = ub ? bvz : vz.y;
      }
      if (strip_ops) {  // moment-tensor injection, amplitude in double (src/fdModel.cpp:681-751)
        const int q1 = c.op_row_start[x + 1];
        for (int q = c.op_row_start[x]; q < q1; ++q) {
          const RowOp op = c.ops[q];
          if (own && (op.z >> 1) == kp) {
            const double w = c.stf[(size_t)op.source * c.nt + c.it];
            const double bb = (double)((op.z & 1) ? b.y : b.x);
            const double amp = ((op.weight * w) * c.dt * bb) * op.rd4;
            const bool odd = op.z & 1;
            float &t = op.field == 0 ? (odd ? vx.y : vx.x) : (odd ? vz.y : vz.x);
            t = (float)(op.sign > 0 ? (double)t + amp : (double)t - amp);
          }
        }
      }
      if (c.point_mode == POINT_ADJOINT && strip_rec) {  // adjoint sources at the receivers (:951-958)
        const int q1 = c.rec_row_start[x + 1];
        for (int q = c.rec_row_start[x]; q < q1; ++q) {
          const int z = c.rec_z[q];
          if (own && (z >> 1) == kp) {
            const size_t t = (size_t)c.rec_id[q] * c.nt + c.it;
            const double bb = (double)((z & 1) ? b.y : b.x);
            const double ax = ((c.dt * bb) * m.ax[t]) * c.rdxdz, az = ((c.dt * bb) * m.az[t]) * c.rdxdz;
            if (z & 1) {
              vx.y = (float)((double)vx.y + ax);
              vz.y = (float)((double)vz.y + az);
            } else {
              vx.x = (float)((double)vx.x + ax);
              vz.x = (float)((double)vz.x + az);
            }
          }
        }
      }
      if (own) {
        st2f(nvx, rv, vx);
        st2f(nvz, rv, vz);
      }
    }

    vx_m2 = vx_m1; vx_m1 = vx_0; vx_0 = vx_p1; vx_p1 = vx_p2;
    vz_m2 = vz_m1; vz_m1 = vz_0; vz_0 = vz_p1;
    xx_m4 = xx_m3; xx_m3 = xx_m2; xx_m2 = xx_m1; xx_m1 = txx;
    xz_m3 = xz_m2; xz_m2 = xz_m1; xz_m1 = txz;
    zz_m2 = zz_m1; zz_m1 = tzz;
  }
}

}  // namespace psv
