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
// float2 pairs; see fused.cuh for the geometry.  Two row sources, like variants 2 and 3 of the FP64 march: loads straight
// from global memory with the next row's loads issued one iteration ahead (fused32_step_kernel), and a TMA bulk-copy
// ring filled by a producer warp (fused32_pipe_kernel, the default: "variant" 3).
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

struct Row32 {  // what one row of the march consumes: vx of row j+2, vz of row j+1, the rest of row j (b of row j-2)
  float2 vx_p2, vz_p1, txx, tzz, txz, lm, mu, b;
};

// Row source 1: global memory, software-pipelined one row ahead.
struct GlobalRows32 {
  const float *ovx, *ovz, *oxx, *ozz, *oxz, *lm, *mu, *b;
  long long ld2n;
  Row32 next;
  __device__ __forceinline__ void load(long long r) {  // rows past the strip end fall into the guard band: harmless
    next.vx_p2 = ld2f(ovx, r + 2 * ld2n);
    next.vz_p1 = ld2f(ovz, r + ld2n);
    next.txx = ld2f(oxx, r);
    next.tzz = ld2f(ozz, r);
    next.txz = ld2f(oxz, r);
    next.lm = ld2f(lm, r);
    next.mu = ld2f(mu, r);
    next.b = ld2f(b, r - 2 * ld2n);
  }
  __device__ __forceinline__ void start(long long r) { load(r); }
  __device__ __forceinline__ void fetch(long long r, Row32 &d) {
    d = next;
    load(r + ld2n);
  }
  __device__ __forceinline__ void release() {}
};

// Row source 2: the shared-memory ring a producer lane fills with bulk-asynchronous copies (8 row segments per slot).
struct PipeRows32 {
  const unsigned char *ring;
  unsigned full0, empty0;
  int stages, seg, stage_bytes, lane_off, slot;
  unsigned phase;
  int lane;
  __device__ __forceinline__ void start(long long) {}
  __device__ __forceinline__ void fetch(long long, Row32 &d) {
    mbar_wait(full0 + 8 * slot, phase);
    const unsigned char *s = ring + (size_t)slot * stage_bytes + lane_off;
    d.vx_p2 = *reinterpret_cast<const float2 *>(s);
    d.vz_p1 = *reinterpret_cast<const float2 *>(s + seg);
    d.txx = *reinterpret_cast<const float2 *>(s + 2 * seg);
    d.tzz = *reinterpret_cast<const float2 *>(s + 3 * seg);
    d.txz = *reinterpret_cast<const float2 *>(s + 4 * seg);
    d.lm = *reinterpret_cast<const float2 *>(s + 5 * seg);
    d.mu = *reinterpret_cast<const float2 *>(s + 6 * seg);
    d.b = *reinterpret_cast<const float2 *>(s + 7 * seg);
  }
  __device__ __forceinline__ void release() {
    __syncwarp();
    if (lane == 0) mbar_arrive(empty0 + 8 * slot);
    if (++slot == stages) {
      slot = 0;
      phase ^= 1;
    }
  }
};

// One warp marches rows [xs, xe) for the z-pairs {kp}; lanes 2..29 own their pair.
template <int MODE, class Rows>
__device__ __forceinline__ void fused32_march(const FusedCommon &c, const FusedModel32 &m, Rows &rows, const int lane,
                                              const int kp, const int xs, const int xe, const bool own) {
  const long long ld2n = c.g.ld >> 1;
  const float *ovx = m.f[c.parity][0], *ovz = m.f[c.parity][1];
  float *nvx = m.f[c.parity ^ 1][0], *nvz = m.f[c.parity ^ 1][1], *nxx = m.f[c.parity ^ 1][2],
        *nzz = m.f[c.parity ^ 1][3], *nxz = m.f[c.parity ^ 1][4];
  const Consts32 k{(float)c.k.sdt, (float)c.k.rdx, (float)c.k.rdz, (float)c.k.c1, (float)c.k.c2};

  const int nz = c.g.nz;
  const int gx0 = c.g.x_off, gnx = c.g.nx_global;
  const int z0 = kp * 2;
  const bool za = z0 >= 2 && z0 < nz - 2;
  const bool zb = z0 + 1 >= 2 && z0 + 1 < nz - 2;
  const float tza = taper_axis_f(c.taper_table, c.g.np, z0, nz), tzb = taper_axis_f(c.taper_table, c.g.np, z0 + 1, nz);

  // Cursors into the row-sorted point lists (as in fused_march): the row of the next receiver / injection operation
  // sits in a register, the strip's receivers ride in registers one list entry per lane.
  int rq = 0, rq_end = 0, oq = 0, oq_end = 0;
  if (c.point_mode != POINT_NONE) {
    rq = c.rec_row_start[xs];
    rq_end = c.rec_row_start[xe];
    if (c.point_mode == POINT_FORWARD) {
      oq = c.op_row_start[xs];
      oq_end = c.op_row_start[xe];
    }
  }
  constexpr int kNoRow = 0x7fffffff;
  int op_next = oq < oq_end ? c.ops[oq].x : kNoRow;
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
        const size_t t = (size_t)lid * c.nt + c.it;
        lax = m.ax[t];
        laz = m.az[t];
      }
    }
  };
  if (rq < rq_end) load_window();
  int rec_next = __shfl_sync(0xffffffffu, lx, 0);
  auto next_receiver = [&]() {
    ++rq;
    if (rq >= rq_end) return kNoRow;
    if (rq - win0 == 32) {
      win0 = rq;
      load_window();
    }
    return __shfl_sync(0xffffffffu, lx, rq - win0);
  };

  int j = xs - 2;
  long long r = (long long)j * ld2n + kp;
  float2 vx_m2 = ld2f(ovx, r - 2 * ld2n), vx_m1 = ld2f(ovx, r - ld2n), vx_0 = ld2f(ovx, r), vx_p1 = ld2f(ovx, r + ld2n);
  float2 vz_m2 = ld2f(ovz, r - 2 * ld2n), vz_m1 = ld2f(ovz, r - ld2n), vz_0 = ld2f(ovz, r);
  const float2 zero2 = make_float2(0.f, 0.f);
  float2 xx_m4 = zero2, xx_m3 = zero2, xx_m2 = zero2, xx_m1 = zero2;
  float2 xz_m3 = zero2, xz_m2 = zero2, xz_m1 = zero2;
  float2 zz_m2 = zero2, zz_m1 = zero2;

  rows.start(r);
#pragma unroll 1
  for (; j < xe + 2; ++j, r += ld2n) {
    Row32 d;
    rows.fetch(r, d);
    const float2 vx_p2 = d.vx_p2, vz_p1 = d.vz_p1;
    float2 txx = d.txx, tzz = d.tzz, txz = d.txz;
    const float2 lm = d.lm, mu = d.mu;
    const float2 b = d.b;
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
      if (c.point_mode == POINT_FORWARD && j == rec_next) {  // receiver sampling, running sum in double
        do {
          const int z = __shfl_sync(0xffffffffu, lz, rq - win0), id = __shfl_sync(0xffffffffu, lid, rq - win0);
          if (own && (z >> 1) == kp) {
            const size_t t = (size_t)id * c.nt + c.it;
            const double sx = __ddiv_rn(__dmul_rn(c.dt, (double)((z & 1) ? vx_0.y : vx_0.x)), c.dxdz);
            const double sz = __ddiv_rn(__dmul_rn(c.dt, (double)((z & 1) ? vz_0.y : vz_0.x)), c.dxdz);
            if (c.rec_increments || c.it == 0) {  // (no read-modify-write of global memory inside the march)
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
      if (x == op_next) {  // moment-tensor injection, amplitude in double (src/fdModel.cpp:681-751)
        do {
          const RowOp op = c.ops[oq];
          if (own && (op.z >> 1) == kp) {
            const double w = c.stf[(size_t)op.source * c.nt + c.it];
            const double bb = (double)((op.z & 1) ? b.y : b.x);
            const double amp = ((op.weight * w) * c.dt * bb) * op.rd4;
            const bool odd = op.z & 1;
            float &t = op.field == 0 ? (odd ? vx.y : vx.x) : (odd ? vz.y : vz.x);
            t = (float)(op.sign > 0 ? (double)t + amp : (double)t - amp);
          }
          ++oq;
          op_next = oq < oq_end ? c.ops[oq].x : kNoRow;
        } while (op_next == x);
      }
      if (c.point_mode == POINT_ADJOINT && x == rec_next) {  // adjoint sources at the receivers (:951-958)
        do {
          const int z = __shfl_sync(0xffffffffu, lz, rq - win0);
          const double sax = __shfl_sync(0xffffffffu, lax, rq - win0), saz = __shfl_sync(0xffffffffu, laz, rq - win0);
          if (own && (z >> 1) == kp) {
            const double bb = (double)((z & 1) ? b.y : b.x);
            const double ax = ((c.dt * bb) * sax) * c.rdxdz, az = ((c.dt * bb) * saz) * c.rdxdz;
            if (z & 1) {
              vx.y = (float)((double)vx.y + ax);
              vz.y = (float)((double)vz.y + az);
            } else {
              vx.x = (float)((double)vx.x + ax);
              vz.x = (float)((double)vz.x + az);
            }
          }
          rec_next = next_receiver();
        } while (rec_next == x);
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
  GlobalRows32 rows{m.f[c.parity][0], m.f[c.parity][1], m.f[c.parity][2], m.f[c.parity][3], m.f[c.parity][4],
                    m.lm, m.mu, m.b, ld2n, Row32{}};
  fused32_march<MODE>(c, m, rows, lane, kp, xs, xe, own);
}

// The FP32 march fed by a TMA bulk-copy ring: NW consumer warps (adjacent z-warps of one strip) + one producer warp
// whose elected lane issues, per row of the march, 8 bulk-asynchronous copies of (NW*28 + 4) float2 pairs into a slot
// of a `stages`-deep shared-memory ring guarded by full/empty mbarriers -- the design of fused_pipe_kernel (fused.cuh)
// at half the bytes.  (The correlation rows read their snapshot / Lame-kernel values straight from global memory.)
template <int NW>
struct PipeGeom32 {
  static constexpr int kOwned = NW * kFusedOwnedPairs;
  static constexpr int kPairs = kOwned + 4;
  static constexpr int kSeg = kPairs * 8;          // bytes of one row segment (multiple of 16: NW*224 + 32)
  static constexpr int kStage = kSeg * kPipeArrays;
};

template <int MODE, int NW>
__global__ void __launch_bounds__((NW + 1) * 32) fused32_pipe_kernel(const FusedCommon c,
                                                                     const FusedModel32 *__restrict__ models,
                                                                     const int stages) {
  using G = PipeGeom32<NW>;
  extern __shared__ __align__(128) unsigned char psv_smem[];
  unsigned long long *bars = reinterpret_cast<unsigned long long *>(psv_smem);  // [stages] full, [stages] empty
  unsigned char *ring = psv_smem + 128 * ((2 * stages * 8 + 127) / 128);
  const unsigned full0 = smem_u32(bars), empty0 = full0 + 8 * stages;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < stages; ++s) {
      mbar_init(full0 + 8 * s, 1);
      mbar_init(empty0 + 8 * s, NW);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const FusedModel32 &m = models[blockIdx.z];
  const long long ld2n = c.g.ld >> 1;
  const int xs = blockIdx.y * c.strip;
  const int xe = min(xs + c.strip, c.g.nx);
  const int kp_tile = blockIdx.x * G::kOwned - 2;  // first pair of the CTA's row segment
  if (warp == NW) {
    if (lane == 0) {
      const float *src[kPipeArrays] = {m.f[c.parity][0], m.f[c.parity][1], m.f[c.parity][2], m.f[c.parity][3],
                                       m.f[c.parity][4], m.lm,           m.mu,             m.b};
      const long long row_off[kPipeArrays] = {2, 1, 0, 0, 0, 0, 0, -2};  // row of each array relative to j
      int slot = 0;
      unsigned phase = 0;
      const unsigned ring_u32 = smem_u32(ring);
      for (int j = xs - 2; j < xe + 2; ++j) {
        mbar_wait(empty0 + 8 * slot, phase ^ 1);
        const unsigned bar = full0 + 8 * slot;
        mbar_expect_tx(bar, G::kStage);
        const unsigned dst = ring_u32 + slot * G::kStage;
#pragma unroll
        for (int a = 0; a < kPipeArrays; ++a)
          bulk_g2s(dst + a * G::kSeg,
                   reinterpret_cast<const float2 *>(src[a]) + ((long long)j + row_off[a]) * ld2n + kp_tile, G::kSeg, bar);
        if (++slot == stages) {
          slot = 0;
          phase ^= 1;
        }
      }
    }
    return;
  }
  PipeRows32 rows{ring, full0, empty0, stages, G::kSeg, G::kStage, (warp * kFusedOwnedPairs + lane) * 8, 0, 0u, lane};
  const int kp = kp_tile + warp * kFusedOwnedPairs + lane;
  const bool own = lane >= 2 && lane < 30 && kp < (int)ld2n;
  fused32_march<MODE>(c, m, rows, lane, kp, xs, xe, own);
}

}  // namespace psv
